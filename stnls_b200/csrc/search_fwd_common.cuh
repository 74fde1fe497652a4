// Shared types and device helpers of the forward search kernels (search_fwd.cu: direct kernel;
// search_fwd_tab.cuh: lattice-table kernel).  Index math follows SURVEY.md appendix A.
#pragma once
#include "common.cuh"

namespace stnls_fwd {

struct FwdParams {
  const float* q;
  const float* k;
  const void* flows;
  float* dists;
  void* inds;
  int* kselect;
  uint8_t* reflect;
  stnls_geom g;
  stnls_search_opts o;
  int L, G, nG, wsz, Kout, W_t;
  int64_t nquery;  // B*HD*Tq*nH*nW
};

template <bool INT>
struct CT {
  typedef float type;
};
template <>
struct CT<true> {
  typedef int type;
};

// per-group (NLS: frame slot st_i ; refine: input offset ki) search-window description
template <typename C>
struct Group {
  int t;   // candidate frame (global index); may be out of range only if ok==false
  C ah, aw, offh, offw;
  int ok;  // refine: input flow legal (refinement_int_kernel.cu:115-117)
};

struct QueryCtx {
  int b, hd, n, t, tl, nh, nw;  // n = b*HD+hd ; tl = t - tq0
  int rh, rw;                   // q-side pixel (strideQ grid)
  int ah, aw;                   // k-side pixel (stride0 grid) == "adj_patch"
  int t_max;
};

template <bool INT, bool REFINE>
__device__ __forceinline__ Group<typename CT<INT>::type> group_params(const FwdParams& P, const QueryCtx& c, int gi) {
  typedef typename CT<INT>::type C;
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  Group<C> gp;
  gp.ok = 1;
  int hd_f = c.hd % o.HD_f;
  C s1 = INT ? (C)(int)o.stride1 : (C)o.stride1;
  C half = (C)((P.wsz - 1) / 2);
  if (!REFINE) {
    int t_next = c.t + gi;
    gp.t = (t_next > c.t_max) ? c.t_max - gi : t_next;
    if (gi >= o.st_offset) {
      int nF = P.W_t - o.st_offset;
      int64_t fb = ((((int64_t)c.b * o.HD_f + hd_f) * g.Tq + c.tl) * nF + (gi - o.st_offset)) * 2;
      int64_t plane = (int64_t)g.nH * g.nW, off = (int64_t)c.nh * g.nW + c.nw;
      const C* fl = reinterpret_cast<const C*>(P.flows);
      C fh = fl[(fb + 1) * plane + off], fw = fl[fb * plane + off];
      gp.ah = bnd((C)(c.ah + fh), g.kH);
      gp.aw = bnd((C)(c.aw + fw), g.kW);
    } else {
      gp.ah = (C)c.ah;
      gp.aw = (C)c.aw;
    }
  } else {
    int64_t fo = ((((((int64_t)c.b * o.HD_f + hd_f) * g.Tq + c.tl) * g.nH + c.nh) * g.nW + c.nw) * o.Ks + gi) * 3;
    const C* fl = reinterpret_cast<const C*>(P.flows);
    C f0 = fl[fo], f1 = fl[fo + 1], f2 = fl[fo + 2];
    if (INT) {
      gp.t = bnd(c.t + (int)f0, g.T);
      gp.ok = (abs((int)f1) < 1e5) && (abs((int)f2) < 1e5);
      gp.ah = bnd((C)(c.ah + f1), g.kH);
      gp.aw = bnd((C)(c.aw + f2), g.kW);
    } else {
      float pt0 = bnd((float)c.t + floorf((float)f0 + 0.5f), g.T);
      gp.t = (int)pt0;  // integer valued; out-of-range values are caught by the validity test
      gp.ok = (fabsf((float)f1) < 1e8) && (fabsf((float)f2) < 1e8);
      float ch = (float)c.ah + (float)f1, cw = (float)c.aw + (float)f2;
      if (gp.ok && P.reflect != nullptr) {
        int64_t ro = fo / 3 * 2;
        P.reflect[ro] = !inb(ch, g.kH);
        P.reflect[ro + 1] = !inb(cw, g.kW);
      }
      gp.ah = (C)bnd(ch, g.kH);
      gp.aw = (C)bnd(cw, g.kW);
    }
  }
  gp.offh = search_off(gp.ah, s1, half, P.wsz, g.kH, o.full_ws != 0);
  gp.offw = search_off(gp.aw, s1, half, P.wsz, g.kW, o.full_ws != 0);
  return gp;
}

template <typename C>
__device__ __forceinline__ Group<C> shfl_group(const Group<C>& m, int src) {
  Group<C> r;
  r.t = __shfl_sync(FULL_MASK, m.t, src);
  r.ah = __shfl_sync(FULL_MASK, m.ah, src);
  r.aw = __shfl_sync(FULL_MASK, m.aw, src);
  r.offh = __shfl_sync(FULL_MASK, m.offh, src);
  r.offw = __shfl_sync(FULL_MASK, m.offw, src);
  r.ok = __shfl_sync(FULL_MASK, m.ok, src);
  return r;
}

// candidate centre inside its group (slot s = wi*wsz + wj)
__device__ __forceinline__ void cand_center(const FwdParams& P, const Group<int>& gp, int s, int& ch, int& cw) {
  int s1 = (int)P.o.stride1;
  int wi = s / P.wsz, wj = s - wi * P.wsz;
  ch = gp.ah + s1 * (wi - gp.offh);
  cw = gp.aw + s1 * (wj - gp.offw);
}
__device__ __forceinline__ void cand_center(const FwdParams& P, const Group<float>& gp, int s, float& ch, float& cw) {
  float s1 = P.o.stride1;
  int wi = s / P.wsz, wj = s - wi * P.wsz;
  ch = fmaf(s1, (float)wi - gp.offh, gp.ah);
  cw = fmaf(s1, (float)wj - gp.offw, gp.aw);
}

// relative offset reported for a candidate (non_local_search_int_kernel.cu:177-179); entries of an
// illegal refine group keep their initial value -1e8 (lib/stnls/search/utils.py:13-18)
template <typename C>
__device__ __forceinline__ void cand_inds(const FwdParams& P, const QueryCtx& c, const Group<C>& gp, int s, C* out) {
  if (!gp.ok) {
    out[0] = out[1] = out[2] = (C)(-1e8);
    return;
  }
  C ch, cw;
  cand_center(P, gp, s, ch, cw);
  out[0] = (C)gp.t - (C)c.t;
  out[1] = ch - (C)c.ah;
  out[2] = cw - (C)c.aw;
}

// ---- patch distances ------------------------------------------------------------------------
template <int DIST>
__device__ __forceinline__ float acc4(float dist, float4 a, float4 b) {
  if (DIST == 0) {
    dist = fmaf(a.x, b.x, dist);
    dist = fmaf(a.y, b.y, dist);
    dist = fmaf(a.z, b.z, dist);
    dist = fmaf(a.w, b.w, dist);
  } else {
    float d;
    d = a.x - b.x; dist = fmaf(d, d, dist);
    d = a.y - b.y; dist = fmaf(d, d, dist);
    d = a.z - b.z; dist = fmaf(d, d, dist);
    d = a.w - b.w; dist = fmaf(d, d, dist);
  }
  return dist;
}

// nls_int.cu:4-83 (pt == 1)
template <int DIST>
__device__ __forceinline__ float patch_dist(const FwdParams& P, const float* qf, const float* kf, const QueryCtx& c,
                                            int ph, int pw) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  float dist = 0.f;
  for (int pi = 0; pi < o.ps; pi++) {
    int d = o.dilation * (pi + o.patch_offset);
    int r1 = c.rh + o.off_Hq + d, p1 = ph + d;
    if (o.reflect_bounds) { r1 = bnd(r1, g.qH); p1 = bnd(p1, g.kH); }
    if (!(inb(r1, g.qH) && inb(p1, g.kH))) continue;
    for (int pj = 0; pj < o.ps; pj++) {
      int e = o.dilation * (pj + o.patch_offset);
      int r2 = c.rw + o.off_Wq + e, p2 = pw + e;
      if (o.reflect_bounds) { r2 = bnd(r2, g.qW); p2 = bnd(p2, g.kW); }
      if (!(inb(r2, g.qW) && inb(p2, g.kW))) continue;
      const float* qp = qf + ((int64_t)r1 * g.qW + r2) * g.Fp;
      const float* kp = kf + ((int64_t)p1 * g.kW + p2) * g.Fp;
      for (int c4 = 0; c4 < g.Fp; c4 += 4) dist = acc4<DIST>(dist, ldg4(qp + c4), ldg4(kp + c4));
    }
  }
  return dist;
}

// nls_bilin2d.cu:3-90 + shared_kernel.cu:738-771 (pt == 1)
template <int DIST>
__device__ __forceinline__ float patch_dist(const FwdParams& P, const float* qf, const float* kf, const QueryCtx& c,
                                            float ph, float pw) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  float dist = 0.f;
  for (int pi = 0; pi < o.ps; pi++) {
    int d = o.dilation * (pi + o.patch_offset);
    int r1 = c.rh + o.off_Hq + d;
    float p1 = ph + (float)d;
    if (o.reflect_bounds) { r1 = bnd(r1, g.qH); p1 = bnd(p1, g.kH); }
    if (!(inb(r1, g.qH) && inb(p1, g.kH))) continue;
    int h0 = __float2int_rz(p1), h1 = __float2int_rz(p1 + 1);
    float gh0 = fmaxf(0.f, 1 - fabsf(h0 - p1)), gh1 = fmaxf(0.f, 1 - fabsf(h1 - p1));
    bool vh0 = inb(h0, g.kH), vh1 = inb(h1, g.kH);
    for (int pj = 0; pj < o.ps; pj++) {
      int e = o.dilation * (pj + o.patch_offset);
      int r2 = c.rw + o.off_Wq + e;
      float p2 = pw + (float)e;
      if (o.reflect_bounds) { r2 = bnd(r2, g.qW); p2 = bnd(p2, g.kW); }
      if (!(inb(r2, g.qW) && inb(p2, g.kW))) continue;
      int w0 = __float2int_rz(p2), w1 = __float2int_rz(p2 + 1);
      float gw0 = fmaxf(0.f, 1 - fabsf(w0 - p2)), gw1 = fmaxf(0.f, 1 - fabsf(w1 - p2));
      bool vw0 = inb(w0, g.kW), vw1 = inb(w1, g.kW);
      float w00 = gh0 * gw0, w01 = gh0 * gw1, w10 = gh1 * gw0, w11 = gh1 * gw1;
      const float* qp = qf + ((int64_t)r1 * g.qW + r2) * g.Fp;
      const float* k00 = kf + ((int64_t)h0 * g.kW + w0) * g.Fp;
      const float* k01 = kf + ((int64_t)h0 * g.kW + w1) * g.Fp;
      const float* k10 = kf + ((int64_t)h1 * g.kW + w0) * g.Fp;
      const float* k11 = kf + ((int64_t)h1 * g.kW + w1) * g.Fp;
      bool v00 = vh0 && vw0, v01 = vh0 && vw1, v10 = vh1 && vw0, v11 = vh1 && vw1;
      for (int c4 = 0; c4 < g.Fp; c4 += 4) {
        float4 pix = f4zero();
        if (v00) { float4 v = ldg4(k00 + c4); pix.x = fmaf(w00, v.x, pix.x); pix.y = fmaf(w00, v.y, pix.y); pix.z = fmaf(w00, v.z, pix.z); pix.w = fmaf(w00, v.w, pix.w); }
        if (v01) { float4 v = ldg4(k01 + c4); pix.x = fmaf(w01, v.x, pix.x); pix.y = fmaf(w01, v.y, pix.y); pix.z = fmaf(w01, v.z, pix.z); pix.w = fmaf(w01, v.w, pix.w); }
        if (v10) { float4 v = ldg4(k10 + c4); pix.x = fmaf(w10, v.x, pix.x); pix.y = fmaf(w10, v.y, pix.y); pix.z = fmaf(w10, v.z, pix.z); pix.w = fmaf(w10, v.w, pix.w); }
        if (v11) { float4 v = ldg4(k11 + c4); pix.x = fmaf(w11, v.x, pix.x); pix.y = fmaf(w11, v.y, pix.y); pix.z = fmaf(w11, v.z, pix.z); pix.w = fmaf(w11, v.w, pix.w); }
        dist = acc4<DIST>(dist, ldg4(qp + c4), pix);
      }
    }
  }
  return dist;
}

// ---- anchor helpers -------------------------------------------------------------------------
// refine: the un-shifted input offset, as reported in slot 0 of its group (anchor_self_kernel.cu:586-590)
template <typename C>
__device__ __forceinline__ void refine_loc(const FwdParams& P, const QueryCtx& c, int gi, C* loc) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  int hd_f = c.hd % o.HD_f;
  int64_t fo = ((((((int64_t)c.b * o.HD_f + hd_f) * g.Tq + c.tl) * g.nH + c.nh) * g.nW + c.nw) * o.Ks + gi) * 3;
  const C* fl = reinterpret_cast<const C*>(P.flows);
  C f0 = fl[fo], f1 = fl[fo + 1], f2 = fl[fo + 2];
  loc[0] = (C)round((double)f0);
  loc[1] = bnd((C)((C)c.rh + f1), g.kH) - (C)c.rh;
  loc[2] = bnd((C)((C)c.rw + f2), g.kW) - (C)c.rw;
}

// anchor key of a candidate: <0 exact match, else L1 distance (only < 10000 competes)
__device__ __forceinline__ float anchor_key(const int* ind, const int* loc) {
  return (ind[0] == loc[0] && ind[1] == loc[1] && ind[2] == loc[2]) ? -1.f : INFINITY;
}
__device__ __forceinline__ float anchor_key(const float* ind, const float* loc) {
  float delta = 0.f;
  delta += fabsf(ind[0] - loc[0]);
  delta += fabsf(ind[1] - loc[1]);
  delta += fabsf(ind[2] - loc[2]);
  if (delta < 1e-8) return -1.f;
  return (delta < 10000.f) ? delta : INFINITY;
}

// refine: slot of group gi that holds the un-shifted input offset (anchor_self_kernel.cu:592-616)
template <typename C>
__device__ __forceinline__ int refine_group_anchor(const FwdParams& P, const QueryCtx& c, const Group<C>& gp, int gi) {
  C loc[3];
  refine_loc<C>(P, c, gi, loc);
  int best = 0;
  float bkey = INFINITY;
  for (int s = 0; s < P.G; s++) {
    C ind[3];
    cand_inds<C>(P, c, gp, s, ind);
    float key = anchor_key(ind, loc);
    if (key < 0.f) return s;
    if (key < bkey) { bkey = key; best = s; }
  }
  return best;
}

// ---- warp top-K list: lane i holds the i-th best (key, pos); smaller key is better, ties -> lower pos
struct TopK {
  float key;
  int pos;
};
__device__ __forceinline__ void topk_insert(TopK& me, float nk, int np, int Klist, int lane) {
  bool newer_better = (nk < me.key) || (nk == me.key && np < me.pos);
  unsigned m = __ballot_sync(FULL_MASK, !newer_better && lane < Klist);
  int ins = __popc(m);
  float sk = __shfl_up_sync(FULL_MASK, me.key, 1);
  int sp = __shfl_up_sync(FULL_MASK, me.pos, 1);
  if (ins < Klist) {
    if (lane == ins) { me.key = nk; me.pos = np; }
    else if (lane > ins) { me.key = sk; me.pos = sp; }
  }
}

}  // namespace stnls_fwd
