// "Lattice table" forward search kernel.
//
// All patch pixels of all candidates of one (query, group) -- group = frame slot of the NLS window or
// input offset of a refinement -- lie on a lattice  pos(m) = base + u*m  per axis (u = 1 for integer
// stride1; 1/2, 1/4 for stride1 = 0.5, 0.25).  One warp owns one query and, per batch of groups:
//   1. axis tables   per (group, axis, m): reflected position -> tap offsets, bilinear weights, validity
//   2. lattice table T[group][channel quad][mh][mw] (shared memory): the key value at every lattice
//      point, bilinear-interpolated ONCE (the reference re-interpolates it for each of the ~ps*ps
//      candidates whose patches cover it); global loads coalesced over channel quads
//   3. candidates    lane = candidate; ps*ps*F_head FMAs reading T and the query patch from shared
//      memory, accumulated in the reference's order (pi -> pj -> channel) so the int path stays
//      bit-identical to the reference kernels
//   4. keys of all L candidates stay in shared memory; the K best are then picked by K rounds of
//      redux.sync min over the warp (ties -> lower candidate index).
// Reference semantics: SURVEY.md appendix A; lib/csrc/search/nls_{int,bilin2d}.cu,
// lib/csrc/shared_kernel.cu:49-60,248-287,738-771, lib/csrc/nn/anchor_self_kernel.cu:24-116,547-650,
// lib/stnls/nn/topk.py:65-123.
#pragma once
#include "search_fwd_tab.cuh"

namespace stnls_fwd {

__device__ __forceinline__ int fdiv(int x, float inv) { return __float2int_rz(((float)x + 0.5f) * inv); }

// order-preserving float -> uint map (smaller float <=> smaller uint); 0xffffffff is reserved for "taken"
__device__ __forceinline__ unsigned f2ord(float f) {
  unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// shared-memory accessors on 32-bit shared-window byte addresses (keeps address math to one IADD/IMAD)
__device__ __forceinline__ float4 lds4(unsigned a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts4(unsigned a, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned lds1(unsigned a) {
  unsigned v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts1(unsigned a, unsigned v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned ldsh(unsigned a) {
  unsigned short v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
  return v;
}
// 16-byte read-only global load at base + 32-bit byte offset
__device__ __forceinline__ float4 ldg4b(const char* base, unsigned off) {
  return __ldg(reinterpret_cast<const float4*>(base + off));
}

__device__ __forceinline__ void fma4(float4& v, float w, const float4 t) {
  v.x = fmaf(w, t.x, v.x); v.y = fmaf(w, t.y, v.y); v.z = fmaf(w, t.z, v.z); v.w = fmaf(w, t.w, v.w);
}

// key value at one lattice point: plain read (int) or 4-tap bilinear, taps outside the frame dropped
// (shared_kernel.cu:738-771; tap order (0,0),(0,1),(1,0),(1,1) and weight gh*gw as in the reference)
template <bool INT>
__device__ __forceinline__ float4 lattice_value(const char* kc, const float4 eh, const float4 ew) {
  const int oh0 = __float_as_int(eh.x), ow0 = __float_as_int(ew.x);   // byte offsets (always inside the frame)
  if (INT) return ldg4b(kc, (unsigned)(oh0 + ow0));
  const int oh1 = __float_as_int(eh.y), ow1 = __float_as_int(ew.y);
  // a tap outside the frame carries weight 0 and a legal offset (axis tables), so no tap branches:
  // fma(0, finite, v) == v, i.e. the same value the reference gets by dropping the tap
  const float4 t00 = ldg4b(kc, (unsigned)(oh0 + ow0)), t01 = ldg4b(kc, (unsigned)(oh0 + ow1));
  const float4 t10 = ldg4b(kc, (unsigned)(oh1 + ow0)), t11 = ldg4b(kc, (unsigned)(oh1 + ow1));
  float4 v = f4zero();
  fma4(v, eh.z * ew.z, t00);
  fma4(v, eh.z * ew.w, t01);
  fma4(v, eh.w * ew.z, t10);
  fma4(v, eh.w * ew.w, t11);
  return v;
}

// PS: patch size (0 = run time); Q4T: channel quads (0 = run time)
template <bool INT, int DIST, bool REFINE, int PS, int Q4T>
__global__ void __launch_bounds__(256, (PS == 3 && Q4T == 4) ? 2 : (Q4T == 0) ? 3 : 4) search_fwd_tab_kernel(const __grid_constant__ FwdParams P,
                                                                const __grid_constant__ TabCfg C) {
  typedef typename CT<INT>::type CC;
  extern __shared__ float4 smem4[];
  const unsigned sb = (unsigned)__cvta_generic_to_shared(smem4);  // all smem traffic goes through lds*/sts* on sb
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int M = C.M;
  const int Q4 = (Q4T > 0) ? Q4T : C.Q4;
  const int ps = (PS > 0) ? PS : o.ps, ps2 = ps * ps;

  // per-CTA decode table of the lattice points: x -> (mh << 8) | mw
  for (int x = threadIdx.x; x < M * M; x += blockDim.x) {
    int mh = fdiv(x, C.inv_M);
    reinterpret_cast<unsigned short*>(smem4)[x] = (unsigned short)((mh << 8) | (x - mh * M));  // plain store, before the barrier
  }
  __syncthreads();

  const int64_t wq = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  if (wq >= P.nquery) return;
  // per-warp smem carve (float4 indices into smem4)
  const int Tb = C.dec_f4 + warp * C.warp_f4;
  const int qb = Tb + C.q_off;
  const int axb = Tb + C.ax_off;              // AxEntry[nb][2][M], one float4 each: (o0, o1, w0, w1)
  const int mskb = (Tb + C.msk_off) * 4;      // uint index: [nb][2] validity masks, then [nb] frame slot
  const int gfrb = mskb + 2 * C.nb;
  const int keyb = (Tb + C.key_off) * 4;      // uint index: [L] ordered keys

  // -- decode query (shared_kernel.cu:76-98) --
  QueryCtx c;
  int64_t r = divmod(wq, g.nW, c.nw);
  r = divmod(r, g.nH, c.nh);
  r = divmod(r, g.Tq, c.tl);
  c.n = (int)r;
  c.b = (int)divmod(r, g.HD, c.hd);
  c.t = c.tl + g.tq0;
  c.ah = (c.nh * o.stride0) % g.kH;
  c.aw = (c.nw * o.stride0) % g.kW;
  if (INT) { c.rh = (c.nh * o.strideQ) % g.qH; c.rw = (c.nw * o.strideQ) % g.qW; }
  else { c.rh = c.ah; c.rw = c.aw; }
  c.t_max = time_tmax(c.t, g.T, o.wt);

  const float* qf = P.q + ((int64_t)c.n * g.Tq + c.tl) * g.qH * g.qW * g.Fp;
  const float* kbase = P.k + (int64_t)c.n * g.Tk * g.kH * g.kW * g.Fp;
  const int64_t kfs = (int64_t)g.kH * g.kW * g.Fp;
  const float invalid = (DIST == 0) ? -INFINITY : INFINITY;
  const bool anchored = o.anchor_self != 0;
  const bool keep_all = o.K <= 0;
  const int64_t obase = wq * P.Kout;
  CC* inds_out = reinterpret_cast<CC*>(P.inds);

  // -- query patch -> smem (once); validity mask of its pixels --
  unsigned qmask;
  {
    bool ok = false;
    int qo = 0;
    if (lane < ps2) {
      int pi = lane / ps, pj = lane - pi * ps;
      int r1 = c.rh + o.off_Hq + o.dilation * (pi + o.patch_offset);
      int r2 = c.rw + o.off_Wq + o.dilation * (pj + o.patch_offset);
      if (o.reflect_bounds) { r1 = bnd(r1, g.qH); r2 = bnd(r2, g.qW); }
      ok = inb(r1, g.qH) && inb(r2, g.qW);
      qo = (r1 * g.qW + r2) * g.Fp;
    }
    qmask = __ballot_sync(FULL_MASK, ok);
    for (int base = 0; base < ps2 * Q4; base += 32) {
      int it = base + lane;
      int p = (Q4T > 0 || C.q4sh >= 0) ? (it >> C.q4sh) : fdiv(it, C.inv_Q4);
      int c4 = it - p * Q4;
      int po = __shfl_sync(FULL_MASK, qo, p & 31);
      if (it < ps2 * Q4 && ((qmask >> p) & 1)) sts4(sb + (qb + it) * 16, ldg4(qf + po + c4 * 4));
    }
  }
  const bool qfull = qmask == ((ps2 >= 32) ? 0xffffffffu : ((1u << ps2) - 1u));

  // window description of the first 32 groups, lane = group: computed once, handed out by shuffles
  const Group<CC> all32 = group_params<INT, REFINE>(P, c, min(lane, P.nG - 1));
  const bool one_chunk = P.nG <= 32;

  // -- pass 1 (NLS anchor): candidate whose offset is (0,0,0) / nearest to it (anchor_self_kernel.cu:62-89)
  int a_nls = 0;
  if (anchored && !REFINE) {
    float bkey = INFINITY;
    int bpos = 0x7fffffff;
    const CC zero[3] = {(CC)0, (CC)0, (CC)0};
    const int nchunks = (P.nG + 31) / 32;
    for (int ch = 0; ch < nchunks; ch++) {
      int gi = ch * 32 + lane;
      Group<CC> mine = all32;
      if (ch > 0) mine = group_params<INT, REFINE>(P, c, min(gi, P.nG - 1));
      int pend = min(P.L, (ch + 1) * 32 * P.G);
      for (int base = ch * 32 * P.G; base < pend; base += 32) {
        int pos = min(base + lane, pend - 1);
        int gq = fdiv(pos, C.inv_G);
        Group<CC> gp = shfl_group(mine, gq - ch * 32);
        int s = pos - gq * P.G;
        int wi = fdiv(s, C.inv_wsz), wj = s - wi * P.wsz;
        CC ind[3];
        ind[0] = (CC)gp.t - (CC)c.t;
        if (INT) {
          ind[1] = (CC)((int)gp.ah + (int)o.stride1 * (wi - (int)gp.offh) - c.ah);
          ind[2] = (CC)((int)gp.aw + (int)o.stride1 * (wj - (int)gp.offw) - c.aw);
        } else {
          ind[1] = (CC)(fmaf(o.stride1, (float)wi - (float)gp.offh, (float)gp.ah) - (float)c.ah);
          ind[2] = (CC)(fmaf(o.stride1, (float)wj - (float)gp.offw, (float)gp.aw) - (float)c.aw);
        }
        float key = anchor_key(ind, zero);
        if (base + lane < pend && (key < bkey || (key == bkey && pos < bpos)) && key < INFINITY) { bkey = key; bpos = pos; }
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      float ok = __shfl_xor_sync(FULL_MASK, bkey, off);
      int op = __shfl_xor_sync(FULL_MASK, bpos, off);
      if (ok < bkey || (ok == bkey && op < bpos)) { bkey = ok; bpos = op; }
    }
    a_nls = (bpos == 0x7fffffff) ? 0 : bpos;
  }

  // -- pass 2: per batch of groups: axis tables -> lattice table -> candidate distances --
  const CC s1c = INT ? (CC)(int)o.stride1 : (CC)o.stride1;
  const int rowsc = g.kW * g.Fp * 4;  // bytes per key row
  for (int g0 = 0; g0 < P.nG; g0 += C.nb) {
    const int nbg = min(C.nb, P.nG - g0);
    Group<CC> mine;
    if (one_chunk) mine = shfl_group(all32, g0 + min(lane, nbg - 1));
    else mine = group_params<INT, REFINE>(P, c, g0 + min(lane, nbg - 1));
    int my_ag = 0;
    if (REFINE && anchored) my_ag = refine_group_anchor<CC>(P, c, mine, g0 + min(lane, nbg - 1));
    if (lane < nbg) sts1(sb + (gfrb + lane) * 4, (unsigned)((mine.ok && resident(mine.t, g)) ? (mine.t - g.tk0) : -1));
    // raw mode is taken for this batch when every window in it is interior: valid frame, no reflection,
    // consecutive in-frame taps and the same interpolation weights at every lattice coordinate
    constexpr bool RAWOK = !INT && DIST == 0 && Q4T > 0 && (PS == 1 || (PS == 3 && Q4T == 4));
    bool useraw = RAWOK && C.raw && qfull &&
                  (__ballot_sync(FULL_MASK, lane >= nbg || (mine.ok && resident(mine.t, g))) == FULL_MASK);
    // 1. axis tables: one 32-lane round per group (both axes, M <= 16) or per (group, axis)
    for (int rd = 0; rd < nbg * C.ax_rounds; rd++) {
      int gl, axis, m;
      if (C.ax_rounds == 1) { gl = rd; axis = lane >> 4; m = lane & 15; }
      else { gl = rd >> 1; axis = rd & 1; m = lane; }
      Group<CC> gp = shfl_group(mine, gl);
      const int lim = axis ? g.kW : g.kH;
      const int sc = axis ? g.Fp * 4 : rowsc;
      const CC a = axis ? gp.aw : gp.ah, off = axis ? gp.offw : gp.offh;
      bool pv = false, reg_e = false;
      int o0 = -1, o1 = -1;
      float w0 = 0.f, w1 = 0.f;
      if (m < M) {
        if (INT) {
          int p = (int)a - (int)s1c * (int)off + o.dilation * o.patch_offset + m;
          if (o.reflect_bounds) p = bnd(p, lim);
          pv = inb(p, lim);
          o0 = pv ? p * sc : 0;  // invalid lattice points are filled from a legal address and never read
        } else {
          float basep = fmaf(-(float)s1c, (float)off, (float)a) + (float)(o.dilation * o.patch_offset);
          float p = fmaf(C.u, (float)m, basep);
          if (o.reflect_bounds) p = bnd(p, lim);
          pv = inb(p, lim);
          int t0 = __float2int_rz(p), t1 = __float2int_rz(p + 1);
          // a tap outside the frame is dropped by the reference (shared_kernel.cu:752-756): here its
          // weight becomes 0 and its offset a legal one, so the lattice loop below needs no tap branches
          const bool v0 = inb(t0, lim), v1 = inb(t1, lim);
          w0 = v0 ? fmaxf(0.f, 1 - fabsf(t0 - p)) : 0.f;
          w1 = v1 ? fmaxf(0.f, 1 - fabsf(t1 - p)) : 0.f;
          o0 = v0 ? t0 * sc : 0;
          o1 = v1 ? t1 * sc : (v0 ? t0 * sc : 0);
          reg_e = pv && v0 && v1 && (o1 == o0 + sc);
        }
        sts4(sb + (axb + (gl * 2 + axis) * M + m) * 16, make_float4(__int_as_float(o0), __int_as_float(o1), w0, w1));
      }
      if (RAWOK && useraw) {   // warp-uniform
        const int src = (C.ax_rounds == 1) ? (lane & 16) : 0;
        const int o00 = __shfl_sync(FULL_MASK, o0, src);
        const float w00 = __shfl_sync(FULL_MASK, w0, src), w10 = __shfl_sync(FULL_MASK, w1, src);
        const bool ok_e = (m >= M) || (reg_e && o0 == o00 + m * sc && w0 == w00 && w1 == w10);
        useraw = __ballot_sync(FULL_MASK, ok_e) == FULL_MASK;
      }
      unsigned bal = __ballot_sync(FULL_MASK, pv);
      if (lane == 0) {
        if (C.ax_rounds == 1) { sts1(sb + (mskb + gl * 2) * 4, bal & 0xffffu); sts1(sb + (mskb + gl * 2 + 1) * 4, bal >> 16); }
        else sts1(sb + (mskb + gl * 2 + axis) * 4, bal);
      }
    }
    __syncwarp();
    // 2. lattice table: item = (mh, mw, channel quad), quads fastest so a warp reads whole pixels
    const int MM = M * M;
    for (int gl = 0; gl < nbg; gl++) {
      if (RAWOK && useraw) break;
      const int fr = (int)lds1(sb + (gfrb + gl) * 4);
      if (fr < 0) continue;
      const unsigned Ah = sb + (axb + (gl * 2) * M) * 16, Aw = Ah + M * 16;
      const int Tg = Tb + gl * C.GS4;
      const char* kf = reinterpret_cast<const char*>(kbase + (int64_t)fr * kfs);
      if (Q4T > 0) {  // 32 % Q4T == 0: the channel quad of a lane never changes
        const int c4 = lane & (Q4T - 1);
        const char* kc = kf + c4 * 16;
        asm volatile("" : "+l"(kc));  // keep the frame pointer in registers (no re-derivation per tap)
        const unsigned Tgc = sb + (Tg + c4 * C.PS4) * 16;
        const unsigned Mp16 = C.Mp * 16;
#pragma unroll 2
        for (int x = lane / Q4T; x < MM; x += 32 / Q4T) {
          const unsigned d = ldsh(sb + x * 2);
          const unsigned mh = d >> 8, mw = d & 255;
          sts4(Tgc + mh * Mp16 + mw * 16, lattice_value<INT>(kc, lds4(Ah + mh * 16), lds4(Aw + mw * 16)));
        }
      } else {
#pragma unroll 2
        for (int it = lane; it < MM * Q4; it += 32) {
          int x, c4;
          if (C.q4sh >= 0) { x = it >> C.q4sh; c4 = it & (Q4 - 1); }
          else { x = fdiv(it, C.inv_Q4); c4 = it - x * Q4; }
          const unsigned d = ldsh(sb + x * 2);
          const unsigned mh = d >> 8, mw = d & 255;
          sts4(sb + (Tg + c4 * C.PS4 + mh * C.Mp + mw) * 16,
               lattice_value<INT>(kf + c4 * 16, lds4(Ah + mh * 16), lds4(Aw + mw * 16)));
        }
      }
    }
    __syncwarp();
    // 3. candidates of this batch
    const unsigned qa = sb + qb * 16, rstep = C.n2 * C.Mp * 16, cstep = C.n2 * 16, pstep = C.PS4 * 16;
    // distance of one candidate with per-pixel validity masks (frame borders, clipped query patches)
    auto dist_masked = [&](unsigned Tc, unsigned hm, unsigned wm) {
      float dist = 0.f;
      for (int pi = 0; pi < ps; pi++) {
        if (!((hm >> (C.n2 * pi)) & 1)) continue;
        for (int pj = 0; pj < ps; pj++) {
          if (!(((wm >> (C.n2 * pj)) & (qmask >> (pi * ps + pj))) & 1)) continue;
          const unsigned tp = Tc + pi * rstep + pj * cstep;
          const unsigned qp = qa + (pi * ps + pj) * Q4 * 16;
          for (int c4 = 0; c4 < Q4; c4++) dist = acc4<DIST>(dist, lds4(qp + c4 * 16), lds4(tp + c4 * pstep));
        }
      }
      return dist;
    };
    // anchor swap (an involution; anchor_self_kernel.cu:101-113), slot-0 / keep-all outputs, key of the top-K
    auto finish = [&](bool active, int gq, const Group<CC>& gp, int ag, int s, float d) {
      const int pos = gq * P.G + s;
      int eff = pos;
      if (anchored) {
        if (REFINE) { if (s == 0) eff = gq * P.G + ag; else if (s == ag) eff = gq * P.G; }
        else { if (pos == 0) eff = a_nls; else if (pos == a_nls) eff = 0; }
      }
      if (keep_all || (anchored && eff == 0)) {
        if (active) {
          const bool to_slot0 = anchored && (REFINE ? (eff % P.G == 0 && ag != 0) : (eff == 0 && a_nls != 0));
          int64_t oi = obase + (keep_all ? eff : 0);
          CC ind[3];
          if (to_slot0) {
            if (REFINE) refine_loc<CC>(P, c, eff / P.G, ind);
            else ind[0] = ind[1] = ind[2] = (CC)0;
          } else cand_inds<CC>(P, c, gp, s, ind);
          P.dists[oi] = d;
          inds_out[oi * 3] = ind[0]; inds_out[oi * 3 + 1] = ind[1]; inds_out[oi * 3 + 2] = ind[2];
          if (P.kselect) P.kselect[oi] = gp.ok ? gq : 1;
        }
        if (keep_all) return;
      }
      if (active && !(anchored && eff == 0))
        sts1(sb + (keyb + eff) * 4, min(f2ord(((DIST == 0) ? -d : d) + 0.f), 0xfffffffeu));
    };

    if (RAWOK && useraw) {
      constexpr int Q4C = Q4T > 0 ? Q4T : 1;
      const int Mr = C.Mr, Wr = C.Wr, WW = Wr * Wr;
      const unsigned scr = sb + (Tb + C.scr_off) * 16;
      const int pixb = g.Fp * 4;
      // 2r. raw key pixels of the (M+1)^2 footprint, one load each
      for (int gl = 0; gl < nbg; gl++) {
        const int fr = (int)lds1(sb + (gfrb + gl) * 4);
        const unsigned Ah = sb + (axb + (gl * 2) * M) * 16, Aw = Ah + M * 16;
        const int base = __float_as_int(lds4(Ah).x) + __float_as_int(lds4(Aw).x);
        const int c4 = lane & (Q4C - 1);
        const char* kc = reinterpret_cast<const char*>(kbase + (int64_t)fr * kfs) + c4 * 16;
        const unsigned Tgc = sb + (Tb + gl * C.GS4 + c4 * C.PS4) * 16;
#pragma unroll 2
        for (int x = lane / Q4C; x < Mr * Mr; x += 32 / Q4C) {
          const int mh = fdiv(x, C.inv_Mr), mw = x - mh * Mr;
          sts4(Tgc + (mh * C.Mp + mw) * 16, ldg4b(kc, (unsigned)(base + mh * rowsc + mw * pixb)));
        }
      }
      __syncwarp();
      // 3r. correlations of the query patch with the raw pixels on the (ws+1)^2 integer grid
      if (PS == 1) {
        const int nPt = nbg * WW;
        for (int base = 0; base < nPt; base += 32) {
          const int e = min(base + lane, nPt - 1);
          const int gl = (C.nb == 1) ? 0 : fdiv(e, C.inv_nPg);
          const int x = e - gl * WW;
          const int a = fdiv(x, C.inv_Wr), b = x - a * Wr;
          const unsigned R = sb + (Tb + gl * C.GS4 + a * C.Mp + b) * 16;
          float cv = 0.f;
#pragma unroll
          for (int c4 = 0; c4 < Q4C; c4++) cv = acc4<DIST>(cv, lds4(qa + c4 * 16), lds4(R + c4 * pstep));
          if (base + lane < nPt) sts1(scr + e * 4, __float_as_uint(cv));
        }
      } else {
        const int nTg = Wr * C.nrr, nT = nbg * nTg;
        for (int base = 0; base < nT; base += 32) {
          const int tb = min(base + lane, nT - 1);
          const int gl = (C.nb == 1) ? 0 : fdiv(tb, C.inv_nTr);
          const int e = tb - gl * nTg;
          int run, a;                                      // grid row fastest: conflict-free reads (odd row pitch)
          if (e < C.nrr * C.rA8) { run = fdiv(e, C.inv_rA8); a = e - run * C.rA8; }
          else { const int e2 = e - C.nrr * C.rA8; run = fdiv(e2, C.inv_rrem); a = C.rA8 + e2 - run * (Wr - C.rA8); }
          const unsigned rd = C.rruns[run];
          const int w0 = rd & 255, cnt = rd >> 8;
          const unsigned T0 = sb + (Tb + gl * C.GS4 + a * C.Mp + w0) * 16;
          float dist[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int pi = 0; pi < 3; pi++) {
            float4 qv[3][Q4C];
#pragma unroll
            for (int pj = 0; pj < 3; pj++)
#pragma unroll
              for (int c4 = 0; c4 < Q4C; c4++) qv[pj][c4] = lds4(qa + ((pi * 3 + pj) * Q4C + c4) * 16);
#pragma unroll
            for (int cc = 0; cc < 6; cc++) {
#pragma unroll
              for (int c4 = 0; c4 < Q4C; c4++) {
                const float4 tv = lds4(T0 + pi * rstep + cc * 16 + c4 * pstep);
#pragma unroll
                for (int j = 0; j < 4; j++)
                  if (cc - j >= 0 && cc - j < 3) dist[j] = acc4<DIST>(dist[j], qv[cc - j][c4], tv);
              }
            }
          }
          if (base + lane < nT) {
#pragma unroll
            for (int j = 0; j < 4; j++)
              if (j < cnt) sts1(scr + (gl * WW + a * Wr + w0 + j) * 4, __float_as_uint(dist[j]));
          }
        }
      }
      __syncwarp();
      // 4r. candidate (wi,wj) = bilinear blend of the four correlations around it (tap order of the reference)
      const int nC = nbg * P.G;
      for (int base = 0; base < nC; base += 32) {
        const bool active = base + lane < nC;
        const int posb = min(base + lane, nC - 1);
        const int gl = (C.nb == 1) ? 0 : fdiv(posb, C.inv_G);
        const int s = posb - gl * P.G;
        const int wi = fdiv(s, C.inv_wsz), wj = s - wi * P.wsz;
        Group<CC> gp = shfl_group(mine, gl);
        const int ag = (REFINE && anchored) ? __shfl_sync(FULL_MASK, my_ag, gl) : 0;
        const unsigned Ah = sb + (axb + (gl * 2) * M) * 16, Aw = Ah + M * 16;
        const float4 eh = lds4(Ah), ew = lds4(Aw);
        const unsigned cb = scr + (gl * WW + wi * Wr + wj) * 4;
        float d = 0.f;
        d = fmaf(eh.z * ew.z, __uint_as_float(lds1(cb)), d);
        d = fmaf(eh.z * ew.w, __uint_as_float(lds1(cb + 4)), d);
        d = fmaf(eh.w * ew.z, __uint_as_float(lds1(cb + Wr * 4)), d);
        d = fmaf(eh.w * ew.w, __uint_as_float(lds1(cb + Wr * 4 + 4)), d);
        finish(active, g0 + gl, gp, ag, s, d);
      }
    } else if (PS == 3 && Q4T == 4 && C.tile) {   // 16 channels per head: the variant compiled with the register budget for it
      // 3a. lane = RUN of up to 3 candidates of one window row whose patches overlap column-wise (candidate
      // j of the run uses lattice columns c = pj + j of the run's 5): every T value is loaded once for the
      // run instead of once per candidate (15 instead of 27 LDS.128 per quad) and the query patch row sits
      // in registers (9 instead of 27 broadcast reads) -- the shared-memory pipe is what bounds this kernel.
      // Each candidate still accumulates pi -> pj -> channel, so the int path stays bit-exact.
      constexpr int Q4C = Q4T > 0 ? Q4T : 1;
      const int nTg = P.wsz * C.nruns, nT = nbg * nTg;
      for (int base = 0; base < nT; base += 32) {
        const bool tact = base + lane < nT;
        const int tb = min(base + lane, nT - 1);
        const int gl = (C.nb == 1) ? 0 : fdiv(tb, C.inv_nTg);
        const int e = tb - gl * nTg;
        int run, wi;                                // window row fastest: conflict-free T reads (odd row pitch)
        if (e < C.nruns * C.tA8) { run = fdiv(e, C.inv_tA8); wi = e - run * C.tA8; }
        else { const int e2 = e - C.nruns * C.tA8; run = fdiv(e2, C.inv_trem); wi = C.tA8 + e2 - run * (P.wsz - C.tA8); }
        const unsigned rd = C.runs[run];
        const int w0 = rd & 255, cnt = rd >> 8;
        const int gq = g0 + gl;
        Group<CC> gp = shfl_group(mine, gl);
        const int ag = (REFINE && anchored) ? __shfl_sync(FULL_MASK, my_ag, gl) : 0;
        CC chh;
        if (INT) chh = (CC)((int)gp.ah + (int)s1c * (wi - (int)gp.offh));
        else chh = (CC)fmaf((float)s1c, (float)wi - (float)gp.offh, (float)gp.ah);
        const bool rowok = tact && gp.ok && resident(gp.t, g) && inb(chh, g.kH);
        const int mh0 = C.n1 * wi;
        const unsigned hm = lds1(sb + (mskb + gl * 2) * 4) >> mh0, wmask = lds1(sb + (mskb + gl * 2 + 1) * 4);
        const bool hfull = qfull && ((hm & C.rpat) == C.rpat);
        const unsigned Tr = sb + (Tb + gl * C.GS4 + mh0 * C.Mp) * 16;
        bool cv[3], fastok = true;
#pragma unroll
        for (int r = 0; r < 3; r++) {
          const int wj = w0 + C.rs * r;
          CC cww;
          if (INT) cww = (CC)((int)gp.aw + (int)s1c * (wj - (int)gp.offw));
          else cww = (CC)fmaf((float)s1c, (float)wj - (float)gp.offw, (float)gp.aw);
          cv[r] = rowok && r < cnt && inb(cww, g.kW);
          const unsigned wm = wmask >> (C.n1 * wj);
          fastok = fastok && (!cv[r] || (hfull && ((wm & C.rpat) == C.rpat)));
        }
        float dr[3] = {invalid, invalid, invalid};
        if (__all_sync(FULL_MASK, fastok)) {
          if (cv[0] || cv[1] || cv[2]) {
            float dist[3] = {0.f, 0.f, 0.f};
            const unsigned T0 = Tr + C.n1 * w0 * 16;
#pragma unroll
            for (int pi = 0; pi < 3; pi++) {
              float4 qv[3][Q4C];
#pragma unroll
              for (int pj = 0; pj < 3; pj++)
#pragma unroll
                for (int c4 = 0; c4 < Q4C; c4++) qv[pj][c4] = lds4(qa + ((pi * 3 + pj) * Q4C + c4) * 16);
#pragma unroll
              for (int cc = 0; cc < 5; cc++) {
#pragma unroll
                for (int c4 = 0; c4 < Q4C; c4++) {
                  const float4 tv = lds4(T0 + pi * rstep + cc * cstep + c4 * pstep);
#pragma unroll
                  for (int j = 0; j < 3; j++)
                    if (cc - j >= 0 && cc - j < 3) dist[j] = acc4<DIST>(dist[j], qv[cc - j][c4], tv);
                }
              }
            }
#pragma unroll
            for (int r = 0; r < 3; r++) if (cv[r]) dr[r] = dist[r];
          }
        } else {
#pragma unroll
          for (int r = 0; r < 3; r++) {
            const int wj = w0 + C.rs * r;
            if (cv[r]) dr[r] = dist_masked(Tr + C.n1 * wj * 16, hm, wmask >> (C.n1 * wj));
          }
        }
#pragma unroll
        for (int r = 0; r < 3; r++) finish(tact && r < cnt, gq, gp, ag, wi * P.wsz + w0 + C.rs * r, dr[r]);
      }
    } else {
    // 3b. lane = candidate
    const int nC = nbg * P.G;
    for (int base = 0; base < nC; base += 32) {
      const bool active = base + lane < nC;
      const int posb = min(base + lane, nC - 1);
      const int gl = (C.nb == 1) ? 0 : fdiv(posb, C.inv_G);
      // lane -> candidate inside the group.  With W8 > 0 the first wsz*W8 lanes walk rows in chunks of W8
      // (a multiple of 8) columns and the rest walks the leftover columns top to bottom, so the eight lanes
      // of one LDS.128 wavefront read eight different 16-byte bank groups of T (row pitch Mp is odd)
      const int e = posb - gl * P.G;
      int wi, wj;
      if (C.W8 == 0) { wi = fdiv(e, C.inv_wsz); wj = e - wi * P.wsz; }
      else if (e < C.nA) { wi = fdiv(e, C.inv_W8); wj = e - wi * C.W8; }
      else { const int e2 = e - C.nA; const int cc = fdiv(e2, C.inv_wsz); wi = e2 - cc * P.wsz; wj = C.W8 + cc; }
      const int s = wi * P.wsz + wj;
      const int gq = g0 + gl;
      Group<CC> gp = shfl_group(mine, gl);
      const int ag = (REFINE && anchored) ? __shfl_sync(FULL_MASK, my_ag, gl) : 0;
      CC chh, cww;
      if (INT) {
        chh = (CC)((int)gp.ah + (int)s1c * (wi - (int)gp.offh));
        cww = (CC)((int)gp.aw + (int)s1c * (wj - (int)gp.offw));
      } else {
        chh = (CC)fmaf((float)s1c, (float)wi - (float)gp.offh, (float)gp.ah);
        cww = (CC)fmaf((float)s1c, (float)wj - (float)gp.offw, (float)gp.aw);
      }
      const bool cvalid = active && gp.ok && resident(gp.t, g) && inb(chh, g.kH) && inb(cww, g.kW);
      const int mh0 = C.n1 * wi, mw0 = C.n1 * wj;
      const unsigned hm = lds1(sb + (mskb + gl * 2) * 4) >> mh0, wm = lds1(sb + (mskb + gl * 2 + 1) * 4) >> mw0;
      const unsigned Tc = sb + (Tb + gl * C.GS4 + mh0 * C.Mp + mw0) * 16;
      const bool allpix = qfull && ((hm & C.rpat) == C.rpat) && ((wm & C.rpat) == C.rpat);
      float d = invalid;
      if (__all_sync(FULL_MASK, allpix || !cvalid)) {
        // every pixel pair of every live candidate is valid: straight-line code
        if (cvalid) {
          float dist = 0.f;
#pragma unroll
          for (int pi = 0; pi < ps; pi++) {
#pragma unroll
            for (int pj = 0; pj < ps; pj++) {
              const unsigned tp = Tc + pi * rstep + pj * cstep;
              const unsigned qp = qa + (pi * ps + pj) * Q4 * 16;
#pragma unroll
              for (int c4 = 0; c4 < Q4; c4++) dist = acc4<DIST>(dist, lds4(qp + c4 * 16), lds4(tp + c4 * pstep));
            }
          }
          d = dist;
        }
      } else if (cvalid) {
        d = dist_masked(Tc, hm, wm);
      }
      finish(active, gq, gp, ag, s, d);
    }
    }
    __syncwarp();
  }
  if (keep_all) return;

  // -- 4. selection: K rounds of warp-wide min over the keys (ties -> lower candidate index) --
  const int Klist = anchored ? P.Kout - 1 : P.Kout;
  const int start = lane + ((anchored && lane == 0) ? 32 : 0);
  // every lane keeps the best TWO of its keys (index ascending, strict <: ties stay with the lower index); a
  // lane re-reads its keys only when it has been the warp's winner twice since its last scan
  unsigned bk, bk2;
  int bi, bi2;
  auto scan2 = [&]() {
    bk = bk2 = 0xffffffffu;
    bi = bi2 = 0x7fffffff;
    for (int i = start; i < P.L; i += 32) {
      const unsigned kv = lds1(sb + (keyb + i) * 4);
      if (kv < bk) { bk2 = bk; bi2 = bi; bk = kv; bi = i; }
      else if (kv < bk2) { bk2 = kv; bi2 = i; }
    }
  };
  scan2();
  bool have2 = true;
  unsigned mykey = 0xffffffffu;
  int mypos = 0x7fffffff;
  for (int kk = 0; kk < Klist; kk++) {
    const unsigned m = __reduce_min_sync(FULL_MASK, bk);
    if (m == 0xffffffffu) break;
    const int widx = (int)__reduce_min_sync(FULL_MASK, (bk == m) ? (unsigned)bi : 0x7fffffffu);
    if (lane == kk) { mykey = m; mypos = widx; }
    if (bi == widx) {
      sts1(sb + (keyb + bi) * 4, 0xffffffffu);
      if (have2) { bk = bk2; bi = bi2; have2 = false; }
      else { scan2(); have2 = true; }
    }
  }

  // -- write the list: recompute the offsets of the kept candidates --
  if (lane < Klist && mypos != 0x7fffffff) {
    int eff = mypos;
    int64_t oi = obase + (anchored ? lane + 1 : lane);
    int gq = eff / P.G, s = eff - gq * P.G;
    Group<CC> gp = group_params<INT, REFINE>(P, c, gq);
    CC ind[3];
    bool done = false;
    if (anchored) {
      if (REFINE) {
        int ag = refine_group_anchor<CC>(P, c, gp, gq);
        if (s == 0 && ag != 0) { refine_loc<CC>(P, c, gq, ind); done = true; }
        else if (s == ag) s = 0;
      } else {
        if (eff == a_nls) { gq = 0; s = 0; gp = group_params<INT, REFINE>(P, c, 0); }
      }
    }
    if (!done) cand_inds<CC>(P, c, gp, s, ind);
    const float kf_ = ord2f(mykey);
    P.dists[oi] = (DIST == 0) ? -kf_ : kf_;
    inds_out[oi * 3] = ind[0]; inds_out[oi * 3 + 1] = ind[1]; inds_out[oi * 3 + 2] = ind[2];
    if (P.kselect) P.kselect[oi] = gp.ok ? gq : 1;
  }
}

template <bool INT, int DIST, bool REFINE, int PS, int Q4T>
int launch_tab_one(const FwdParams& P, const TabCfg& C, cudaStream_t st) {
  size_t smem = ((size_t)C.dec_f4 + (size_t)C.wpb * C.warp_f4) * 16;
  auto kern = search_fwd_tab_kernel<INT, DIST, REFINE, PS, Q4T>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return -1;
  }
  int64_t blocks = (P.nquery + C.wpb - 1) / C.wpb;
  kern<<<(unsigned)blocks, C.wpb * 32, smem, st>>>(P, C);
  return 0;
}

template <bool INT, int DIST, bool REFINE, int PS>
int launch_tab_q4(const FwdParams& P, const TabCfg& C, cudaStream_t st) {
  if (C.Q4 == 2) return launch_tab_one<INT, DIST, REFINE, PS, 2>(P, C, st);
  if (C.Q4 == 4) return launch_tab_one<INT, DIST, REFINE, PS, 4>(P, C, st);
  return launch_tab_one<INT, DIST, REFINE, PS, 0>(P, C, st);
}

template <bool INT, bool REFINE>
int launch_tab_all(const FwdParams& P, const TabCfg& C, cudaStream_t st) {
  const int ps = P.o.ps;
  if (P.o.dist_type == 0) {
    if (ps == 1) return launch_tab_q4<INT, 0, REFINE, 1>(P, C, st);
    if (ps == 3) return launch_tab_q4<INT, 0, REFINE, 3>(P, C, st);
    return launch_tab_q4<INT, 0, REFINE, 0>(P, C, st);
  }
  if (ps == 1) return launch_tab_q4<INT, 1, REFINE, 1>(P, C, st);
  if (ps == 3) return launch_tab_q4<INT, 1, REFINE, 3>(P, C, st);
  return launch_tab_q4<INT, 1, REFINE, 0>(P, C, st);
}

}  // namespace stnls_fwd
