// Search backward: gradients for q, k and (float) flows.  One thread per (query, channel quad);
// grad_q is accumulated over the K neighbours in registers and flushed once per patch pixel,
// grad_k is scattered with 16-byte vector REDs into the channels-last accumulator.
// Replaces:
//   non_local_search_int_vid_backward_kernel   lib/csrc/search/non_local_search_int_kernel.cu:273-355
//   nls_bwd_vid_kernel / nls_bwd_vidflows_kernel lib/csrc/search/non_local_search_bilin2d_kernel.cu:285-371,464-598
//   refinement_vidflows_backward_kernel        lib/csrc/search/refinement_bilin2d_kernel.cu:359-491
#include "common.cuh"

namespace {

struct BwdParams {
  float* gq;
  float* gk;
  float* gflows;
  const float* q;
  const float* k;
  const float* flows;
  const float* gd;
  const float* gi;
  const void* inds;
  const int* kselect;
  const uint8_t* reflect;
  stnls_geom g;
  stnls_search_opts o;
  int K, mode, nquad, W_t;
  int64_t nthreads;
};

// MODE 0: int ; 1: bilinear ; 2: bilinear + flow grads (NLS layout) ; 3: same, refine layout
template <int MODE, int DIST, int PS>
__global__ void __launch_bounds__(128) search_bwd_kernel(const __grid_constant__ BwdParams P) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P.nthreads) return;
  const int c4 = (int)(idx % P.nquad) * 4;
  int64_t wq = idx / P.nquad;
  int64_t r = wq;
  const int nw = (int)(r % g.nW); r /= g.nW;
  const int nh = (int)(r % g.nH); r /= g.nH;
  const int tl = (int)(r % g.Tq); r /= g.Tq;
  const int n = (int)r, hd = n % g.HD, b = n / g.HD;
  const int t = tl + g.tq0;
  const int ah = (nh * o.stride0) % g.kH, aw = (nw * o.stride0) % g.kW;
  const int rh = (MODE == 0) ? (nh * o.strideQ) % g.qH : ah;
  const int rw = (MODE == 0) ? (nw * o.strideQ) % g.qW : aw;
  const int hd_f = hd % o.HD_f;
  const int ps = (PS > 0) ? PS : o.ps;

  const float* qf = P.q + ((int64_t)n * g.Tq + tl) * g.qH * g.qW * g.Fp + c4;
  float* gqf = P.gq + ((int64_t)n * g.Tq + tl) * g.qH * g.qW * g.Fp + c4;
  const int64_t kfs = (int64_t)g.kH * g.kW * g.Fp;
  const float* kb = P.k + (int64_t)n * g.Tk * kfs + c4;
  float* gkb = P.gk + (int64_t)n * g.Tk * kfs + c4;

  float4 acc[(PS > 0) ? PS * PS : 1];
#pragma unroll
  for (int i = 0; i < ((PS > 0) ? PS * PS : 1); i++) acc[i] = f4zero();

  for (int ki = 0; ki < P.K; ki++) {
    const int64_t oi = wq * P.K + ki;
    const float weight = P.gd[oi];
    if (MODE == 0) {
      const int* ind = reinterpret_cast<const int*>(P.inds) + oi * 3;
      const int p0 = bnd(t + ind[0], g.T);
      const int ph = ah + ind[1], pw = aw + ind[2];
      if (!inb(p0, g.T)) continue;
      const float* kf = kb + (int64_t)(p0 - g.tk0) * kfs;
      float* gkf = gkb + (int64_t)(p0 - g.tk0) * kfs;
#pragma unroll
      for (int pi = 0; pi < ps; pi++) {
        int d = o.dilation * (pi + o.patch_offset);
        int r1 = rh + o.off_Hq + d, p1 = ph + d;
        if (o.reflect_bounds) { r1 = bnd(r1, g.qH); p1 = bnd(p1, g.kH); }
        if (!(inb(r1, g.qH) && inb(p1, g.kH))) continue;
#pragma unroll
        for (int pj = 0; pj < ps; pj++) {
          int e = o.dilation * (pj + o.patch_offset);
          int r2 = rw + o.off_Wq + e, p2 = pw + e;
          if (o.reflect_bounds) { r2 = bnd(r2, g.qW); p2 = bnd(p2, g.kW); }
          if (!(inb(r2, g.qW) && inb(p2, g.kW))) continue;
          const int64_t qo = ((int64_t)r1 * g.qW + r2) * g.Fp, ko = ((int64_t)p1 * g.kW + p2) * g.Fp;
          float4 x0 = ldg4(qf + qo), x1 = ldg4(kf + ko), a0, a1;
          if (DIST == 0) { a0 = f4scale(x1, weight); a1 = f4scale(x0, weight); }
          else {
            float w2 = 2 * weight;
            a0 = make_float4(w2 * (x0.x - x1.x), w2 * (x0.y - x1.y), w2 * (x0.z - x1.z), w2 * (x0.w - x1.w));
            a1 = make_float4(-a0.x, -a0.y, -a0.z, -a0.w);
          }
          if (PS > 0) { float4& A = acc[pi * PS + pj]; A.x += a0.x; A.y += a0.y; A.z += a0.z; A.w += a0.w; }
          else red_add4(gqf + qo, a0);
          red_add4(gkf + ko, a1);
        }
      }
      continue;
    }
    // ---------------- float modes ----------------
    const float* ind = reinterpret_cast<const float*>(P.inds) + oi * 3;
    const float pp0 = (float)t + ind[0], pph = (float)ah + ind[1], ppw = (float)aw + ind[2];
    float iw0 = 0.f, iw1 = 0.f;
    int sHc = 1, sWc = 1, kj = 0;
    int64_t ro = 0;
    if (MODE >= 2) { iw0 = P.gi[oi * 3 + 1]; iw1 = P.gi[oi * 3 + 2]; }
    if (MODE == 3) {
      kj = P.kselect[oi];
      ro = (((((int64_t)b * o.HD_f + hd_f) * g.Tq + tl) * g.nH + nh) * g.nW + nw) * o.Ks + kj;
      sHc = P.reflect[ro * 2] ? -1 : 1;
      sWc = P.reflect[ro * 2 + 1] ? -1 : 1;
      if (c4 == 0) {
        atomicAdd(P.gflows + ro * 3 + 1, sHc * iw0);
        atomicAdd(P.gflows + ro * 3 + 2, sWc * iw1);
      }
    }
    if (MODE >= 2) {
      if (!(inb(pp0, g.T) && inb(pph, g.kH) && inb(ppw, g.kW))) continue;
    }
    const float p0 = bnd(pp0, g.T);
    if (!inb(p0, g.T)) continue;
    const int pt0 = __float2int_rn(p0);
    const float* kf = kb + (int64_t)(pt0 - g.tk0) * kfs;
    float* gkf = gkb + (int64_t)(pt0 - g.tk0) * kfs;
    float dfw = 0.f, dfh = 0.f;  // accumulated d/d(flow_w), d/d(flow_h)
#pragma unroll
    for (int pi = 0; pi < ps; pi++) {
      int d = o.dilation * (pi + o.patch_offset);
      int r1 = rh + o.off_Hq + d;
      float p1 = pph + (float)d;
      const int signH = inb(p1, g.kH) ? 1 : -1;
      if (o.reflect_bounds) { r1 = bnd(r1, g.qH); p1 = bnd(p1, g.kH); }
      if (!(inb(r1, g.qH) && inb(p1, g.kH))) continue;
#pragma unroll
      for (int pj = 0; pj < ps; pj++) {
        int e = o.dilation * (pj + o.patch_offset);
        int r2 = rw + o.off_Wq + e;
        float p2 = ppw + (float)e;
        const int signW = inb(p2, g.kW) ? 1 : -1;
        if (o.reflect_bounds) { r2 = bnd(r2, g.qW); p2 = bnd(p2, g.kW); }
        if (!(inb(r2, g.qW) && inb(p2, g.kW))) continue;
        const Taps tp = make_taps(p1, p2, g.kH, g.kW);
        const int64_t qo = ((int64_t)r1 * g.qW + r2) * g.Fp;
        const float4 x0 = ldg4(qf + qo);
        const float4 x1 = interp4(kf - c4, tp, g.kW, g.Fp, c4);
        float4 a0, a1;
        if (DIST == 0) { a0 = f4scale(x1, weight); a1 = f4scale(x0, weight); }
        else {
          float w2 = weight * 2;
          a0 = make_float4(w2 * (x0.x - x1.x), w2 * (x0.y - x1.y), w2 * (x0.z - x1.z), w2 * (x0.w - x1.w));
          a1 = make_float4(-a0.x, -a0.y, -a0.z, -a0.w);
        }
        if (PS > 0) { float4& A = acc[pi * PS + pj]; A.x += a0.x; A.y += a0.y; A.z += a0.z; A.w += a0.w; }
        else red_add4(gqf + qo, a0);
#pragma unroll
        for (int ix = 0; ix < 2; ix++) {
#pragma unroll
          for (int jx = 0; jx < 2; jx++) {
            if (!tp.vh[ix] || !tp.vw[jx]) continue;
            const int64_t ko = ((int64_t)tp.h[ix] * g.kW + tp.w[jx]) * g.Fp;
            red_add4(gkf + ko, f4scale(a1, tp.gh[ix] * tp.gw[jx]));
            if (MODE >= 2) {  // update_dFlows, shared_kernel.cu:916-964
              const float4 v = ldg4(kf + ko);
              const float dot = v.x * a1.x + v.y * a1.y + v.z * a1.z + v.w * a1.w;
              const float gH = (jx == 0) ? -tp.gh[ix] : tp.gh[ix];
              const float gW = (ix == 0) ? -tp.gw[jx] : tp.gw[jx];
              dfw += gH * dot * signW;
              dfh += gW * dot * signH;
            }
          }
        }
      }
    }
    if (MODE == 2) {  // non_local_search_bilin2d_kernel.cu:555-594
      const int dt = (int)pp0 - t;
      const int dto = time_tmax(t, g.T, o.wt) - t;
      const int si = (dt >= 0) ? (dt - o.st_offset) : (dto - dt - o.st_offset);
      if (si < 0) continue;
      const int nF = P.W_t - o.st_offset;
      const int64_t fb = ((((int64_t)b * o.HD_f + hd_f) * g.Tq + tl) * nF + si) * 2;
      const int64_t plane = (int64_t)g.nH * g.nW, off = (int64_t)nh * g.nW + nw;
      const float hi = (float)ah + P.flows[(fb + 1) * plane + off];
      const float wi = (float)aw + P.flows[fb * plane + off];
      const int sH = inb(hi, g.kH) ? 1 : -1, sW = inb(wi, g.kW) ? 1 : -1;
      float addw = sW * dfw, addh = sH * dfh;
      if (c4 == 0) { addw += sW * iw1; addh += sH * iw0; }
      atomicAdd(P.gflows + fb * plane + off, addw);
      atomicAdd(P.gflows + (fb + 1) * plane + off, addh);
    } else if (MODE == 3) {  // bwd_flow_assign_v2, shared_kernel.cu:1001-1026
      atomicAdd(P.gflows + ro * 3 + 2, sWc * dfw);
      atomicAdd(P.gflows + ro * 3 + 1, sHc * dfh);
    }
  }

  if (PS > 0) {
#pragma unroll
    for (int pi = 0; pi < PS; pi++) {
      int r1 = rh + o.off_Hq + o.dilation * (pi + o.patch_offset);
      if (o.reflect_bounds) r1 = bnd(r1, g.qH);
      if (!inb(r1, g.qH)) continue;
#pragma unroll
      for (int pj = 0; pj < PS; pj++) {
        int r2 = rw + o.off_Wq + o.dilation * (pj + o.patch_offset);
        if (o.reflect_bounds) r2 = bnd(r2, g.qW);
        if (!inb(r2, g.qW)) continue;
        red_add4(gqf + ((int64_t)r1 * g.qW + r2) * g.Fp, acc[pi * PS + pj]);
      }
    }
  }
}

template <int MODE, int DIST>
void launch_bwd(const BwdParams& P, cudaStream_t st) {
  const int tpb = 128;
  unsigned blocks = (unsigned)((P.nthreads + tpb - 1) / tpb);
  if (P.o.ps == 1) search_bwd_kernel<MODE, DIST, 1><<<blocks, tpb, 0, st>>>(P);
  else if (P.o.ps == 3) search_bwd_kernel<MODE, DIST, 3><<<blocks, tpb, 0, st>>>(P);
  else search_bwd_kernel<MODE, DIST, 0><<<blocks, tpb, 0, st>>>(P);
}

}  // namespace

extern "C" int stnls_search_backward(float* gq_cl, float* gk_cl, float* grad_flows, const float* q_cl,
                                     const float* k_cl, const float* flows, const float* grad_dists,
                                     const float* grad_inds, const void* inds, const int* kselect,
                                     const uint8_t* reflect, int K, const stnls_geom* g,
                                     const stnls_search_opts* o, stnls_stream_t stream) {
  STNLS_REQUIRE(gq_cl && gk_cl && q_cl && k_cl && grad_dists && inds && g && o, "search_backward: null argument");
  STNLS_REQUIRE(g->Fp % 4 == 0, "search_backward: Fp must be a multiple of 4");
  STNLS_REQUIRE(o->dist_type == 0 || o->dist_type == 1, "Uknown distance type. Must be 0 (product) or 1 (l2)");
  BwdParams P;
  P.gq = gq_cl; P.gk = gk_cl; P.gflows = grad_flows; P.q = q_cl; P.k = k_cl; P.flows = flows;
  P.gd = grad_dists; P.gi = grad_inds; P.inds = inds; P.kselect = kselect; P.reflect = reflect;
  P.g = *g; P.o = *o; P.K = K; P.nquad = g->Fp / 4; P.W_t = 2 * o->wt + 1;
  const bool refine = o->wr > 0;
  if (o->itype_int) P.mode = 0;
  else if (!grad_flows) P.mode = 1;
  else P.mode = refine ? 3 : 2;
  if (P.mode >= 2) STNLS_REQUIRE(grad_inds != nullptr, "search_backward: grad_inds required with grad_flows");
  if (P.mode == 2) STNLS_REQUIRE(flows != nullptr, "search_backward: flows required with grad_flows");
  if (P.mode == 3) STNLS_REQUIRE(kselect && reflect, "search_backward: kselect/reflect required for refine flow grads");
  P.nthreads = (int64_t)g->B * g->HD * g->Tq * g->nH * g->nW * P.nquad;
  if (P.nthreads == 0 || K == 0) return STNLS_OK;
  STNLS_REQUIRE(P.nthreads / 128 < 2147483647LL, "search_backward: too many threads for one launch");
  cudaStream_t st = (cudaStream_t)stream;
  switch (P.mode * 2 + (o->dist_type ? 1 : 0)) {
    case 0: launch_bwd<0, 0>(P, st); break;
    case 1: launch_bwd<0, 1>(P, st); break;
    case 2: launch_bwd<1, 0>(P, st); break;
    case 3: launch_bwd<1, 1>(P, st); break;
    case 4: launch_bwd<2, 0>(P, st); break;
    case 5: launch_bwd<2, 1>(P, st); break;
    case 6: launch_bwd<3, 0>(P, st); break;
    case 7: launch_bwd<3, 1>(P, st); break;
  }
  STNLS_CHECK_LAUNCH("search_backward");
  return STNLS_OK;
}
