// Search backward: gradients for q, k and (float) flows.  One thread per (query, channel quad);
// grad_q is accumulated over the K neighbours in registers and flushed once per patch pixel,
// grad_k is scattered with 16-byte vector REDs into the channels-last accumulator.
// Replaces:
//   non_local_search_int_vid_backward_kernel   lib/csrc/search/non_local_search_int_kernel.cu:273-355
//   nls_bwd_vid_kernel / nls_bwd_vidflows_kernel lib/csrc/search/non_local_search_bilin2d_kernel.cu:285-371,464-598
//   refinement_vidflows_backward_kernel        lib/csrc/search/refinement_bilin2d_kernel.cu:359-491
#include "common.cuh"
#ifndef SB_MINB
#define SB_MINB 4
#endif
#include <cstdlib>

namespace {

__device__ __forceinline__ float dot4(const float4 a, const float4 b) {
  return fmaf(a.w, b.w, fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)));
}
__device__ __forceinline__ void fma4(float4& y, float a, const float4 x) {
  y.x = fmaf(a, x.x, y.x); y.y = fmaf(a, x.y, y.y); y.z = fmaf(a, x.z, y.z); y.w = fmaf(a, x.w, y.w);
}

struct BwdParams {
  float* gq;
  float* gk;
  float* gflows;
  const float* q;
  const float* k;
  const float* flows;
  const float* gd;
  const float* gw;       // fused normalise backward: grad_dists is not read, it is computed from the
  const float* w;        //   gradient of the weights and the weights: gd[k] = sm_scale * w[k] * (gw[k] - sum_j w[j] gw[j])
  float sm_scale;
  const float* gi;
  const void* inds;
  const int* kselect;
  const uint8_t* reflect;
  stnls_geom g;
  stnls_search_opts o;
  int K, mode, nquad, W_t, fast3, ksplit, qagg, rolled;
  int idx32;                                  // launch shape of frame_grid(): thread index decoded with the FastDivs
  unsigned per_frame, frames;
  unsigned gx, gy, gz;
  FastDiv fquad, fK, fnW, fTq, fHD;
  int64_t nthreads;
};

// MODE 0: int ; 1: bilinear ; 2: bilinear + flow grads (NLS layout) ; 3: same, refine layout
// PS > 0: compile-time patch size (grad_q of the patch is kept in registers over the K neighbours)
// KSPLIT: one thread per (query, neighbour, quad) instead of a loop over the K neighbours
template <int MODE, int DIST, int PS, bool FAST, bool KSPLIT>
__global__ void __launch_bounds__(128, (FAST && !KSPLIT) ? 3 : SB_MINB) search_bwd_kernel(const __grid_constant__ BwdParams P) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  int c4, nw, nh, tl, hd, n, b;
  int64_t wq;
  int k_lo = 0, k_hi = P.K;
  if (P.idx32) {   // grid.x = the threads of one frame, grid.y/z = frames (common.cuh frame_grid): multiply-high divisions
    unsigned u = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned fr = blockIdx.y + blockIdx.z * gridDim.y;
    if (u >= P.per_frame || fr >= P.frames) return;
    u = fdivmod(u, P.fquad, c4);
    if (KSPLIT) { u = fdivmod(u, P.fK, k_lo); k_hi = k_lo + 1; }
    nh = (int)fdivmod(u, P.fnW, nw);
    n = (int)fdivmod(fr, P.fTq, tl);
    b = (int)fdivmod((unsigned)n, P.fHD, hd);
    wq = ((int64_t)fr * g.nH + nh) * g.nW + nw;
  } else {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.nthreads) return;
    wq = divmod(idx, P.nquad, c4);
    if (KSPLIT) { wq = divmod(wq, P.K, k_lo); k_hi = k_lo + 1; }
    int64_t r = divmod(wq, g.nW, nw);
    r = divmod(r, g.nH, nh);
    r = divmod(r, g.Tq, tl);
    n = (int)r;
    b = (int)divmod(r, g.HD, hd);
  }
  c4 *= 4;
  const int t = tl + g.tq0;
  const int ah = (nh * o.stride0) % g.kH, aw = (nw * o.stride0) % g.kW;
  const int rh = (MODE == 0) ? (nh * o.strideQ) % g.qH : ah;
  const int rw = (MODE == 0) ? (nw * o.strideQ) % g.qW : aw;
  const int hd_f = hd % o.HD_f;
  const int ps = (PS > 0) ? PS : o.ps;
  const unsigned qmask = P.qagg ? ((P.nquad >= 32 ? 0xffffffffu : ((1u << P.nquad) - 1u)) << ((threadIdx.x & 31) & ~(P.nquad - 1))) : 0u;

  const float* qf = P.q + ((int64_t)n * g.Tq + tl) * g.qH * g.qW * g.Fp + c4;
  float* gqf = P.gq + ((int64_t)n * g.Tq + tl) * g.qH * g.qW * g.Fp + c4;
  const int64_t kfs = (int64_t)g.kH * g.kW * g.Fp;
  const float* kb = P.k + (int64_t)n * g.Tk * kfs + c4;
  float* gkb = P.gk + (int64_t)n * g.Tk * kfs + c4;

  // grad_q of the query patch: kept in registers over the K neighbours (loop form) or sent straight
  // to memory (KSPLIT: one neighbour per thread)
  constexpr bool ACCQ = (PS > 0) && !KSPLIT;
  float4 acc[ACCQ ? PS * PS : 1];
#pragma unroll
  for (int i = 0; i < (ACCQ ? PS * PS : 1); i++) acc[i] = f4zero();

  // grad_dists of neighbour oi: read, or (fused normalise backward) the arithmetic of softmax_bwd_kernel (normz.cu) in
  // the same order, so both routes give the same bits
  const bool smb = P.gw != nullptr;
  float sm_dot = 0.f;
  if (smb) {
    const float* wr = P.w + wq * P.K;
    const float* gr = P.gw + wq * P.K;
    for (int j = 0; j < P.K; j++) sm_dot = fmaf(__ldg(wr + j), __ldg(gr + j), sm_dot);
  }
  auto gdist = [&](int64_t oi) -> float {
    if (!smb) return P.gd[oi];
    return __fmul_rn(__fmul_rn(P.sm_scale, __ldg(P.w + oi)), __fsub_rn(__ldg(P.gw + oi), sm_dot));
  };

  if (MODE == 0) {  // nls_int.cu:88-167
    for (int ki = k_lo; ki < k_hi; ki++) {
      const int64_t oi = wq * P.K + ki;
      const float weight = gdist(oi);
      const int* ind = reinterpret_cast<const int*>(P.inds) + oi * 3;
      const int p0 = bnd(t + ind[0], g.T);
      const int ph = ah + ind[1], pw = aw + ind[2];
      if (!resident(p0, g)) continue;
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
          if (ACCQ) { float4& A = acc[ACCQ ? pi * PS + pj : 0]; A.x += a0.x; A.y += a0.y; A.z += a0.z; A.w += a0.w; }
          else red_add4(gqf + qo, a0);
          red_add4(gkf + ko, a1);
        }
      }
    }
  } else {
    // ---------------- float modes ----------------
    struct Pre {
      float weight, pp0, pph, ppw, iw0, iw1;
      int sHc, sWc, pt0;
      int64_t ro;
      bool valid;
    };
    // per-neighbour preamble; `first` = do the once-only index-gradient atomics of the refine layout
    auto preamble = [&](int ki, bool first) {
      Pre z;
      const int64_t oi = wq * P.K + ki;
      z.weight = gdist(oi);
      const float* ind = reinterpret_cast<const float*>(P.inds) + oi * 3;
      z.pp0 = (float)t + ind[0]; z.pph = (float)ah + ind[1]; z.ppw = (float)aw + ind[2];
      z.iw0 = z.iw1 = 0.f;
      z.sHc = z.sWc = 1;
      z.ro = 0;
      z.pt0 = 0;
      z.valid = false;
      if (MODE >= 2) { z.iw0 = P.gi[oi * 3 + 1]; z.iw1 = P.gi[oi * 3 + 2]; }
      if (MODE == 3) {
        const int kj = P.kselect[oi];
        z.ro = (((((int64_t)b * o.HD_f + hd_f) * g.Tq + tl) * g.nH + nh) * g.nW + nw) * o.Ks + kj;
        z.sHc = P.reflect[z.ro * 2] ? -1 : 1;
        z.sWc = P.reflect[z.ro * 2 + 1] ? -1 : 1;
        if (first && c4 == 0) {
          atomicAdd(P.gflows + z.ro * 3 + 1, z.sHc * z.iw0);
          atomicAdd(P.gflows + z.ro * 3 + 2, z.sWc * z.iw1);
        }
      }
      if (MODE >= 2) {
        if (!(inb(z.pp0, g.T) && inb(z.pph, g.kH) && inb(z.ppw, g.kW))) return z;
      }
      const float p0 = bnd(z.pp0, g.T);
      z.pt0 = __float2int_rn(p0);
      if (!inb(p0, g.T) || !resident(z.pt0, g)) return z;
      z.valid = true;
      return z;
    };
    // flow-gradient tail of one neighbour (non_local_search_bilin2d_kernel.cu:555-594 ; shared_kernel.cu:1001-1026)
    auto tail = [&](const Pre& z, float dfw, float dfh) {
      if (MODE >= 2 && P.qagg) {   // one pair of atomics per (query, neighbour): sum the channel-quad lanes first
        for (int off = P.nquad >> 1; off > 0; off >>= 1) {
          dfw += __shfl_xor_sync(qmask, dfw, off);
          dfh += __shfl_xor_sync(qmask, dfh, off);
        }
        if (c4 != 0) return;
      }
      if (MODE == 2) {
        const int dt = (int)z.pp0 - t;
        const int dto = time_tmax(t, g.T, o.wt) - t;
        const int si = (dt >= 0) ? (dt - o.st_offset) : (dto - dt - o.st_offset);
        if (si < 0) return;
        const int nF = P.W_t - o.st_offset;
        const int64_t fb = ((((int64_t)b * o.HD_f + hd_f) * g.Tq + tl) * nF + si) * 2;
        const int64_t plane = (int64_t)g.nH * g.nW, off = (int64_t)nh * g.nW + nw;
        const float hi = (float)ah + P.flows[(fb + 1) * plane + off];
        const float wi = (float)aw + P.flows[fb * plane + off];
        const int sH = inb(hi, g.kH) ? 1 : -1, sW = inb(wi, g.kW) ? 1 : -1;
        float addw = sW * dfw, addh = sH * dfh;
        if (c4 == 0) { addw += sW * z.iw1; addh += sH * z.iw0; }
        atomicAdd(P.gflows + fb * plane + off, addw);
        atomicAdd(P.gflows + (fb + 1) * plane + off, addh);
      } else if (MODE == 3) {
        atomicAdd(P.gflows + z.ro * 3 + 2, z.sWc * dfw);
        atomicAdd(P.gflows + z.ro * 3 + 1, z.sHc * dfh);
      }
    };
    // generic patch loop: every pixel validated / reflected on its own (nls_bilin2d.cu:95-291)
    auto generic = [&](const Pre& z) {
      const float weight = z.weight;
      const float* kf = kb + (int64_t)(z.pt0 - g.tk0) * kfs;
      float* gkf = gkb + (int64_t)(z.pt0 - g.tk0) * kfs;
      float dfw = 0.f, dfh = 0.f;
#pragma unroll
      for (int pi = 0; pi < ps; pi++) {
        int d = o.dilation * (pi + o.patch_offset);
        int r1 = rh + o.off_Hq + d;
        float p1 = z.pph + (float)d;
        const int signH = inb(p1, g.kH) ? 1 : -1;
        if (o.reflect_bounds) { r1 = bnd(r1, g.qH); p1 = bnd(p1, g.kH); }
        if (!(inb(r1, g.qH) && inb(p1, g.kH))) continue;
#pragma unroll
        for (int pj = 0; pj < ps; pj++) {
          int e = o.dilation * (pj + o.patch_offset);
          int r2 = rw + o.off_Wq + e;
          float p2 = z.ppw + (float)e;
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
          if (ACCQ) { float4& A = acc[ACCQ ? pi * PS + pj : 0]; A.x += a0.x; A.y += a0.y; A.z += a0.z; A.w += a0.w; }
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
      tail(z, dfw, dfh);
    };

    // footprint fast path (ps == 3, dilation == 1): the 9 bilinear reads of an interior neighbour
    // patch share a 4x4 pixel footprint -> 16 loads + 16 vector REDs instead of 36 + 36.  Pairs that
    // do not qualify are deferred (bit mask) so a warp does not run both paths for every k.
    const int po = o.patch_offset;
    const bool fast3 = FAST && (PS == 3) && P.fast3 && (P.K <= 32) && (o.off_Hq == 0) && (o.off_Wq == 0) &&
                       (rh + po >= 0) && (rh + po + 2 <= g.qH - 1) && (rw + po >= 0) && (rw + po + 2 <= g.qW - 1);
    if (!fast3) {
      for (int ki = k_lo; ki < k_hi; ki++) {
        const Pre z = preamble(ki, true);
        if (z.valid) generic(z);
      }
    } else {
      unsigned slow = 0;
      for (int ki = k_lo; ki < k_hi; ki++) {
        const Pre z = preamble(ki, true);
        if (!z.valid) continue;
        const int h0 = __float2int_rz(z.pph + (float)po), w0 = __float2int_rz(z.ppw + (float)po);
        float gh[3][2], gw[3][2];
        bool ok = true;
#pragma unroll
        for (int p = 0; p < 3; p++) {
          const float n1 = z.pph + (float)(p + po), n2 = z.ppw + (float)(p + po);
          const int a0 = __float2int_rz(n1), a1 = __float2int_rz(n1 + 1);
          const int b0 = __float2int_rz(n2), b1 = __float2int_rz(n2 + 1);
          ok = ok && (n1 >= 0) && (a0 == h0 + p) && (a1 == h0 + p + 1) && (a1 <= g.kH - 1);
          ok = ok && (n2 >= 0) && (b0 == w0 + p) && (b1 == w0 + p + 1) && (b1 <= g.kW - 1);
          gh[p][0] = fmaxf(0.f, 1 - fabsf(a0 - n1)); gh[p][1] = fmaxf(0.f, 1 - fabsf(a1 - n1));
          gw[p][0] = fmaxf(0.f, 1 - fabsf(b0 - n2)); gw[p][1] = fmaxf(0.f, 1 - fabsf(b1 - n2));
        }
        if (!ok) { slow |= 1u << ki; continue; }
        const float weight = z.weight;
        const int64_t fo = (int64_t)(z.pt0 - g.tk0) * kfs + ((int64_t)h0 * g.kW + w0) * g.Fp;
        const float* vr = kb + fo;
        float* gr_ = gkb + fo;
        const int64_t rs = (int64_t)g.kW * g.Fp;
        float dfw = 0.f, dfh = 0.f;
        float4 A[4], B[4], accA[4], accB[4];
#pragma unroll
        for (int j = 0; j < 4; j++) { A[j] = ldg4(vr + j * g.Fp); accA[j] = f4zero(); }
        if (KSPLIT && P.rolled) {
          // one neighbour per thread: the three patch rows as a ROLLED loop -- a third of the straight-line code, for
          // a launch whose warps each run the path once (instruction fetch, not reuse, feeds them)
#pragma unroll 1
          for (int pi = 0; pi < 3; pi++) {
            const float n1 = z.pph + (float)(pi + po);
            const int a0 = __float2int_rz(n1);
            const float g0 = fmaxf(0.f, 1 - fabsf(a0 - n1)), g1 = fmaxf(0.f, 1 - fabsf((a0 + 1) - n1));
#pragma unroll
            for (int j = 0; j < 4; j++) { B[j] = ldg4(vr + (pi + 1) * rs + j * g.Fp); accB[j] = f4zero(); }
#pragma unroll
            for (int pj = 0; pj < 3; pj++) {
              const int64_t qo = ((int64_t)(rh + pi + po) * g.qW + (rw + pj + po)) * g.Fp;
              const float4 x0 = ldg4(qf + qo);
              const float w00 = g0 * gw[pj][0], w01 = g0 * gw[pj][1];
              const float w10 = g1 * gw[pj][0], w11 = g1 * gw[pj][1];
              float4 x1 = f4zero();
              fma4(x1, w00, A[pj]); fma4(x1, w01, A[pj + 1]); fma4(x1, w10, B[pj]); fma4(x1, w11, B[pj + 1]);
              float4 a0_, a1;
              if (DIST == 0) { a0_ = f4scale(x1, weight); a1 = f4scale(x0, weight); }
              else {
                const float w2 = weight * 2;
                a0_ = make_float4(w2 * (x0.x - x1.x), w2 * (x0.y - x1.y), w2 * (x0.z - x1.z), w2 * (x0.w - x1.w));
                a1 = make_float4(-a0_.x, -a0_.y, -a0_.z, -a0_.w);
              }
              red_add4(gqf + qo, a0_);
              fma4(accA[pj], w00, a1); fma4(accA[pj + 1], w01, a1);
              fma4(accB[pj], w10, a1); fma4(accB[pj + 1], w11, a1);
              if (MODE >= 2) {
                const float d00 = dot4(A[pj], a1), d01 = dot4(A[pj + 1], a1);
                const float d10 = dot4(B[pj], a1), d11 = dot4(B[pj + 1], a1);
                dfw += g0 * (d01 - d00) + g1 * (d11 - d10);
                dfh += gw[pj][0] * (d10 - d00) + gw[pj][1] * (d11 - d01);
              }
            }
#pragma unroll
            for (int j = 0; j < 4; j++) { red_add4(gr_ + pi * rs + j * g.Fp, accA[j]); A[j] = B[j]; accA[j] = accB[j]; }
          }
#pragma unroll
          for (int j = 0; j < 4; j++) red_add4(gr_ + 3 * rs + j * g.Fp, accA[j]);
          tail(z, dfw, dfh);
          continue;
        }
#pragma unroll
        for (int pi = 0; pi < 3; pi++) {
#pragma unroll
          for (int j = 0; j < 4; j++) { B[j] = ldg4(vr + (pi + 1) * rs + j * g.Fp); accB[j] = f4zero(); }
#pragma unroll
          for (int pj = 0; pj < 3; pj++) {
            const int64_t qo = ((int64_t)(rh + pi + po) * g.qW + (rw + pj + po)) * g.Fp;
            const float4 x0 = ldg4(qf + qo);
            const float w00 = gh[pi][0] * gw[pj][0], w01 = gh[pi][0] * gw[pj][1];
            const float w10 = gh[pi][1] * gw[pj][0], w11 = gh[pi][1] * gw[pj][1];
            float4 x1 = f4zero();
            fma4(x1, w00, A[pj]); fma4(x1, w01, A[pj + 1]); fma4(x1, w10, B[pj]); fma4(x1, w11, B[pj + 1]);
            float4 a0, a1;
            if (DIST == 0) { a0 = f4scale(x1, weight); a1 = f4scale(x0, weight); }
            else {
              const float w2 = weight * 2;
              a0 = make_float4(w2 * (x0.x - x1.x), w2 * (x0.y - x1.y), w2 * (x0.z - x1.z), w2 * (x0.w - x1.w));
              a1 = make_float4(-a0.x, -a0.y, -a0.z, -a0.w);
            }
            if (ACCQ) { float4& Aq = acc[ACCQ ? pi * 3 + pj : 0]; Aq.x += a0.x; Aq.y += a0.y; Aq.z += a0.z; Aq.w += a0.w; }
            else red_add4(gqf + qo, a0);
            fma4(accA[pj], w00, a1); fma4(accA[pj + 1], w01, a1);
            fma4(accB[pj], w10, a1); fma4(accB[pj + 1], w11, a1);
            if (MODE >= 2) {  // update_dFlows, shared_kernel.cu:916-964 (all signs +1 in the interior)
              const float d00 = dot4(A[pj], a1), d01 = dot4(A[pj + 1], a1);
              const float d10 = dot4(B[pj], a1), d11 = dot4(B[pj + 1], a1);
              dfw += gh[pi][0] * (d01 - d00) + gh[pi][1] * (d11 - d10);
              dfh += gw[pj][0] * (d10 - d00) + gw[pj][1] * (d11 - d01);
            }
          }
#pragma unroll
          for (int j = 0; j < 4; j++) { red_add4(gr_ + pi * rs + j * g.Fp, accA[j]); A[j] = B[j]; accA[j] = accB[j]; }
        }
#pragma unroll
        for (int j = 0; j < 4; j++) red_add4(gr_ + 3 * rs + j * g.Fp, accA[j]);
        tail(z, dfw, dfh);
      }
      while (slow) {
        const int ki = __ffs(slow) - 1;
        slow &= slow - 1;
        const Pre z = preamble(ki, false);
        if (z.valid) generic(z);
      }
    }
  }

  if (ACCQ) {
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
        red_add4(gqf + ((int64_t)r1 * g.qW + r2) * g.Fp, acc[ACCQ ? pi * PS + pj : 0]);
      }
    }
  }
}


template <int MODE, int DIST, bool KSPLIT>
void launch_bwd_k(const BwdParams& P, cudaStream_t st) {
  const int tpb = 128;
  const dim3 blocks(P.gx, P.gy, P.gz);
  if (P.o.ps == 1) search_bwd_kernel<MODE, DIST, 1, false, KSPLIT><<<blocks, tpb, 0, st>>>(P);
  else if (P.o.ps == 3 && P.fast3 && MODE != 0) search_bwd_kernel<MODE, DIST, 3, true, KSPLIT><<<blocks, tpb, 0, st>>>(P);
  else if (P.o.ps == 3) search_bwd_kernel<MODE, DIST, 3, false, KSPLIT><<<blocks, tpb, 0, st>>>(P);
  else search_bwd_kernel<MODE, DIST, 0, false, KSPLIT><<<blocks, tpb, 0, st>>>(P);
}
template <int MODE, int DIST>
void launch_bwd(const BwdParams& P, cudaStream_t st) {
  if (P.ksplit) launch_bwd_k<MODE, DIST, true>(P, st);
  else launch_bwd_k<MODE, DIST, false>(P, st);
}

}  // namespace

static int search_backward_impl(float* gq_cl, float* gk_cl, float* grad_flows, const float* q_cl,
                                const float* k_cl, const float* flows, const float* grad_dists,
                                const float* grad_inds, const void* inds, const int* kselect,
                                const uint8_t* reflect, int K, const stnls_geom* g,
                                const stnls_search_opts* o, stnls_stream_t stream, const float* grad_weights,
                                const float* weights, float sm_scale) {
  STNLS_REQUIRE(gq_cl && gk_cl && q_cl && k_cl && (grad_dists || (grad_weights && weights)) && inds && g && o,
                "search_backward: null argument");
  STNLS_REQUIRE(g->Fp % 4 == 0, "search_backward: Fp must be a multiple of 4");
  STNLS_REQUIRE(o->dist_type == 0 || o->dist_type == 1, "Uknown distance type. Must be 0 (product) or 1 (l2)");
  if (int rc = stnls_check_geom(g, -1, "search_backward")) return rc;
  BwdParams P;
  P.gq = gq_cl; P.gk = gk_cl; P.gflows = grad_flows; P.q = q_cl; P.k = k_cl; P.flows = flows;
  P.gd = grad_dists; P.gw = grad_dists ? nullptr : grad_weights; P.w = weights; P.sm_scale = sm_scale; P.gi = grad_inds; P.inds = inds; P.kselect = kselect; P.reflect = reflect;
  P.g = *g; P.o = *o; P.K = K; P.nquad = g->Fp / 4; P.W_t = 2 * o->wt + 1;
  P.fast3 = (o->ps == 3 && o->dilation == 1 && !getenv("STNLS_NO_FAST3")) ? 1 : 0;
  const bool refine = o->wr > 0;
  if (o->itype_int) P.mode = 0;
  else if (!grad_flows) P.mode = 1;
  else P.mode = refine ? 3 : 2;
  if (P.mode >= 2) STNLS_REQUIRE(grad_inds != nullptr, "search_backward: grad_inds required with grad_flows");
  if (P.mode == 2) STNLS_REQUIRE(flows != nullptr, "search_backward: flows required with grad_flows");
  if (P.mode == 3) STNLS_REQUIRE(kselect && reflect, "search_backward: kselect/reflect required for refine flow grads");
  P.nthreads = (int64_t)g->B * g->HD * g->Tq * g->nH * g->nW * P.nquad;
  P.qagg = (P.nquad > 1 && P.nquad <= 32 && (P.nquad & (P.nquad - 1)) == 0 && getenv("STNLS_QAGG")) ? 1 : 0;   // opt-in: measured SLOWER
  // (B200, r02_l vs r02_g: wpsum_backward c4s 74.9 -> 83.8 ms, search_backward 91.9 -> 98.5 ms; C2 0.294 -> 0.309 ms): the
  // per-pair atomics were never the limiter, the butterflies and the convergence they force are pure overhead
  if (P.nthreads == 0 || K == 0) return STNLS_OK;
  // float modes, small launches: one thread per (query, neighbour, quad)
  // measured on B200 (profiles/README.md): with ps=3 the split form wins while the whole problem is a few
  // waves (C2: 0.65 -> 0.53 ms) and loses once the launch is large (1080p/4: 91 ms loop vs 132 ms split)
  const char* env_ks = getenv("STNLS_SB_KSPLIT");
  P.ksplit = (P.mode != 0 && o->ps >= 2 && P.nthreads < (1 << 20)) ? 1 : 0;
  if (env_ks) P.ksplit = (env_ks[0] == '1' && P.mode != 0) ? 1 : 0;
  if (P.ksplit) P.nthreads *= K;
  { const char* e = getenv("STNLS_SB_ROLLED"); P.rolled = (e && e[0] == '0') ? 0 : 1; }   // C2: 0.551 -> 0.513 ms
  STNLS_REQUIRE(P.nthreads / 128 < 2147483647LL, "search_backward: too many threads for one launch");
  {
    const int64_t per_frame = (int64_t)g->nH * g->nW * P.nquad * (P.ksplit ? K : 1), frames = (int64_t)g->B * g->HD * g->Tq;
    dim3 grid;
    P.idx32 = (!getenv("STNLS_NO_FRAMEGRID") && frame_grid(per_frame, frames, 128, grid)) ? 1 : 0;
    if (!P.idx32) grid = dim3((unsigned)((P.nthreads + 127) / 128));
    P.gx = grid.x; P.gy = grid.y; P.gz = grid.z;
    P.per_frame = (unsigned)per_frame; P.frames = (unsigned)frames;
    P.fquad = make_fastdiv(P.nquad); P.fK = make_fastdiv(K); P.fnW = make_fastdiv(g->nW);
    P.fTq = make_fastdiv(g->Tq); P.fHD = make_fastdiv(g->HD);
  }
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

extern "C" int stnls_search_backward(float* gq_cl, float* gk_cl, float* grad_flows, const float* q_cl,
                                     const float* k_cl, const float* flows, const float* grad_dists,
                                     const float* grad_inds, const void* inds, const int* kselect,
                                     const uint8_t* reflect, int K, const stnls_geom* g,
                                     const stnls_search_opts* o, stnls_stream_t stream) {
  STNLS_REQUIRE(grad_dists, "search_backward: null argument");
  return search_backward_impl(gq_cl, gk_cl, grad_flows, q_cl, k_cl, flows, grad_dists, grad_inds, inds, kselect, reflect, K,
                              g, o, stream, nullptr, nullptr, 0.f);
}

extern "C" int stnls_search_backward_sm(float* gq_cl, float* gk_cl, float* grad_flows, const float* q_cl,
                                        const float* k_cl, const float* flows, const float* grad_weights,
                                        const float* weights, float scale, const float* grad_inds, const void* inds,
                                        const int* kselect, const uint8_t* reflect, int K, const stnls_geom* g,
                                        const stnls_search_opts* o, stnls_stream_t stream) {
  STNLS_REQUIRE(grad_weights && weights, "search_backward_sm: null argument");
  return search_backward_impl(gq_cl, gk_cl, grad_flows, q_cl, k_cl, flows, nullptr, grad_inds, inds, kselect, reflect, K, g, o,
                              stream, grad_weights, weights, scale);
}
