// Weighted patch sum (== NonLocalGatherAdd) forward / backward on the channels-last layout.
// Replaces:
//   gather_add_forward_kernel          lib/csrc/agg/gather_add_kernel.cu:15-136
//   gather_add_int_backward_kernel     lib/csrc/agg/gather_add_kernel.cu:201-319
//   gather_add_bilin2d_backward_kernel lib/csrc/agg/gather_add_kernel.cu:369-514 (+ shared_kernel.cu:773-827)
// The forward is written as a GATHER (one thread per output pixel and channel quad, looping over the
// <= ps*ps covering queries and their K neighbours): no atomics, no zero-fill, the division by the
// patch-overlap count fused in.  The backward keeps grad_weights / grad_flows in registers and
// scatters grad_v with 16-byte vector REDs.
#include "common.cuh"
#ifndef WB_MINB
#define WB_MINB 4
#endif
#include <cstdlib>

namespace {

struct AggParams {
  float* out;        // fwd: out_cl ; bwd: gv_cl
  float* gweights;
  float* gflows;
  const float* gout;
  const float* v;
  const float* weights;
  const void* flows;
  const int* counts_h;
  const int* counts_w;
  stnls_geom g;
  stnls_agg_opts o;
  int nquad;
  int stack;         // 1: NonLocalGather -- no sum over K, slab (n*K + k) of out / gout belongs to neighbour k
  int qagg;          // backward: channel quads are a power of two -> per-pair sums are reduced across the quad lanes
  const float* dists;   // forward, fused normalise step: weights = softmax(sm_scale * dists[row, :K]) computed in the
  float* weights_out;   //   kernel (and written here for the backward); rows of dists have Kin >= K entries
  float sm_scale;
  int Kin;
  int idx32;         // backward: launch shape of frame_grid() (common.cuh), thread index decoded with the FastDivs
  unsigned per_frame, frames;
  FastDiv fquad, fK, fnW, fTq;
  int64_t nthreads;
};

template <bool INT>
__global__ void __launch_bounds__(128) wpsum_fwd_kernel(const __grid_constant__ AggParams P) {
  const stnls_geom& g = P.g;
  const stnls_agg_opts& o = P.o;
  const int H = g.kH, W = g.kW;
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P.nthreads) return;
  const int c4 = (int)(idx % P.nquad) * 4;
  int64_t r = idx / P.nquad;
  const int w = (int)(r % W); r /= W;
  const int h = (int)(r % H); r /= H;
  const int tl = (int)(r % g.Tq); r /= g.Tq;
  int k_lo = 0, k_hi = o.K;
  if (P.stack) { k_lo = (int)(r % o.K); k_hi = k_lo + 1; r /= o.K; }   // thread order [n][k][t][h][w][quad]
  const int n = (int)r;
  const int t = tl + g.tq0;
  const int64_t vfs = (int64_t)H * W * g.Fp;
  const float* vb = P.v + (int64_t)n * g.Tk * vfs + c4;
  float4 acc = f4zero();
  for (int pi = 0; pi < o.ps; pi++) {
    const int d = o.dilation * (pi + o.patch_offset);
    const int hq = h - d;
    if (hq < 0 || (hq % o.stride0) != 0) continue;
    const int nh = hq / o.stride0;
    if (nh >= g.nH) continue;
    for (int pj = 0; pj < o.ps; pj++) {
      const int e = o.dilation * (pj + o.patch_offset);
      const int wq = w - e;
      if (wq < 0 || (wq % o.stride0) != 0) continue;
      const int nw = wq / o.stride0;
      if (nw >= g.nW) continue;
      const int64_t qb = ((((int64_t)n * g.Tq + tl) * g.nH + nh) * g.nW + nw) * o.K;
      for (int ki = k_lo; ki < k_hi; ki++) {
        const float weight = P.weights[qb + ki];
        if (INT) {
          const int* f = reinterpret_cast<const int*>(P.flows) + (qb + ki) * 3;
          int n0 = t + f[0], n1 = hq + f[1], n2 = wq + f[2];
          if (!((abs(n1) < 1e7) && (abs(n2) < 1e7))) continue;
          n0 = bnd(n0, g.T); n1 = bnd(n1, H); n2 = bnd(n2, W);
          n1 += d; n2 += e;
          if (o.reflect_bounds) { n1 = bnd(n1, H); n2 = bnd(n2, W); }
          if (!(resident(n0, g) && inb(n1, H) && inb(n2, W))) continue;
          const float4 pix = ldg4(vb + (int64_t)(n0 - g.tk0) * vfs + ((int64_t)n1 * W + n2) * g.Fp);
          acc.x += weight * pix.x; acc.y += weight * pix.y; acc.z += weight * pix.z; acc.w += weight * pix.w;
        } else {
          const float* f = reinterpret_cast<const float*>(P.flows) + (qb + ki) * 3;
          float n0 = (float)t + f[0], n1 = (float)hq + f[1], n2 = (float)wq + f[2];
          if (!((fabsf(n1) < 1e7) && (fabsf(n2) < 1e7))) continue;
          n0 = bnd(n0, g.T); n1 = bnd(n1, H); n2 = bnd(n2, W);
          n1 += (float)d; n2 += (float)e;
          if (o.reflect_bounds) { n1 = bnd(n1, H); n2 = bnd(n2, W); }
          if (!(resident(n0, g) && inb(n1, H) && inb(n2, W))) continue;
          const Taps tp = make_taps(n1, n2, H, W);
          const float4 pix = interp4(vb - c4 + (int64_t)((int)n0 - g.tk0) * vfs, tp, W, g.Fp, c4);
          acc.x += weight * pix.x; acc.y += weight * pix.y; acc.z += weight * pix.z; acc.w += weight * pix.w;
        }
      }
    }
  }
  const float cnt = (float)(P.counts_h[h] * P.counts_w[w]);
  float4 o4 = make_float4(acc.x / cnt, acc.y / cnt, acc.z / cnt, acc.w / cnt);
  const int64_t slab = P.stack ? (int64_t)n * o.K + k_lo : n;
  *reinterpret_cast<float4*>(P.out + ((slab * g.Tq + tl) * H + h) * W * g.Fp + (int64_t)w * g.Fp + c4) = o4;
}

__device__ __forceinline__ float dot4(const float4 a, const float4 b) {
  return fmaf(a.w, b.w, fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)));
}
__device__ __forceinline__ void axpy4(float4& y, float a, const float4 x) {
  y.x = fmaf(a, x.x, y.x); y.y = fmaf(a, x.y, y.y); y.z = fmaf(a, x.z, y.z); y.w = fmaf(a, x.w, y.w);
}

// ---- tiled forward for ps == 3, dilation == 1 (the README / 1080p shapes) ------------------------------
// The gather kernel above redoes the whole (query, neighbour) setup -- flow read, reflection, bilinear taps --
// for each of the 9 output pixels a patch covers.  Here a CTA owns a TH x TW tile of output pixels of one
// frame: thread = (query whose patch touches the tile, channel quad) loops over the K neighbours ONCE,
// keeps its 3x3 weighted patch in registers (interior neighbours: one shared 4x4 footprint, 16 loads), and
// the patches are then summed into a shared-memory tile in 9 conflict-free phases (for a fixed patch pixel
// the map query -> output pixel is injective), divided by the overlap counts and stored once.  Queries on
// the tile rim are recomputed by the neighbouring CTA instead of exchanged -- still no atomics.
struct TileCfg {
  int TH, TW, nQh, nQw, tiles_w;
};

template <bool INT>
__global__ void __launch_bounds__(512) wpsum_fwd_tile_kernel(const __grid_constant__ AggParams P,
                                                             const __grid_constant__ TileCfg C) {
  extern __shared__ float4 tile[];   // [TH][TW][nquad]
  const stnls_geom& g = P.g;
  const stnls_agg_opts& o = P.o;
  const int H = g.kH, W = g.kW, Q4 = P.nquad, po = o.patch_offset, s0 = o.stride0;
  const int tile_w = blockIdx.x % C.tiles_w, tile_h = blockIdx.x / C.tiles_w;
  const int tl = blockIdx.y % g.Tq, n = blockIdx.y / g.Tq;
  const int t = tl + g.tq0;
  const int h0 = tile_h * C.TH, w0 = tile_w * C.TW;
  for (int i = threadIdx.x; i < C.TH * C.TW * Q4; i += blockDim.x) tile[i] = f4zero();
  __syncthreads();

  const int c4 = threadIdx.x % Q4;
  const int ql = threadIdx.x / Q4;
  const int qlh = ql / C.nQw, qlw = ql - qlh * C.nQw;
  const int xh = h0 - (2 + po), xw = w0 - (2 + po);            // first reference pixel whose patch reaches the tile
  const int nh = (xh <= 0 ? 0 : (xh + s0 - 1) / s0) + qlh, nw = (xw <= 0 ? 0 : (xw + s0 - 1) / s0) + qlw;
  const int rh = nh * s0, rw = nw * s0;                         // stride0 grids never wrap: nh*s0 <= H-1
  const bool active = qlh < C.nQh && nh < g.nH && nw < g.nW && rh + po <= h0 + C.TH - 1 && rw + po <= w0 + C.TW - 1;

  float4 acc[9];
#pragma unroll
  for (int i = 0; i < 9; i++) acc[i] = f4zero();
  if (active) {
    const int64_t vfs = (int64_t)H * W * g.Fp;
    const float* vb = P.v + (int64_t)n * g.Tk * vfs + c4 * 4;
    const int64_t qb = ((((int64_t)n * g.Tq + tl) * g.nH + nh) * g.nW + nw) * o.K;
    const int64_t rs = (int64_t)W * g.Fp;
    unsigned slow = 0;
    // per-pixel path of one neighbour (reflected / clipped patches): gather_add_kernel.cu:76-123
    auto slow_path = [&](float weight, float n0, float c1, float c2) {
      const float* vf = vb + (int64_t)((int)n0 - g.tk0) * vfs;
#pragma unroll
      for (int pi = 0; pi < 3; pi++) {
        float n1 = c1 + (float)(pi + po);
        if (o.reflect_bounds) n1 = bnd(n1, H);
        if (!inb(n1, H)) continue;
#pragma unroll
        for (int pj = 0; pj < 3; pj++) {   // unrolled: acc[] keeps compile-time indices and stays in registers
          float n2 = c2 + (float)(pj + po);
          if (o.reflect_bounds) n2 = bnd(n2, W);
          if (!inb(n2, W)) continue;
          const Taps tp = make_taps(n1, n2, H, W);
          axpy4(acc[pi * 3 + pj], weight, interp4(vf - c4 * 4, tp, W, g.Fp, c4 * 4));
        }
      }
    };
    // fused normalise step: the query's softmax statistics, the same arithmetic in the same order as
    // softmax_fwd_kernel (normz.cu) so that both routes give the same bits; the tile that holds the query's reference
    // pixel writes the weights for the backward pass
    const bool sm = P.dists != nullptr;
    const float* drow = sm ? P.dists + (qb / o.K) * P.Kin : nullptr;
    float sm_m = 0.f, sm_inv = 0.f;
    if (sm) {
      float m = -INFINITY, ssum = 0.f;
      for (int k = 0; k < o.K; k++) m = fmaxf(m, __fmul_rn(P.sm_scale, __ldg(drow + k)));
      for (int k = 0; k < o.K; k++) ssum = __fadd_rn(ssum, expf(__fsub_rn(__fmul_rn(P.sm_scale, __ldg(drow + k)), m)));
      sm_m = m; sm_inv = 1.f / ssum;
    }
    auto wload = [&](int ki) -> float {
      if (!sm) return __ldg(P.weights + qb + ki);
      return __fmul_rn(expf(__fsub_rn(__fmul_rn(P.sm_scale, __ldg(drow + ki)), sm_m)), sm_inv);
    };
    const bool wowner = sm && c4 == 0 && rh >= h0 && rh < h0 + C.TH && rw >= w0 && rw < w0 + C.TW;
    // the (weight, offset) of neighbour k+1 is fetched while neighbour k is processed
    const int* fi = reinterpret_cast<const int*>(P.flows) + qb * 3;
    float wnext = wload(0);
    int x0 = __ldg(fi), x1 = __ldg(fi + 1), x2 = __ldg(fi + 2);
    for (int ki = 0; ki < o.K; ki++) {
      const float weight = wnext;
      const int y0 = x0, y1 = x1, y2 = x2;
      if (wowner) P.weights_out[qb + ki] = weight;
      if (ki + 1 < o.K) {
        wnext = wload(ki + 1);
        x0 = __ldg(fi + (ki + 1) * 3); x1 = __ldg(fi + (ki + 1) * 3 + 1); x2 = __ldg(fi + (ki + 1) * 3 + 2);
      }
      if (INT) {
        int n0 = t + y0, c1 = rh + y1, c2 = rw + y2;
        if (!((abs(c1) < 1e7) && (abs(c2) < 1e7))) continue;
        n0 = bnd(n0, g.T); c1 = bnd(c1, H); c2 = bnd(c2, W);
        if (!resident(n0, g)) continue;
        const float* vf = vb + (int64_t)(n0 - g.tk0) * vfs;
#pragma unroll
        for (int pi = 0; pi < 3; pi++) {
          int n1 = c1 + pi + po;
          if (o.reflect_bounds) n1 = bnd(n1, H);
          if (!inb(n1, H)) continue;
#pragma unroll
          for (int pj = 0; pj < 3; pj++) {
            int n2 = c2 + pj + po;
            if (o.reflect_bounds) n2 = bnd(n2, W);
            if (!inb(n2, W)) continue;
            axpy4(acc[pi * 3 + pj], weight, ldg4(vf + (int64_t)n1 * rs + (int64_t)n2 * g.Fp));
          }
        }
      } else {
        float n0 = (float)t + __int_as_float(y0), c1 = (float)rh + __int_as_float(y1), c2 = (float)rw + __int_as_float(y2);
        if (!((fabsf(c1) < 1e7) && (fabsf(c2) < 1e7))) continue;
        n0 = bnd(n0, g.T); c1 = bnd(c1, H); c2 = bnd(c2, W);
        if (!resident(n0, g)) continue;
        const float* vf = vb + (int64_t)((int)n0 - g.tk0) * vfs;
        // interior neighbour: all 9 patch pixels unreflected, their taps inside one 4x4 footprint
        const int hh = __float2int_rz(c1 + (float)po), ww = __float2int_rz(c2 + (float)po);
        float gh[3][2], gw[3][2];
        bool ok = true;
#pragma unroll
        for (int p = 0; p < 3; p++) {
          const float n1 = c1 + (float)(p + po), n2 = c2 + (float)(p + po);
          const int a0 = __float2int_rz(n1), a1 = __float2int_rz(n1 + 1);
          const int b0 = __float2int_rz(n2), b1 = __float2int_rz(n2 + 1);
          ok = ok && (n1 >= 0) && (a0 == hh + p) && (a1 == hh + p + 1) && (a1 <= H - 1) && (n1 <= (float)(H - 1));
          ok = ok && (n2 >= 0) && (b0 == ww + p) && (b1 == ww + p + 1) && (b1 <= W - 1) && (n2 <= (float)(W - 1));
          gh[p][0] = fmaxf(0.f, 1 - fabsf(a0 - n1)); gh[p][1] = fmaxf(0.f, 1 - fabsf(a1 - n1));
          gw[p][0] = fmaxf(0.f, 1 - fabsf(b0 - n2)); gw[p][1] = fmaxf(0.f, 1 - fabsf(b1 - n2));
        }
        if (ok) {
          const float* vr = vf + (int64_t)hh * rs + (int64_t)ww * g.Fp;
          float4 A[4], B[4];
#pragma unroll
          for (int j = 0; j < 4; j++) A[j] = ldg4(vr + j * g.Fp);
#pragma unroll
          for (int pi = 0; pi < 3; pi++) {
#pragma unroll
            for (int j = 0; j < 4; j++) B[j] = ldg4(vr + (pi + 1) * rs + j * g.Fp);
#pragma unroll
            for (int pj = 0; pj < 3; pj++) {
              float4 pix = f4zero();   // tap order and weights gh*gw as shared_kernel.cu:738-771
              axpy4(pix, gh[pi][0] * gw[pj][0], A[pj]); axpy4(pix, gh[pi][0] * gw[pj][1], A[pj + 1]);
              axpy4(pix, gh[pi][1] * gw[pj][0], B[pj]); axpy4(pix, gh[pi][1] * gw[pj][1], B[pj + 1]);
              axpy4(acc[pi * 3 + pj], weight, pix);
            }
#pragma unroll
            for (int j = 0; j < 4; j++) A[j] = B[j];
          }
        } else if (o.K <= 32) {
          slow |= 1u << ki;   // deferred: a warp then runs the per-pixel path only as often as its worst lane needs
        } else {
          slow_path(weight, n0, c1, c2);
        }
      }
    }
    while (slow) {
      const int ki = __ffs(slow) - 1;
      slow &= slow - 1;
      const float* f = reinterpret_cast<const float*>(P.flows) + (qb + ki) * 3;
      float n0 = bnd((float)t + __ldg(f), g.T), c1 = bnd((float)rh + __ldg(f + 1), H), c2 = bnd((float)rw + __ldg(f + 2), W);
      slow_path(wload(ki), n0, c1, c2);
    }
  }
  // 9 phases: patch pixel (pi,pj) of every query lands on a different tile pixel
#pragma unroll
  for (int pp = 0; pp < 9; pp++) {
    const int h = rh + pp / 3 + po - h0, w = rw + pp % 3 + po - w0;
    if (active && h >= 0 && h < C.TH && w >= 0 && w < C.TW) {
      float4& d = tile[(h * C.TW + w) * Q4 + c4];
      d.x += acc[pp].x; d.y += acc[pp].y; d.z += acc[pp].z; d.w += acc[pp].w;
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < C.TH * C.TW * Q4; i += blockDim.x) {
    const int q4 = i % Q4, px = i / Q4;
    const int h = h0 + px / C.TW, w = w0 + px % C.TW;
    if (h >= H || w >= W) continue;
    const float cnt = (float)(P.counts_h[h] * P.counts_w[w]);
    const float4 a = tile[i];
    *reinterpret_cast<float4*>(P.out + ((((int64_t)n * g.Tq + tl) * H + h) * W + w) * g.Fp + q4 * 4) =
        make_float4(a.x / cnt, a.y / cnt, a.z / cnt, a.w / cnt);
  }
}


// Sum over the channel-quad lanes of one (query, neighbour): the `nq` lanes (a power of two <= 32, adjacent, all on the
// same control path because every branch above depends on (query, neighbour) only) hold partial dot products; after
// the butterfly lane c4 == 0 issues the ONE store / atomic of the pair -- the reference issues ps*ps atomics per pair
// and channel (gather_add_kernel.cu:315), the round-1 kernel one per channel quad.
__device__ __forceinline__ float quad_sum(float v, int nq, unsigned mask) {
  for (int off = nq >> 1; off > 0; off >>= 1) v += __shfl_xor_sync(mask, v, off);
  return v;
}

// FAST3: ps == 3 && dilation == 1.  The 9 bilinear reads of one neighbour patch then share a 4x4
// pixel footprint (all patch pixels have the same fractional position), so an interior (query, k)
// pair reads 16 value pixels and issues 16 vector REDs instead of 36 + 36.
// KSPLIT: one thread per (query, neighbour, quad) instead of a loop over the K neighbours -- K times
// more threads in flight for a kernel that is latency-bound, not bandwidth-bound
template <bool INT, bool FAST3, bool KSPLIT>
__global__ void __launch_bounds__(128, WB_MINB) wpsum_bwd_kernel(const __grid_constant__ AggParams P) {
  const stnls_geom& g = P.g;
  const stnls_agg_opts& o = P.o;
  const int H = g.kH, W = g.kW;
  int c4, nw, nh, tl, n;
  int64_t wq;
  int k_lo = 0, k_hi = o.K;
  if (P.idx32) {
    unsigned u = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned fr = blockIdx.y + blockIdx.z * gridDim.y;
    if (u >= P.per_frame || fr >= P.frames) return;
    u = fdivmod(u, P.fquad, c4);
    if (KSPLIT) { u = fdivmod(u, P.fK, k_lo); k_hi = k_lo + 1; }
    nh = (int)fdivmod(u, P.fnW, nw);
    n = (int)fdivmod(fr, P.fTq, tl);
    wq = ((int64_t)fr * g.nH + nh) * g.nW + nw;
  } else {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P.nthreads) return;
    wq = divmod(idx, P.nquad, c4);
    if (KSPLIT) { wq = divmod(wq, o.K, k_lo); k_hi = k_lo + 1; }
    int64_t r = divmod(wq, g.nW, nw);
    r = divmod(r, g.nH, nh);
    r = divmod(r, g.Tq, tl);
    n = (int)r;
  }
  c4 *= 4;
  const int t = tl + g.tq0;
  const int rh = (nh * o.stride0) % H, rw = (nw * o.stride0) % W;
  const int64_t vfs = (int64_t)H * W * g.Fp;
  const float* vb = P.v + (int64_t)n * g.Tk * vfs + c4;
  float* gvb = P.out + (int64_t)n * g.Tk * vfs + c4;
  const unsigned qmask = P.qagg ? ((P.nquad >= 32 ? 0xffffffffu : ((1u << P.nquad) - 1u)) << ((threadIdx.x & 31) & ~(P.nquad - 1))) : 0u;
  const float* gof = P.gout + ((P.stack ? (int64_t)n * o.K + k_lo : (int64_t)n) * g.Tq + tl) * vfs + c4;  // stack => KSPLIT

  // generic path for one neighbour: every patch pixel validated / reflected on its own
  // (gather_add_kernel.cu:201-319 int, :369-514 + shared_kernel.cu:773-827 float)
  auto generic = [&](int ki) {
    const int64_t oi = wq * o.K + ki;
    const float weight = P.weights[oi];
    float accw = 0.f, fh = 0.f, fw = 0.f;
    if (INT) {
      const int* f = reinterpret_cast<const int*>(P.flows) + oi * 3;
      int n0 = t + f[0], c1 = rh + f[1], c2 = rw + f[2];
      if (!((abs(c1) < 1e7) && (abs(c2) < 1e7))) return;
      n0 = bnd(n0, g.T); c1 = bnd(c1, H); c2 = bnd(c2, W);
      if (!resident(n0, g)) return;
      const float* vf = vb + (int64_t)(n0 - g.tk0) * vfs;
      float* gvf = gvb + (int64_t)(n0 - g.tk0) * vfs;
      for (int pi = 0; pi < o.ps; pi++) {
        const int d = o.dilation * (pi + o.patch_offset);
        const int r1 = rh + d;
        int n1 = c1 + d;
        if (o.reflect_bounds) n1 = bnd(n1, H);
        if (!(inb(r1, H) && inb(n1, H))) continue;
        for (int pj = 0; pj < o.ps; pj++) {
          const int e = o.dilation * (pj + o.patch_offset);
          const int r2 = rw + e;
          int n2 = c2 + e;
          if (o.reflect_bounds) n2 = bnd(n2, W);
          if (!(inb(r2, W) && inb(n2, W))) continue;
          const float4 gr = ldg4(gof + ((int64_t)r1 * W + r2) * g.Fp);
          const int64_t vo = ((int64_t)n1 * W + n2) * g.Fp;
          const float4 pv = ldg4(vf + vo);
          red_add4(gvf + vo, f4scale(gr, weight));
          accw += gr.x * pv.x + gr.y * pv.y + gr.z * pv.z + gr.w * pv.w;
        }
      }
    } else {
      const float* f = reinterpret_cast<const float*>(P.flows) + oi * 3;
      float n0 = (float)t + f[0], c1 = (float)rh + f[1], c2 = (float)rw + f[2];
      if (!((fabsf(c1) < 1e7) && (fabsf(c2) < 1e7))) return;
      n0 = bnd(n0, g.T);
      const int sH0 = inb(c1, H) ? 1 : -1;
      c1 = bnd(c1, H);
      const int sW0 = inb(c2, W) ? 1 : -1;
      c2 = bnd(c2, W);
      if (!resident(n0, g)) return;
      const int nti = (int)n0;
      const float* vf = vb + (int64_t)(nti - g.tk0) * vfs;
      float* gvf = gvb + (int64_t)(nti - g.tk0) * vfs;
      for (int pi = 0; pi < o.ps; pi++) {
        const int d = o.dilation * (pi + o.patch_offset);
        const int r1 = rh + d;
        float n1 = c1 + (float)d;
        const int signH = inb(n1, H) ? sH0 : -sH0;
        if (o.reflect_bounds) n1 = bnd(n1, H);
        if (!(inb(r1, H) && inb(n1, H))) continue;
        for (int pj = 0; pj < o.ps; pj++) {
          const int e = o.dilation * (pj + o.patch_offset);
          const int r2 = rw + e;
          float n2 = c2 + (float)e;
          const int signW = inb(n2, W) ? sW0 : -sW0;
          if (o.reflect_bounds) n2 = bnd(n2, W);
          if (!(inb(r2, W) && inb(n2, W))) continue;
          const float4 gr = ldg4(gof + ((int64_t)r1 * W + r2) * g.Fp);
          const float4 wg = f4scale(gr, weight);
          Taps tp = make_taps(n1, n2, H, W);
          if (P.stack) {  // NonLocalGather's backward reflects a tap outside the frame (gather_bilin2d.cu:188-189)
#pragma unroll
            for (int x = 0; x < 2; x++) {
              tp.h[x] = bnd(tp.h[x], H); tp.vh[x] = inb(tp.h[x], H);
              tp.w[x] = bnd(tp.w[x], W); tp.vw[x] = inb(tp.w[x], W);
            }
          }
          float aH = 0.f, aW = 0.f;
#pragma unroll
          for (int ix = 0; ix < 2; ix++) {
#pragma unroll
            for (int jx = 0; jx < 2; jx++) {
              if (!tp.vh[ix] || !tp.vw[jx]) continue;
              const int64_t vo = ((int64_t)tp.h[ix] * W + tp.w[jx]) * g.Fp;
              const float4 pv = ldg4(vf + vo);
              const float dot = gr.x * pv.x + gr.y * pv.y + gr.z * pv.z + gr.w * pv.w;
              accw += tp.gh[ix] * tp.gw[jx] * dot;
              aW += (jx == 0) ? -tp.gh[ix] * dot : tp.gh[ix] * dot;
              aH += (ix == 0) ? -tp.gw[jx] * dot : tp.gw[jx] * dot;
              red_add4(gvf + vo, f4scale(wg, tp.gh[ix] * tp.gw[jx]));
            }
          }
          fh += weight * aH * signH;
          fw += weight * aW * signW;
        }
      }
    }
    if (P.qagg) {   // one atomic per (query, neighbour): sum the channel-quad lanes first
      accw = quad_sum(accw, P.nquad, qmask);
      if (!INT) { fh = quad_sum(fh, P.nquad, qmask); fw = quad_sum(fw, P.nquad, qmask); }
      if (c4 != 0) return;
    }
    atomicAdd(P.gweights + oi, accw);
    if (!INT) {
      atomicAdd(P.gflows + oi * 3 + 1, fh);
      atomicAdd(P.gflows + oi * 3 + 2, fw);
    }
  };

  if (!FAST3) {
    for (int ki = k_lo; ki < k_hi; ki++) generic(ki);
    return;
  }

  // FAST3 (float, ps == 3, dilation == 1): interior (query, k) pairs take the footprint path; the
  // others are deferred (bit mask) and handled afterwards so a warp does not run both paths per k
  const int po = o.patch_offset;
  const bool refok = (rh + po >= 0) && (rh + po + 2 <= H - 1) && (rw + po >= 0) && (rw + po + 2 <= W - 1);
  if (!refok || o.K > 32) {
    for (int ki = k_lo; ki < k_hi; ki++) generic(ki);
    return;
  }
  float4 grp[9];  // the query's own 3x3 patch of grad_out, read once for all K neighbours
#pragma unroll
  for (int pi = 0; pi < 3; pi++)
#pragma unroll
    for (int pj = 0; pj < 3; pj++) grp[pi * 3 + pj] = ldg4(gof + ((int64_t)(rh + pi + po) * W + (rw + pj + po)) * g.Fp);
  unsigned slow = 0;
  for (int ki = k_lo; ki < k_hi; ki++) {
    const int64_t oi = wq * o.K + ki;
    const float* f = reinterpret_cast<const float*>(P.flows) + oi * 3;
    float n0 = (float)t + f[0], c1 = (float)rh + f[1], c2 = (float)rw + f[2];
    // centre inside the frame (no reflection) and every tap of the 3x3 patch inside a 4x4 footprint
    n0 = bnd(n0, g.T);
    bool ok = resident(n0, g) && (c1 >= 0) && (c2 >= 0) && (c1 <= (float)(H - 1)) && (c2 <= (float)(W - 1));
    const int h0 = __float2int_rz(c1 + (float)po), w0 = __float2int_rz(c2 + (float)po);
    float gh[3][2], gw[3][2];
#pragma unroll
    for (int p = 0; p < 3; p++) {
      const float n1 = c1 + (float)(p + po), n2 = c2 + (float)(p + po);
      const int a0 = __float2int_rz(n1), a1 = __float2int_rz(n1 + 1);
      const int b0 = __float2int_rz(n2), b1 = __float2int_rz(n2 + 1);
      ok = ok && (n1 >= 0) && (a0 == h0 + p) && (a1 == h0 + p + 1) && (a1 <= H - 1);
      ok = ok && (n2 >= 0) && (b0 == w0 + p) && (b1 == w0 + p + 1) && (b1 <= W - 1);
      gh[p][0] = fmaxf(0.f, 1 - fabsf(a0 - n1)); gh[p][1] = fmaxf(0.f, 1 - fabsf(a1 - n1));
      gw[p][0] = fmaxf(0.f, 1 - fabsf(b0 - n2)); gw[p][1] = fmaxf(0.f, 1 - fabsf(b1 - n2));
    }
    if (!ok) { slow |= 1u << ki; continue; }
    const float weight = P.weights[oi];
    const int nti = (int)n0;
    const int64_t fo = (int64_t)(nti - g.tk0) * vfs + ((int64_t)h0 * W + w0) * g.Fp;
    const float* vr = vb + fo;
    float* gr_ = gvb + fo;
    const int64_t rs = (int64_t)W * g.Fp;
    float4 A[4], B[4], accA[4], accB[4];
#pragma unroll
    for (int j = 0; j < 4; j++) { A[j] = ldg4(vr + j * g.Fp); accA[j] = f4zero(); }
    float accw = 0.f, aH = 0.f, aW = 0.f;
#pragma unroll
    for (int pi = 0; pi < 3; pi++) {
#pragma unroll
      for (int j = 0; j < 4; j++) { B[j] = ldg4(vr + (pi + 1) * rs + j * g.Fp); accB[j] = f4zero(); }
#pragma unroll
      for (int pj = 0; pj < 3; pj++) {
        const float4 gr = grp[pi * 3 + pj];
        const float d00 = dot4(gr, A[pj]), d01 = dot4(gr, A[pj + 1]);
        const float d10 = dot4(gr, B[pj]), d11 = dot4(gr, B[pj + 1]);
        const float w00 = gh[pi][0] * gw[pj][0], w01 = gh[pi][0] * gw[pj][1];
        const float w10 = gh[pi][1] * gw[pj][0], w11 = gh[pi][1] * gw[pj][1];
        accw += w00 * d00 + w01 * d01 + w10 * d10 + w11 * d11;
        aW += gh[pi][0] * (d01 - d00) + gh[pi][1] * (d11 - d10);
        aH += gw[pj][0] * (d10 - d00) + gw[pj][1] * (d11 - d01);
        const float4 wg = f4scale(gr, weight);
        axpy4(accA[pj], w00, wg); axpy4(accA[pj + 1], w01, wg);
        axpy4(accB[pj], w10, wg); axpy4(accB[pj + 1], w11, wg);
      }
#pragma unroll
      for (int j = 0; j < 4; j++) { red_add4(gr_ + pi * rs + j * g.Fp, accA[j]); A[j] = B[j]; accA[j] = accB[j]; }
    }
#pragma unroll
    for (int j = 0; j < 4; j++) red_add4(gr_ + 3 * rs + j * g.Fp, accA[j]);
    if (P.qagg) {
      accw = quad_sum(accw, P.nquad, qmask);
      aH = quad_sum(aH, P.nquad, qmask);
      aW = quad_sum(aW, P.nquad, qmask);
      if (c4 != 0) continue;
    }
    atomicAdd(P.gweights + oi, accw);
    atomicAdd(P.gflows + oi * 3 + 1, weight * aH);
    atomicAdd(P.gflows + oi * 3 + 2, weight * aW);
  }
  while (slow) {
    const int ki = __ffs(slow) - 1;
    slow &= slow - 1;
    generic(ki);
  }
}


int agg_check(const stnls_geom* g, const stnls_agg_opts* o, const char* who) {
  STNLS_REQUIRE(g && o, "%s: null argument", who);
  STNLS_REQUIRE(g->Fp % 4 == 0 && g->Fp >= g->F, "%s: Fp must be a multiple of 4 and >= F", who);
  STNLS_REQUIRE(g->qH == g->kH && g->qW == g->kW, "%s: output size must equal input size", who);
  STNLS_REQUIRE(o->ps > 0 && o->stride0 > 0 && o->K >= 0, "%s: bad ps/stride0/K", who);
  return stnls_check_geom(g, -1, who);
}

}  // namespace

// dists != NULL: fused normalise step -- `weights` is then an OUTPUT (softmax(scale * dists[:, :K])); *launches = kernels run
static int agg_forward(float* out_cl, const float* v_cl, const float* weights, const void* flows,
                       const int* counts_h, const int* counts_w, const stnls_geom* g,
                       const stnls_agg_opts* o, stnls_stream_t stream, int stack, const float* dists = nullptr,
                       int Kin = 0, float scale = 0.f, int* launches = nullptr) {
  if (int rc = agg_check(g, o, "wpsum_forward")) return rc;
  STNLS_REQUIRE(out_cl && v_cl && weights && flows && counts_h && counts_w, "wpsum_forward: null argument");
  if (launches) *launches = 0;
  AggParams P = {};
  P.out = out_cl; P.v = v_cl; P.weights = weights; P.flows = flows; P.counts_h = counts_h; P.counts_w = counts_w;
  P.g = *g; P.o = *o; P.nquad = g->Fp / 4;
  P.stack = stack;
  P.nthreads = (int64_t)g->B * g->HD * g->Tq * g->kH * g->kW * P.nquad * (stack ? o->K : 1);
  if (P.nthreads == 0) return STNLS_OK;
  STNLS_REQUIRE(P.nthreads / 128 < 2147483647LL, "wpsum_forward: too many threads for one launch");
  unsigned blocks = (unsigned)((P.nthreads + 127) / 128);
  cudaStream_t st = (cudaStream_t)stream;
  if (launches) *launches = 1;
  // tile kernel: the largest tile with (queries touching the tile) x quads <= 512 threads
  static const int dims[5][2] = {{16, 32}, {16, 16}, {12, 16}, {8, 16}, {8, 8}};
  TileCfg C = {};
  for (int i = 0; i < 5; i++) {
    C.TH = dims[i][0]; C.TW = dims[i][1];
    C.nQh = (C.TH + 1) / o->stride0 + 1; C.nQw = (C.TW + 1) / o->stride0 + 1;
    if (C.nQh * C.nQw * P.nquad <= 512) break;
  }
  const int threads = (C.nQh * C.nQw * P.nquad + 31) / 32 * 32;
  const int64_t ny = (int64_t)g->B * g->HD * g->Tq;
  const bool tile_ok = !stack && o->ps == 3 && o->dilation == 1 && P.nquad <= 8 && !getenv("STNLS_WPF_NO_TILE") &&
                       threads <= 512 && ny <= 65535;
  if (dists) {
    STNLS_REQUIRE(o->K > 0 && Kin >= o->K, "wpsum_forward_sm: rows of dists must hold at least K entries (K=%d, Kin=%d)", o->K, Kin);
    if (tile_ok && !getenv("STNLS_WPF_NO_FUSED_SM")) {
      P.dists = dists; P.weights_out = const_cast<float*>(weights); P.sm_scale = scale; P.Kin = Kin;
    } else {   // the gather kernel would redo the row statistics once per covered pixel: normalise first, two launches
      if (int rc = stnls_softmax_forward(const_cast<float*>(weights), dists, ny * g->nH * g->nW, o->K, Kin, scale, stream)) return rc;
      if (launches) *launches = 2;
    }
  }
  if (tile_ok) {
    C.tiles_w = (g->kW + C.TW - 1) / C.TW;
    const int tiles_h = (g->kH + C.TH - 1) / C.TH;
    dim3 grid((unsigned)(C.tiles_w * tiles_h), (unsigned)ny);
    const size_t smem = (size_t)C.TH * C.TW * P.nquad * 16;
    if (o->itype_int) wpsum_fwd_tile_kernel<true><<<grid, threads, smem, st>>>(P, C);
    else wpsum_fwd_tile_kernel<false><<<grid, threads, smem, st>>>(P, C);
    STNLS_CHECK_LAUNCH("wpsum_forward(tile)");
    return STNLS_OK;
  }
  if (o->itype_int) wpsum_fwd_kernel<true><<<blocks, 128, 0, st>>>(P);
  else wpsum_fwd_kernel<false><<<blocks, 128, 0, st>>>(P);
  STNLS_CHECK_LAUNCH("wpsum_forward");
  return STNLS_OK;
}

extern "C" int stnls_wpsum_forward(float* out_cl, const float* v_cl, const float* weights, const void* flows,
                                   const int* counts_h, const int* counts_w, const stnls_geom* g,
                                   const stnls_agg_opts* o, stnls_stream_t stream) {
  return agg_forward(out_cl, v_cl, weights, flows, counts_h, counts_w, g, o, stream, 0);
}
extern "C" int stnls_wpsum_forward_sm(float* out_cl, float* weights, const float* v_cl, const float* dists, int Kin,
                                      float scale, const void* flows, const int* counts_h, const int* counts_w,
                                      const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream, int* launches) {
  STNLS_REQUIRE(dists, "wpsum_forward_sm: null argument");
  return agg_forward(out_cl, v_cl, weights, flows, counts_h, counts_w, g, o, stream, 0, dists, Kin, scale, launches);
}
extern "C" int stnls_gather_forward(float* stack_cl, const float* v_cl, const float* weights, const void* flows,
                                    const int* counts_h, const int* counts_w, const stnls_geom* g,
                                    const stnls_agg_opts* o, stnls_stream_t stream) {
  return agg_forward(stack_cl, v_cl, weights, flows, counts_h, counts_w, g, o, stream, 1);
}

static int agg_backward(float* gv_cl, float* grad_weights, float* grad_flows, const float* gout_cl,
                        const float* v_cl, const float* weights, const void* flows,
                        const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream, int stack) {
  if (int rc = agg_check(g, o, "wpsum_backward")) return rc;
  STNLS_REQUIRE(gv_cl && grad_weights && gout_cl && v_cl && weights && flows, "wpsum_backward: null argument");
  STNLS_REQUIRE(o->itype_int || grad_flows, "wpsum_backward: grad_flows required for itype float");
  AggParams P = {};
  P.out = gv_cl; P.gweights = grad_weights; P.gflows = grad_flows; P.gout = gout_cl; P.v = v_cl;
  P.weights = weights; P.flows = flows; P.g = *g; P.o = *o; P.nquad = g->Fp / 4;
  P.nthreads = (int64_t)g->B * g->HD * g->Tq * g->nH * g->nW * P.nquad;
  P.qagg = (P.nquad > 1 && P.nquad <= 32 && (P.nquad & (P.nquad - 1)) == 0 && getenv("STNLS_QAGG")) ? 1 : 0;   // opt-in: measured SLOWER
  // (B200, r02_l vs r02_g: wpsum_backward c4s 74.9 -> 83.8 ms, search_backward 91.9 -> 98.5 ms; C2 0.294 -> 0.309 ms): the
  // per-pair atomics were never the limiter, the butterflies and the convergence they force are pure overhead
  if (P.nthreads == 0 || o->K == 0) return STNLS_OK;
  STNLS_REQUIRE(P.nthreads / 128 < 2147483647LL, "wpsum_backward: too many threads for one launch");
  unsigned blocks = (unsigned)((P.nthreads + 127) / 128);
  cudaStream_t st = (cudaStream_t)stream;
  const bool fast3 = (o->ps == 3 && o->dilation == 1 && !getenv("STNLS_NO_FAST3"));
  P.stack = stack;
  const bool ksplit = stack || (o->ps >= 2 && getenv("STNLS_WPB_NO_KSPLIT") == nullptr);  // default on: 0.47 -> 0.29 ms on C2
  if (ksplit) {
    P.nthreads *= o->K;
    STNLS_REQUIRE(P.nthreads / 128 < 2147483647LL, "wpsum_backward: too many threads for one launch");
    blocks = (unsigned)((P.nthreads + 127) / 128);
  }
  dim3 grid(blocks);
  {
    const int64_t per_frame = (int64_t)g->nH * g->nW * P.nquad * (ksplit ? o->K : 1), frames = (int64_t)g->B * g->HD * g->Tq;
    P.idx32 = (!getenv("STNLS_NO_FRAMEGRID") && frame_grid(per_frame, frames, 128, grid)) ? 1 : 0;
    if (!P.idx32) grid = dim3(blocks);
    P.per_frame = (unsigned)per_frame; P.frames = (unsigned)frames;
    P.fquad = make_fastdiv(P.nquad); P.fK = make_fastdiv(o->K); P.fnW = make_fastdiv(g->nW); P.fTq = make_fastdiv(g->Tq);
  }
  if (o->itype_int) {
    if (ksplit) wpsum_bwd_kernel<true, false, true><<<grid, 128, 0, st>>>(P);
    else wpsum_bwd_kernel<true, false, false><<<grid, 128, 0, st>>>(P);
  } else if (fast3) {
    if (ksplit) wpsum_bwd_kernel<false, true, true><<<grid, 128, 0, st>>>(P);
    else wpsum_bwd_kernel<false, true, false><<<grid, 128, 0, st>>>(P);
  } else {
    if (ksplit) wpsum_bwd_kernel<false, false, true><<<grid, 128, 0, st>>>(P);
    else wpsum_bwd_kernel<false, false, false><<<grid, 128, 0, st>>>(P);
  }
  STNLS_CHECK_LAUNCH("wpsum_backward");
  return STNLS_OK;
}

extern "C" int stnls_wpsum_backward(float* gv_cl, float* grad_weights, float* grad_flows, const float* gout_cl,
                                    const float* v_cl, const float* weights, const void* flows,
                                    const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream) {
  return agg_backward(gv_cl, grad_weights, grad_flows, gout_cl, v_cl, weights, flows, g, o, stream, 0);
}
extern "C" int stnls_gather_backward(float* gv_cl, float* grad_weights, float* grad_flows, const float* gstack_cl,
                                     const float* v_cl, const float* weights, const void* flows,
                                     const stnls_geom* g, const stnls_agg_opts* o, stnls_stream_t stream) {
  return agg_backward(gv_cl, grad_weights, grad_flows, gstack_cl, v_cl, weights, flows, g, o, stream, 1);
}
