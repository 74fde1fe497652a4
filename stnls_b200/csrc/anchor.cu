// Per-frame anchoring of the un-reduced candidate list: self_action = "anchor_each" /
// "anchor_and_remove_ref_frame" of NonLocalSearch (lib/stnls/search/impl/non_local_search.py:96-110).
// Behaviour follows anchor_self_time_kernel (lib/csrc/nn/anchor_self_kernel.cu:187-307): for every
// (query, frame group st) the candidate whose index equals the flow-predicted location `loc` is swapped into
// slot 0 of the group, and slot 0's index is overwritten with `loc`:
//   loc_t = t_next - t_ref,  loc_h = reflect(h_ref + flow_h) - h_ref,  loc_w likewise   (st >= st_offset)
//   loc   = (t_next - t_ref, 0, 0)                                                        (st <  st_offset)
// int indices: first exact match, else slot 0.  float indices: first candidate with |delta|_1 < 1e-8, else the
// first candidate with the smallest |delta|_1 (< 1e4), else slot 0.
// One warp per (query, group): lanes stride over the S candidates, a shuffle reduction picks the winner.
// HBM-bound: 16 S bytes read + 32 bytes written per row.
#include <type_traits>
#include "common.cuh"

namespace {

struct AnchorTimeParams {
  float* dists;       // (N, Q, G, S)
  void* inds;         // (N, Q, G, S, 3)
  const void* flows;  // (B, HD_f, T, nflow, 2, nH, nW)
  int64_t rows;       // N * Q * G
  int HD, HD_f, T, G, S, nflow, st_offset, wt, stride0, qH, qW, kH, kW, nH, nW;
};

template <typename IT>
__global__ void __launch_bounds__(256) anchor_time_kernel(AnchorTimeParams P) {
  const int lane = threadIdx.x & 31;
  const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= P.rows) return;
  const int st = (int)(row % P.G);
  const int64_t nq = row / P.G;
  const int Q = P.T * P.nH * P.nW;
  const int qi = (int)(nq % Q);
  const int n = (int)(nq / Q);
  const int bi = n / P.HD, hi_f = (n % P.HD) % P.HD_f;

  // reference pixel of the query (shared_kernel.cu:76-98)
  const int nHW = P.nH * P.nW;
  const int ti = qi / nHW;
  const int rem = qi - ti * nHW;
  const int nh = rem / P.nW;
  const int h_ref = (nh * P.stride0) % P.qH;
  const int w_ref = (((rem - nh * P.nW) % P.nW) * P.stride0) % P.qW;
  const int n_hi = h_ref / P.stride0, n_wi = w_ref / P.stride0;

  const int t_max = time_tmax(ti, P.T, P.wt);
  int t_next = ti + st;
  t_next = (t_next > t_max) ? t_max - st : t_next;

  IT loc[3];
  loc[0] = (IT)(t_next - ti);
  if (st >= P.st_offset) {
    const IT* f = (const IT*)P.flows +
                  ((((int64_t)bi * P.HD_f + hi_f) * P.T + ti) * P.nflow + (st - P.st_offset)) * 2 * (int64_t)nHW;
    const int64_t o = (int64_t)n_hi * P.nW + n_wi;
    IT lh = (IT)h_ref + f[nHW + o];
    IT lw = (IT)w_ref + f[o];
    loc[1] = bnd(lh, P.kH) - (IT)h_ref;
    loc[2] = bnd(lw, P.kW) - (IT)w_ref;
  } else {
    loc[1] = 0;
    loc[2] = 0;
  }

  IT* ind = (IT*)P.inds + row * P.S * 3;
  float* dist = P.dists + row * P.S;
  // key = (delta, index), smallest wins; an exact match has delta -1
  float best = 10000.f;
  int best_i = 0x7fffffff;
  for (int s = lane; s < P.S; s += 32) {
    float delta;
    if constexpr (std::is_integral<IT>::value) {
      const bool eq = ind[s * 3] == loc[0] && ind[s * 3 + 1] == loc[1] && ind[s * 3 + 2] == loc[2];
      delta = eq ? -1.f : 20000.f;
    } else {
      delta = 0.f;
#pragma unroll
      for (int ix = 0; ix < 3; ix++) delta += fabsf((float)ind[s * 3 + ix] - (float)loc[ix]);
      if (delta < 1e-8f) delta = -1.f;
    }
    if (delta < best) { best = delta; best_i = s; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ob = __shfl_xor_sync(FULL_MASK, best, off);
    const int oi = __shfl_xor_sync(FULL_MASK, best_i, off);
    if (ob < best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
  }
  const int m = (best_i == 0x7fffffff) ? 0 : best_i;
  if (lane == 0) {
    const float d0 = dist[0];
    dist[0] = dist[m];
    dist[m] = d0;
    const IT t0 = ind[0], t1 = ind[1], t2 = ind[2];
    ind[0] = loc[0]; ind[1] = loc[1]; ind[2] = loc[2];
    ind[m * 3] = t0; ind[m * 3 + 1] = t1; ind[m * 3 + 2] = t2;
  }
}

}  // namespace

extern "C" int stnls_anchor_self_time(float* dists, void* inds, const void* flows, int B, int HD, int HD_f,
                                      int T, int G, int S, int nflow, int wt, int stride0, int qH, int qW,
                                      int kH, int kW, int itype_int, stnls_stream_t stream) {
  STNLS_REQUIRE(dists && inds && flows, "anchor_self_time: null argument");
  STNLS_REQUIRE(B > 0 && HD > 0 && HD_f > 0 && T > 0 && G > 0 && S > 0 && nflow > 0 && stride0 > 0,
                "anchor_self_time: sizes must be positive");
  STNLS_REQUIRE(G == 2 * wt + 1 || G == 2 * wt, "anchor_self_time: %d groups for wt=%d (need 2wt+1 or 2wt)", G, wt);
  STNLS_REQUIRE(G - nflow <= 1 && G - nflow >= -1, "anchor_self_time: %d flows for %d groups", nflow, G);
  AnchorTimeParams P;
  P.dists = dists; P.inds = inds; P.flows = flows;
  P.HD = HD; P.HD_f = HD_f; P.T = T; P.G = G; P.S = S; P.nflow = nflow; P.st_offset = G - nflow;
  P.wt = wt; P.stride0 = stride0; P.qH = qH; P.qW = qW; P.kH = kH; P.kW = kW;
  P.nH = (kH - 1) / stride0 + 1; P.nW = (kW - 1) / stride0 + 1;
  P.rows = (int64_t)B * HD * T * P.nH * P.nW * G;
  const int64_t blocks = (P.rows * 32 + 255) / 256;
  STNLS_REQUIRE(blocks < 2147483647LL, "anchor_self_time: too many rows for one launch");
  cudaStream_t st = (cudaStream_t)stream;
  if (itype_int) anchor_time_kernel<int><<<(unsigned)blocks, 256, 0, st>>>(P);
  else anchor_time_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(P);
  STNLS_CHECK_LAUNCH("anchor_self_time");
  return STNLS_OK;
}
