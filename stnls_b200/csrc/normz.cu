// Fused normalise step between the search and the aggregation:
//   weights = softmax(scale * sign * dists[..., :k_agg], -1),  sign = -1 for l2 distances
// Replaces the torch op chain of lib/stnls/normz/softmax.py:21-42 (slice -> neg -> mul -> Softmax) and of the
// README (`softmax(10*dists,-1)`, README.md:51) by one kernel per direction.  One thread per query row, the
// K <= 64 values in registers; invalid candidates (-inf after the sign flip) get weight 0, a row with no
// valid candidate gives NaN exactly as th.softmax does.  HBM-bound: (Kin + K) * 4 bytes per row.
#include "common.cuh"

namespace {

constexpr int MAXK = 64;

template <int KT>
__global__ void __launch_bounds__(256) softmax_fwd_kernel(float* __restrict__ out, const float* __restrict__ dists,
                                                          int64_t rows, int K, int Kin, float scale) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const int Kc = KT > 0 ? KT : K;
  const float* x = dists + r * Kin;
  float v[KT > 0 ? KT : MAXK];
  float m = -INFINITY;
#pragma unroll
  for (int k = 0; k < (KT > 0 ? KT : MAXK); k++) {
    if (k < Kc) { v[k] = scale * __ldg(x + k); m = fmaxf(m, v[k]); }
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < (KT > 0 ? KT : MAXK); k++) {
    if (k < Kc) { v[k] = expf(v[k] - m); s += v[k]; }
  }
  const float inv = 1.f / s;
  float* o = out + r * Kc;
#pragma unroll
  for (int k = 0; k < (KT > 0 ? KT : MAXK); k++) {
    if (k < Kc) o[k] = v[k] * inv;
  }
}

// grad_dists[k] = scale * w[k] * (g[k] - sum_j g[j] w[j]) for k < K, 0 for K <= k < Kin
template <int KT>
__global__ void __launch_bounds__(256) softmax_bwd_kernel(float* __restrict__ gdists, const float* __restrict__ gw,
                                                          const float* __restrict__ w, int64_t rows, int K, int Kin,
                                                          float scale) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const int Kc = KT > 0 ? KT : K;
  const float* wr = w + r * Kc;
  const float* gr = gw + r * Kc;
  float wv[KT > 0 ? KT : MAXK], gv[KT > 0 ? KT : MAXK];
  float dot = 0.f;
#pragma unroll
  for (int k = 0; k < (KT > 0 ? KT : MAXK); k++) {
    if (k < Kc) { wv[k] = __ldg(wr + k); gv[k] = __ldg(gr + k); dot = fmaf(wv[k], gv[k], dot); }
  }
  float* o = gdists + r * Kin;
#pragma unroll
  for (int k = 0; k < (KT > 0 ? KT : MAXK); k++) {
    if (k < Kc) o[k] = scale * wv[k] * (gv[k] - dot);
  }
  for (int k = Kc; k < Kin; k++) o[k] = 0.f;
}

}  // namespace

extern "C" int stnls_softmax_forward(float* weights, const float* dists, int64_t rows, int K, int Kin,
                                     float scale, stnls_stream_t stream) {
  STNLS_REQUIRE(weights && dists, "softmax_forward: null argument");
  STNLS_REQUIRE(K > 0 && K <= MAXK && Kin >= K, "softmax_forward: need 0 < K <= %d and Kin >= K (K=%d, Kin=%d)", MAXK, K, Kin);
  if (rows == 0) return STNLS_OK;
  STNLS_REQUIRE(rows / 256 < 2147483647LL, "softmax_forward: too many rows for one launch");
  const unsigned blocks = (unsigned)((rows + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (K == 10) softmax_fwd_kernel<10><<<blocks, 256, 0, st>>>(weights, dists, rows, K, Kin, scale);
  else if (K == 9) softmax_fwd_kernel<9><<<blocks, 256, 0, st>>>(weights, dists, rows, K, Kin, scale);
  else if (K == 20) softmax_fwd_kernel<20><<<blocks, 256, 0, st>>>(weights, dists, rows, K, Kin, scale);
  else softmax_fwd_kernel<0><<<blocks, 256, 0, st>>>(weights, dists, rows, K, Kin, scale);
  STNLS_CHECK_LAUNCH("softmax_forward");
  return STNLS_OK;
}

extern "C" int stnls_softmax_backward(float* grad_dists, const float* grad_weights, const float* weights,
                                      int64_t rows, int K, int Kin, float scale, stnls_stream_t stream) {
  STNLS_REQUIRE(grad_dists && grad_weights && weights, "softmax_backward: null argument");
  STNLS_REQUIRE(K > 0 && K <= MAXK && Kin >= K, "softmax_backward: need 0 < K <= %d and Kin >= K (K=%d, Kin=%d)", MAXK, K, Kin);
  if (rows == 0) return STNLS_OK;
  STNLS_REQUIRE(rows / 256 < 2147483647LL, "softmax_backward: too many rows for one launch");
  const unsigned blocks = (unsigned)((rows + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (K == 10) softmax_bwd_kernel<10><<<blocks, 256, 0, st>>>(grad_dists, grad_weights, weights, rows, K, Kin, scale);
  else if (K == 9) softmax_bwd_kernel<9><<<blocks, 256, 0, st>>>(grad_dists, grad_weights, weights, rows, K, Kin, scale);
  else if (K == 20) softmax_bwd_kernel<20><<<blocks, 256, 0, st>>>(grad_dists, grad_weights, weights, rows, K, Kin, scale);
  else softmax_bwd_kernel<0><<<blocks, 256, 0, st>>>(grad_dists, grad_weights, weights, rows, K, Kin, scale);
  STNLS_CHECK_LAUNCH("softmax_backward");
  return STNLS_OK;
}
