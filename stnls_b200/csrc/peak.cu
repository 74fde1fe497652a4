// fp32 FMA-pipe microbenchmark: the roofline denominator of the search forward, MEASURED in the same run as the
// kernels it is compared with.  Every thread runs 16 independent FFMA chains of `iters` steps, so
// the FMA pipe is the only limiter; flops per launch = 2 * 16 * iters * 256 * blocks.
#include "common.cuh"

namespace {
__global__ void __launch_bounds__(256) ffma_peak_kernel(float* sink, int iters) {
  float a[16];
  const float x = 1.0f + 1e-7f * (float)threadIdx.x, y = 1e-9f * (float)blockIdx.x;
#pragma unroll
  for (int i = 0; i < 16; i++) a[i] = (float)i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = fmaf(a[i], x, y);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; i++) s += a[i];
  if (s == 123.456f) sink[0] = s;   // never true: keeps the chains alive
}
}  // namespace

extern "C" int stnls_ffma_peak(float* sink, int blocks, int iters, stnls_stream_t stream) {
  STNLS_REQUIRE(sink && blocks > 0 && iters > 0, "ffma_peak: bad argument");
  ffma_peak_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(sink, iters);
  STNLS_CHECK_LAUNCH("ffma_peak");
  return STNLS_OK;
}
