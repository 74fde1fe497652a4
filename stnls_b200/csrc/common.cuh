// Shared device helpers for the sm_100a kernels.  Index math follows the behavioural spec in
// SURVEY.md appendix A (reference: lib/csrc/shared_kernel.cu:49-60,76-98,143-161,248-312).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/stnls_b200.h"

#define FULL_MASK 0xffffffffu

// -------- errors ---------------------------------------------------------------------------
void stnls_set_error(const char* fmt, ...);
// host: frame-shard fields of a geometry are consistent (0 <= tq0, tq0+Tq <= T, 0 <= tk0, tk0+Tk <= T); `wt` >= 0
// additionally requires the resident key frames to cover the temporal windows of the owned query frames
int stnls_check_geom(const stnls_geom* g, int wt, const char* who);
#define STNLS_REQUIRE(cond, ...)            \
  do {                                      \
    if (!(cond)) {                          \
      stnls_set_error(__VA_ARGS__);         \
      return STNLS_EINVAL;                  \
    }                                       \
  } while (0)
#define STNLS_CHECK_LAUNCH(name)                                             \
  do {                                                                       \
    cudaError_t e_ = cudaGetLastError();                                     \
    if (e_ != cudaSuccess) {                                                 \
      stnls_set_error("%s: %s", name, cudaGetErrorString(e_));               \
      return STNLS_ECUDA;                                                    \
    }                                                                        \
  } while (0)

// -------- index decode: x = q * d + rem, x >= 0.  The 64-bit division is a ~100-instruction software routine;
// every launch this library makes has fewer than 2^32 work items, so the 32-bit path is the one that runs.
__device__ __forceinline__ int64_t divmod(int64_t x, int d, int& rem) {
  if (x <= 0xffffffffLL) {
    const unsigned q = (unsigned)x / (unsigned)d;
    rem = (int)((unsigned)x - q * (unsigned)d);
    return (int64_t)q;
  }
  const int64_t q = x / d;
  rem = (int)(x - q * d);
  return q;
}

// Division by a launch-constant divisor as multiply-high + shift (3 instructions instead of the ~20 of a 32-bit
// hardware-less division): valid for dividends below 2^31.  mul == 0 encodes d == 1.
struct FastDiv {
  unsigned d, mul, shr;
};
inline FastDiv make_fastdiv(int d_) {
  FastDiv f;
  f.d = (unsigned)d_; f.mul = 0; f.shr = 0;
  if (d_ > 1) {
    unsigned l = 0;
    while ((1ull << l) < (unsigned long long)d_) l++;          // ceil(log2 d)
    const unsigned p = 31 + l;
    f.mul = (unsigned)((((unsigned long long)1 << p) + f.d - 1) / f.d);
    f.shr = p - 32;
  }
  return f;
}
// Launch shape of the per-(query, ...) kernels: grid.x covers the threads of ONE frame (n, t) -- fewer than 2^31, so the
// index inside the frame is decoded with FastDivs -- and grid.y/z enumerate the frames.  false: a frame alone has 2^31
// threads or more (the caller falls back to the flat 64-bit index).
inline bool frame_grid(int64_t per_frame, int64_t frames, int tpb, dim3& grid) {
  if (per_frame <= 0 || frames <= 0 || per_frame >= (1LL << 31) || frames >= (1LL << 30)) return false;
  const unsigned gy = (unsigned)(frames < 32768 ? frames : 32768);
  const unsigned gz = (unsigned)((frames + gy - 1) / gy);
  if (gz > 65535) return false;
  grid = dim3((unsigned)((per_frame + tpb - 1) / tpb), gy, gz);
  return true;
}
#ifdef __CUDACC__
__device__ __forceinline__ unsigned fdivmod(unsigned x, const FastDiv& f, int& rem) {
  const unsigned q = f.mul ? (__umulhi(x, f.mul) >> f.shr) : x;
  rem = (int)(x - q * f.d);
  return q;
}
#endif

// -------- reflect-once bounds (shared_kernel.cu:49-60) ---------------------------------------
__device__ __forceinline__ int bnd(int v, int lim) {
  if (v < 0) return -v;
  if (v > lim - 1) return 2 * (lim - 1) - v;
  return v;
}
__device__ __forceinline__ float bnd(float v, int lim) {
  if (v < 0) return -v;
  if (v > (float)(lim - 1)) return (float)(2 * (lim - 1)) - v;
  return v;
}
__device__ __forceinline__ bool inb(int v, int lim) { return (v >= 0) && (v <= lim - 1); }
__device__ __forceinline__ bool inb(float v, int lim) { return (v >= 0) && (v <= (float)(lim - 1)); }
// frame t (global index) is one of the key/value frames this rank holds, [tk0, tk0+Tk).  Every kernel tests this
// before it forms a frame pointer: a frame inside the clip but outside the resident halo counts as invalid
// (skipped) instead of being read or written out of bounds.  Unsharded: tk0 = 0, Tk = T, i.e. inb(t, T).
__device__ __forceinline__ bool resident(int t, const stnls_geom& g) { return (unsigned)(t - g.tk0) < (unsigned)g.Tk; }
__device__ __forceinline__ bool resident(float t, const stnls_geom& g) {
  return (t >= (float)g.tk0) && (t <= (float)(g.tk0 + g.Tk - 1));
}

// -------- temporal window (shared_kernel.cu:307-312) -----------------------------------------
__device__ __forceinline__ int time_tmax(int ti, int T, int wt) {
  int t_shift = min(0, ti - wt) + max(0, ti + wt - (T - 1));
  return min(T - 1, ti + wt - t_shift);
}

// -------- search-window placement (shared_kernel.cu:248-287) ---------------------------------
__device__ __forceinline__ int search_off(int a, int s1, int wsHalf, int ws, int lim, bool full_ws) {
  if (!full_ws) return wsHalf;
  int off = wsHalf;
  if ((a - s1 * wsHalf) < 0) off = (int)floor(a / (1. * s1));
  int aMax = a + s1 * ((ws - 1) - off);
  if (aMax > (lim - 1)) off = (int)ceil((a - (lim - 1)) / (1. * s1) + (ws - 1));
  return off;
}
__device__ __forceinline__ float search_off(float a, float s1, float wsHalf, int ws, int lim, bool full_ws) {
  if (!full_ws) return wsHalf;
  float off = wsHalf;
  if (fmaf(-s1, wsHalf, a) < 0) off = (float)floor(a / (1. * s1));
  float aMax = fmaf(s1, ((float)(ws - 1) - off), a);
  if (aMax > (float)(lim - 1)) off = (float)ceil((a - (float)(lim - 1)) / (1. * s1) + (ws - 1));
  return roundf(off);
}

// -------- bilinear taps (shared_kernel.cu:738-771): tap = trunc(x + {0,1}), w = max(0,1-|tap-x|),
//          taps outside the frame are dropped (not clamped, not reflected) ------------------------
struct Taps {
  int h[2], w[2];
  float gh[2], gw[2];
  bool vh[2], vw[2];
};
__device__ __forceinline__ Taps make_taps(float hi, float wi, int H, int W) {
  Taps t;
#pragma unroll
  for (int x = 0; x < 2; x++) {
    t.h[x] = __float2int_rz(hi + x);
    t.gh[x] = fmaxf(0.f, 1 - fabsf(t.h[x] - hi));
    t.vh[x] = inb(t.h[x], H);
    t.w[x] = __float2int_rz(wi + x);
    t.gw[x] = fmaxf(0.f, 1 - fabsf(t.w[x] - wi));
    t.vw[x] = inb(t.w[x], W);
  }
  return t;
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

// vector reduction into global memory: one 16-byte RED instead of four scalar atomics (sm_90+)
__device__ __forceinline__ void red_add4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w));   // no "memory" clobber: the accumulators never alias what a kernel reads, so loads may be
                            // scheduled across the REDs (they are fire-and-forget)
}

__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ float4 f4scale(float4 a, float s) {
  return make_float4(a.x * s, a.y * s, a.z * s, a.w * s);
}
__device__ __forceinline__ float f4sum(float4 a) { return (a.x + a.y) + (a.z + a.w); }

// interpolate one channel quad at (taps) from a cl frame: frame + (h*W + w)*Fp + c4
__device__ __forceinline__ float4 interp4(const float* frame, const Taps& t, int W, int Fp, int c4) {
  float4 pix = f4zero();
#pragma unroll
  for (int ix = 0; ix < 2; ix++) {
#pragma unroll
    for (int jx = 0; jx < 2; jx++) {
      if (!t.vh[ix] || !t.vw[jx]) continue;
      float w = t.gh[ix] * t.gw[jx];
      float4 v = ldg4(frame + ((int64_t)t.h[ix] * W + t.w[jx]) * Fp + c4);
      pix.x = fmaf(w, v.x, pix.x);
      pix.y = fmaf(w, v.y, pix.y);
      pix.z = fmaf(w, v.z, pix.z);
      pix.w = fmaf(w, v.w, pix.w);
    }
  }
  return pix;
}
