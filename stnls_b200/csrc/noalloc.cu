// Allocation-free forward search on PLANAR videos (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// The ordinary route is three calls with two caller-allocated scratch buffers:
//     stnls_pack_cl(q_cl <- q),  stnls_pack_cl(k_cl <- k),  stnls_search_forward(q_cl, k_cl, ...)
// Here the two channels-last copies live in ONE file-scope device array of STNLS_NOALLOC_MAX_BYTES that every GPU
// which loads the library holds until exit.  The same pack kernel and the same search kernels run on it, so the
// results are the ones of the ordinary route bit for bit.
//   * Nothing has to start at zero: the pack step writes every float the search reads (pad channels as zeros), so the
//     contents a previous call left behind are never observed.
//   * Every call on a GPU shares the array: calls that overlap on that GPU (two streams, two host threads) RACE.
//     Serialise them (one stream, or an event between them).
//   * No cudaMalloc / cudaFree / allocating library routine is called; the device address of the array is looked up
//     once per device with cudaGetSymbolAddress (in host code the array's NAME is only a host-side stand-in).
#include <mutex>

#include "../../include/stnls_b200.h"
#include "common.cuh"

namespace {

constexpr size_t kScratchFloats = (size_t)STNLS_NOALLOC_MAX_BYTES / sizeof(float);
static __device__ __align__(256) float g_noalloc_scratch[kScratchFloats];

// never launched: names the array in device code so that it is kept in the module whatever the host side does
__global__ void noalloc_keep_kernel(float* out) {
  if (out) out[0] = g_noalloc_scratch[0];
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
float* g_scratch_ptr[kMaxDevices];   // per device, the array's DEVICE address (nullptr: not looked up yet)

// the device address of the scratch array on the CURRENT device
int scratch_for_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    stnls_set_error("search_forward_planar_noalloc: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return STNLS_ECUDA;
  }
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_ptr[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      stnls_set_error("search_forward_planar_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return STNLS_ECUDA;
    }
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return STNLS_OK;
}

inline size_t round256(size_t bytes) { return (bytes + 255) & ~(size_t)255; }

}  // namespace

extern "C" int stnls_search_forward_planar_noalloc(const float* q, const float* k, const void* flows, float* dists,
                                                   void* inds, const stnls_geom* g, const stnls_search_opts* o,
                                                   stnls_stream_t stream) {
  STNLS_REQUIRE(q && k && dists && inds && g && o, "search_forward_planar_noalloc: null argument");
  STNLS_REQUIRE(g->B > 0 && g->HD > 0 && g->T > 0 && g->Tq > 0 && g->Tk > 0 && g->F > 0 && g->Fp >= g->F && g->Fp % 4 == 0 &&
                    g->qH > 0 && g->qW > 0 && g->kH > 0 && g->kW > 0,
                "search_forward_planar_noalloc: empty or inconsistent geometry");
  STNLS_REQUIRE(g->tq0 >= 0 && g->tq0 + g->Tq <= g->T && g->tk0 >= 0 && g->tk0 + g->Tk <= g->T,
                "search_forward_planar_noalloc: frame ranges [%d,%d) / [%d,%d) outside the clip of %d frames", g->tq0,
                g->tq0 + g->Tq, g->tk0, g->tk0 + g->Tk, g->T);
  const size_t N = (size_t)g->B * g->HD;
  const size_t q_bytes = round256(N * g->Tq * g->qH * g->qW * g->Fp * sizeof(float));
  const size_t k_bytes = round256(N * g->Tk * g->kH * g->kW * g->Fp * sizeof(float));
  STNLS_REQUIRE(q_bytes + k_bytes <= (size_t)STNLS_NOALLOC_MAX_BYTES,
                "search_forward_planar_noalloc: the channels-last copies of q and k need %zu bytes, the built-in scratch "
                "holds %zu (STNLS_NOALLOC_MAX_BYTES); use stnls_pack_cl + stnls_search_forward with your own buffers",
                q_bytes + k_bytes, (size_t)STNLS_NOALLOC_MAX_BYTES);
  float* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc != STNLS_OK) return rc;
  float* q_cl = scratch;
  float* k_cl = scratch + q_bytes / sizeof(float);
  // planar videos (B, T, HD*F, H, W), contiguous: element (b, hd, t, c, h, w)
  const int64_t Fall = (int64_t)g->HD * g->F;
  const int64_t qhw = (int64_t)g->qH * g->qW, khw = (int64_t)g->kH * g->kW;
  rc = stnls_pack_cl(q_cl, q, g->B, g->HD, g->tq0, g->Tq, g->F, g->Fp, g->qH, g->qW, (int64_t)g->T * Fall * qhw,
                     (int64_t)g->F * qhw, Fall * qhw, qhw, nullptr, nullptr, stream);
  if (rc != STNLS_OK) return rc;
  rc = stnls_pack_cl(k_cl, k, g->B, g->HD, g->tk0, g->Tk, g->F, g->Fp, g->kH, g->kW, (int64_t)g->T * Fall * khw,
                     (int64_t)g->F * khw, Fall * khw, khw, nullptr, nullptr, stream);
  if (rc != STNLS_OK) return rc;
  return stnls_search_forward(q_cl, k_cl, flows, dists, inds, nullptr, nullptr, g, o, stream);
}
