// Host side of the tile forward kernel (search_fwd_tile.cuh): applicability test, shared-memory carve, the TMA tensor
// map over the channels-last key video, launch.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cstdlib>
#include <cstring>
#include "search_fwd_tile.cuh"

namespace stnls_fwd {

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime (no link-time dependency on libcuda)
EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    (void)cudaGetLastError();
  }
  return fn;
}

int round_up(int x, int m) { return (x + m - 1) / m * m; }

// shared memory one CTA needs with `nbuf` box buffers; fills the carve
int carve(const FwdParams& P, TileCfg& C, int nbuf) {
  const int Fp = P.g.Fp, QW = 32 / Fp;
  C.nbuf = nbuf;
  C.box_tx = C.BH * C.BW * Fp * 4;
  C.box_bytes = round_up(C.box_tx, 128);
  int off = nbuf * C.box_bytes;
  C.off_desc = off; off += TILE_NQ * P.W_t * 16;
  C.off_pk = off; off += TILE_NQP * TILE_CAP * 4;   // per-query pool: keys ...
  C.off_pi = off; off += TILE_NQP * TILE_CAP * 4;   // ... and candidate indices
  C.QT = (TILE_TQ - 1) * P.o.stride0 + P.o.ps;
  C.off_qt = off; off += round_up(C.QT * C.QT * Fp * 4, 16);
  // transpose scratch: candidates of one flush x (Fp + 4) floats per query; the query stride keeps the sub-groups of a
  // warp on different banks
  const int ws = P.o.ws, half = P.o.stride1 == 0.5f;
  const int WE = half ? (ws + 1) / 2 : ws, WO = half ? ws / 2 : ws;
  int nv = 0;   // most candidates any (sub-)window hands over per flush
  for (int a = 0; a < 2; a++)
    for (int b = 0; b < 2; b++) {
      const int wh = a ? WO : WE, ww = b ? WO : WE;
      if (wh > 0 && ww > 0 && tile_flush_size(wh, ww, Fp) > nv) nv = tile_flush_size(wh, ww, Fp);
    }
  C.cand_pitch = Fp + 4;
  C.q_stride = nv * C.cand_pitch;
  while ((C.q_stride & 31) != (Fp & 31)) C.q_stride++;
  C.scr_warp = QW * C.q_stride;
  const bool rawdot = P.o.ps == 1 && P.o.dist_type == 0 && !half;   // that variant sums channels inside a lane: no scratch
  C.off_scr = off; off += rawdot ? 0 : (TILE_THREADS / 32) * C.scr_warp * 4;
  C.off_misc = off; off += (4 + TILE_MAXWT * 8 + TILE_MAXWT * 2 + 5 * TILE_NQP) * 4;
  C.off_akey = off; off += P.o.anchor_self ? TILE_NQ * P.W_t * 4 : 0;
  C.off_apos = off; off += P.o.anchor_self ? TILE_NQ * P.W_t * 4 : 0;
  C.smem_bytes = round_up(off, 128);
  return C.smem_bytes;
}

}  // namespace

int launch_tile(const FwdParams& P, cudaStream_t st) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  const char* env = getenv("STNLS_FWD_TILE");
  if (env && env[0] == '0') return 0;
  // ---- does this path apply? (everything else keeps the table / direct kernels) ----
  if (o.itype_int || o.wr > 0 || P.W_t <= 0 || P.W_t > TILE_MAXWT) return 0;
  if (o.dilation != 1 || !o.reflect_bounds || o.K <= 0 || o.off_Hq != 0 || o.off_Wq != 0) return 0;
  if (g.qH != g.kH || g.qW != g.kW || o.strideQ != o.stride0 || o.stride0 < 1) return 0;
  if (!(o.ps == 1 || o.ps == 3) || !(o.stride1 == 1.0f || o.stride1 == 0.5f)) return 0;
  if (!(g.Fp == 8 || g.Fp == 16) || o.ws > 15 || !(o.ws & 1)) return 0;
  const int LK = P.Kout - (o.anchor_self ? 1 : 0);
  if (LK < 1 || LK > 32) return 0;
  if ((int64_t)g.kH * g.kW * g.Fp >= (1LL << 31)) return 0;
  EncodeTiledFn enc = encode_tiled();
  if (!enc) return 0;

  TileCfg C;
  memset(&C, 0, sizeof(C));
  const bool half = o.stride1 == 0.5f;
  C.NRq = (half ? (o.ws + 1) / 2 : o.ws) + o.ps;
  const char* env_rb = getenv("STNLS_FWD_TILE_RB");
  C.RB = env_rb ? atoi(env_rb) : 6;
  if (C.RB < 0) C.RB = 0;
  C.BH = C.BW = (TILE_TQ - 1) * o.stride0 + C.NRq + C.RB;
  if (C.BH > 256) return 0;
  // one reflection must be enough for every pixel a valid candidate touches, and the mirror sources must lie inside
  // the box: frames smaller than the box keep the table kernel
  if (g.kH < C.BH || g.kW < C.BW) return 0;
  C.LK = LK;
  C.tiles_h = (g.nH + TILE_TQ - 1) / TILE_TQ;
  C.tiles_w = (g.nW + TILE_TQ - 1) / TILE_TQ;
  const int64_t blocks = (int64_t)C.tiles_h * C.tiles_w * g.Tq * g.B * g.HD;
  if (blocks <= 0 || blocks >= 2147483647LL) return 0;
  // two box buffers (the next frame slot loads while this one is searched) when two CTAs per SM still fit
  // (113 KB each), else one buffer; one CTA per SM as the last resort
  const char* env_nb = getenv("STNLS_FWD_TILE_NBUF");
  bool planned = false;
  for (int limit = 113 * 1024; limit <= 227 * 1024 && !planned; limit += 114 * 1024)
    for (int nb = 2; nb >= 1 && !planned; nb--) {
      if (env_nb && atoi(env_nb) != nb) continue;
      if (carve(P, C, nb) + 1024 <= limit) planned = true;
    }
  if (!planned) return 0;

  // ---- tensor map: k_cl as (Fp, kW, kH, N*Tk), box (Fp, BW, BH, 1), zero fill outside ----
  CUtensorMap map;
  const cuuint64_t gdim[4] = {(cuuint64_t)g.Fp, (cuuint64_t)g.kW, (cuuint64_t)g.kH, (cuuint64_t)g.B * g.HD * g.Tk};
  const cuuint64_t gstr[3] = {(cuuint64_t)g.Fp * 4, (cuuint64_t)g.kW * g.Fp * 4, (cuuint64_t)g.kH * g.kW * g.Fp * 4};
  const cuuint32_t box[4] = {(cuuint32_t)g.Fp, (cuuint32_t)C.BW, (cuuint32_t)C.BH, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  if ((reinterpret_cast<uintptr_t>(P.k) & 15) != 0) return 0;
  const CUresult rc = enc(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(P.k), gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) return 0;
  return o.dist_type == 0 ? launch_tile_prod(P, C, map, st) : launch_tile_l2(P, C, map, st);
}

}  // namespace stnls_fwd
