// "Tile" forward search (float NonLocalSearch): a CTA owns an 8x8 tile of queries of one (batch, head, frame).
//
//   * geometry once per (query, frame slot), one THREAD each (window placement, validity masks, anchor slot) --
//     the table kernel spent a whole warp per query on it;
//   * per frame slot the raw key pixels every query of the tile can touch -- one box of the channels-last key
//     video, tile + search reach + flow spread -- are staged in shared memory by ONE TMA load
//     (cp.async.bulk.tensor over a 4-D tensor map of k_cl, mbarrier completion, double-buffered over the frame
//     slots where shared memory allows).  Rows/columns outside the frame arrive as zeros and are filled with the
//     MIRRORED pixels, so reflected patch pixels (shared_kernel.cu:49-60) need no special case:
//     bilinear interpolation of the mirrored image at p equals interpolation of the image at reflect(p);
//   * lane = CHANNEL, 32/Fp queries per warp.  A lane walks the lattice rows of its query's window keeping two
//     raw key rows in registers (rolling), interpolates each lattice point once (separable, 4 flops) and feeds it
//     to the <= ps*ps candidates that cover it: per (query, frame, channel) ~(ws+ps)^2 LDS.32 and
//     4 (ws+ps-1)^2 + (1|2) ps^2 ws^2 flops -- no interpolated table in shared memory, no per-query global loads;
//   * the per-channel partial sums of finished candidate rows are transposed through a small per-warp scratch,
//     one lane per candidate sums the Fp partials; candidates that are not worse than the query's K-th best so far
//     are APPENDED (ballot + popc slot) to the query's 64-entry pool in shared memory.  Only when a pool is about to
//     overflow, and once after the last frame slot, the query's Fp lanes compact it to the K best, sorted
//     (tile_merge: lanes sort their share in registers, K butterfly-min rounds, ties to the lower candidate index)
//     and the K-th key becomes the new admission threshold -- about three compactions per query;
//   * stride1 = 0.5: candidates split by the parity of their window row / column into four integer-spaced
//     sub-windows, each with its own fractional phase.
// A query whose window does not fit the staged box for some frame slot (flows that vary by more than the budget
// inside one tile), or whose flow is not finite, is handed to direct_query() (search_fwd_direct.cuh) by the same
// CTA: complete semantics, no restriction.  Distances are summed channel-parallel, so this kernel is for
// itype="float" only (tolerance 1e-4); the int path keeps the bit-exact table kernel.
// Reference semantics: SURVEY.md appendix A; lib/csrc/search/non_local_search_bilin2d_kernel.cu:20-193,
// lib/csrc/search/nls_bilin2d.cu:3-90, lib/csrc/shared_kernel.cu:248-287,738-771,
// lib/csrc/nn/anchor_self_kernel.cu:24-116, lib/stnls/nn/topk.py:65-123.
#pragma once
#include <cuda.h>
#include "search_fwd_direct.cuh"

namespace stnls_fwd {

struct TileCfg {
  int BH, BW;              // staged key box, pixels
  int NRq;                 // raw rows / columns one query's window touches
  int RB;                  // budget for the spread of the flows inside one tile, pixels
  int nbuf;                // box buffers (2: the next frame slot loads while this one is searched)
  int tiles_h, tiles_w;
  int LK;                  // entries of the per-query list (K, or K-1 when slot 0 is the anchor)
  int QT;                  // staged query tile: QT x QT pixels
  int off_qt;
  int box_bytes;           // one buffer, multiple of 128
  int box_tx;              // bytes one TMA box delivers (BH*BW*Fp*4)
  int off_desc, off_pk, off_pi, off_scr, off_misc, off_akey, off_apos;   // byte offsets into dynamic shared memory
  int scr_warp;            // floats of transpose scratch per warp
  int cand_pitch;          // floats between two candidates in the scratch (Fp + 4: conflict-free LDS.128)
  int q_stride;            // floats between the scratch areas of two queries of a warp
  int smem_bytes;
};

// host side (search_fwd_tile_host.cu): builds the tensor map + configuration and launches; returns
//   1 launched, 0 this path does not apply to (P) -- fall through to the other kernels, -1 CUDA error
int launch_tile(const FwdParams& P, cudaStream_t st);
// one per instantiation unit
int launch_tile_prod(const FwdParams& P, const TileCfg& C, const CUtensorMap& map, cudaStream_t st);
int launch_tile_l2(const FwdParams& P, const TileCfg& C, const CUtensorMap& map, cudaStream_t st);

#ifdef __CUDACC__

constexpr int TILE_TQ = 8, TILE_NQ = 64, TILE_THREADS = 256, TILE_MAXWT = 16;
constexpr int TILE_NQP = 65;   // per-query arrays: 64 queries + a dummy entry (index 64) for lanes without a query

// ---- mbarrier / TMA (PTX) ---------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(unsigned a, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned a, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned a, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
      : "=r"(ok)
      : "r"(a), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void tma_load_box(unsigned dst, const CUtensorMap* map, unsigned mbar, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(mbar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ unsigned f2ord_t(float f) {
  unsigned b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f_t(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// ---- the per-lane core: one channel of one query, one integer-spaced (sub-)window of WH x WW candidates ----------
// bp: this lane's channel of the raw key pixel (r0, c0) in the staged box; lattice point (x, y) sits at raw position
// (r0 + x + fh, c0 + y + fw).  Rows of finished candidates are handed to `flush` in groups of G rows.
template <int DIST>
__device__ __forceinline__ float acc1(float acc, float qv, float kv) {
  if (DIST == 0) return fmaf(qv, kv, acc);
  const float d = qv - kv;
  return fmaf(d, d, acc);
}

// Finished candidates are handed to `flush(lin0, nv, vals)` in batches of FL (lin0 = row-major index of vals[0] inside
// this sub-window): FL = the whole window when it is small, else the number of lanes that sum channels (FP), so that
// every reader lane has a candidate in every flush (rows of 9 candidates would leave 7 of 16 lanes idle).
__host__ __device__ constexpr int tile_flush_size(int WH, int WW, int FP) { return (WH * WW <= 2 * FP) ? WH * WW : FP; }

template <int WH, int WW, int PS, int DIST, int FP, int FL, class Flush>
__device__ __forceinline__ void tile_core(const float* __restrict__ bp, const int rowpitch, const float fh, const float fw,
                                          const float (&qv)[PS * PS], Flush&& flush) {
  constexpr int MR = WH + PS - 1, MC = WW + PS - 1;
  constexpr int PMAX = FL + WW;   // pending finished candidates: fewer than FL left over + one new row
  const float gh0 = 1.f - fh, gw0 = 1.f - fw;
  float prev[MC + 1], cur[MC + 1];
#pragma unroll
  for (int y = 0; y <= MC; y++) prev[y] = bp[y * FP];
  float acc[PS][WW];
#pragma unroll
  for (int s = 0; s < PS; s++)
#pragma unroll
    for (int j = 0; j < WW; j++) acc[s][j] = 0.f;
  float pend[PMAX];
  int np = 0, lin0 = 0;   // compile-time values once the row loop is unrolled
#pragma unroll
  for (int x = 0; x < MR; x++) {
    const float* rp = bp + (x + 1) * rowpitch;
#pragma unroll
    for (int y = 0; y <= MC; y++) cur[y] = rp[y * FP];
    float L[MC];
    {
      float v0 = fmaf(fh, cur[0], gh0 * prev[0]);
#pragma unroll
      for (int y = 0; y < MC; y++) {
        const float v1 = fmaf(fh, cur[y + 1], gh0 * prev[y + 1]);
        L[y] = fmaf(fw, v1, gw0 * v0);
        v0 = v1;
      }
    }
#pragma unroll
    for (int pi = 0; pi < PS; pi++) {
      const int i = x - pi;
      if (i >= 0 && i < WH) {
#pragma unroll
        for (int j = 0; j < WW; j++)
#pragma unroll
          for (int pj = 0; pj < PS; pj++) acc[i % PS][j] = acc1<DIST>(acc[i % PS][j], qv[pi * PS + pj], L[j + pj]);
      }
    }
    if (x >= PS - 1) {
      const int i = x - (PS - 1);   // candidate row i is complete
#pragma unroll
      for (int j = 0; j < WW; j++) {
        pend[np + j] = acc[i % PS][j];
        acc[i % PS][j] = 0.f;
      }
      np += WW;
#pragma unroll
      for (int rep = 0; rep < (PMAX + FL - 1) / FL; rep++) {
        if (np >= FL || (i == WH - 1 && np > 0)) {
          const int nv = np < FL ? np : FL;
          flush(lin0, nv, pend);
#pragma unroll
          for (int j = 0; j < PMAX - FL; j++) pend[j] = pend[j + FL];
          lin0 += nv;
          np -= nv;
        }
      }
    }
#pragma unroll
    for (int y = 0; y <= MC; y++) prev[y] = cur[y];
  }
}

// The same core with the steady-state lattice rows as a ROLLED loop over groups of PS rows (needs WH % PS == 0): inside
// a group the accumulator of a candidate row (i % PS) and the fill of the pending batch are compile-time again -- a group
// starts with an empty batch and ends by flushing what is left -- so nothing is indexed dynamically, and the code is
// (PS - 1) + PS lattice rows and two or three flush sites instead of WH + PS - 1 rows and WH * WW / FL sites.  At
// ws = 9, ps = 3 the unrolled core is ~5 k instructions that every warp streams through once per (query, frame): ncu
// shows `no_instruction` as its second stall reason (profiles/r02_z_ncu_final.json).
template <int WH, int WW, int PS, int DIST, int FP, int FL, class Flush>
__device__ __forceinline__ void tile_core_rolled(const float* __restrict__ bp, const int rowpitch, const float fh,
                                                 const float fw, const float (&qv)[PS * PS], Flush&& flush) {
  static_assert(PS > 1 && WH % PS == 0 && WH >= 2 * PS, "tile_core_rolled: whole groups of PS candidate rows");
  constexpr int MC = WW + PS - 1, G = WH / PS;
  constexpr int PMAX = FL + WW;
  const float gh0 = 1.f - fh, gw0 = 1.f - fw;
  float prev[MC + 1], cur[MC + 1];
#pragma unroll
  for (int y = 0; y <= MC; y++) prev[y] = bp[y * FP];
  float acc[PS][WW];
#pragma unroll
  for (int s = 0; s < PS; s++)
#pragma unroll
    for (int j = 0; j < WW; j++) acc[s][j] = 0.f;
  float L[MC];
  auto lattice_row = [&](const float* rp) {   // raw row -> interpolated lattice row L, raw row becomes `prev`
#pragma unroll
    for (int y = 0; y <= MC; y++) cur[y] = rp[y * FP];
    float v0 = fmaf(fh, cur[0], gh0 * prev[0]);
#pragma unroll
    for (int y = 0; y < MC; y++) {
      const float v1 = fmaf(fh, cur[y + 1], gh0 * prev[y + 1]);
      L[y] = fmaf(fw, v1, gw0 * v0);
      v0 = v1;
    }
#pragma unroll
    for (int y = 0; y <= MC; y++) prev[y] = cur[y];
  };
  // head: lattice rows 0 .. PS-2 feed candidate rows 0 .. PS-2, none is complete yet
#pragma unroll
  for (int x = 0; x < PS - 1; x++) {
    lattice_row(bp + (x + 1) * rowpitch);
#pragma unroll
    for (int pi = 0; pi <= x; pi++)
#pragma unroll
      for (int j = 0; j < WW; j++)
#pragma unroll
        for (int pj = 0; pj < PS; pj++) acc[x - pi][j] = acc1<DIST>(acc[x - pi][j], qv[pi * PS + pj], L[j + pj]);
  }
#pragma unroll 1
  for (int g = 0; g < G; g++) {
    float pend[PMAX];
    int np = 0;                       // compile-time inside the group
    int lin0 = g * PS * WW;
    const float* rp = bp + (PS + PS * g) * rowpitch;   // raw row x + 1 of lattice row x = PS - 1 + PS * g
#pragma unroll
    for (int r = 0; r < PS; r++) {
      lattice_row(rp + r * rowpitch);
#pragma unroll
      for (int pi = 0; pi < PS; pi++) {
        const int s = r + PS - 1 - pi;          // candidate row PS * g + s
        if (s < PS || g + 1 < G) {               // s >= PS: a row of the NEXT group -- there is none after the last
#pragma unroll
          for (int j = 0; j < WW; j++)
#pragma unroll
            for (int pj = 0; pj < PS; pj++) acc[s % PS][j] = acc1<DIST>(acc[s % PS][j], qv[pi * PS + pj], L[j + pj]);
        }
      }
      // candidate row PS * g + r is complete
#pragma unroll
      for (int j = 0; j < WW; j++) {
        pend[np + j] = acc[r][j];
        acc[r][j] = 0.f;
      }
      np += WW;
#pragma unroll
      for (int rep = 0; rep < (PMAX + FL - 1) / FL; rep++) {
        if (np >= FL || (r == PS - 1 && np > 0)) {
          const int nv = np < FL ? np : FL;
          flush(lin0, nv, pend);
#pragma unroll
          for (int j = 0; j < PMAX - FL; j++) pend[j] = pend[j + FL];
          lin0 += nv;
          np -= nv;
        }
      }
    }
  }
}

// ---- pool compaction: the K best of a query's pool, sorted, to the head of the pool ------------------------------
// Called by whole warps; each sub-group of GS lanes works on its own query (pk/px/cnt/thr are that query's); the group
// minima are warp-wide redux.sync with the other groups' lanes neutral.  Deliberately NOT inlined: it runs a few times per query,
// from inside the unrolled core, and must not add to the core's register pressure.
constexpr int TILE_CAP = 64;   // pool entries per query

template <int GS>
__device__ __noinline__ void tile_merge(unsigned* pk, int* px, int* cnt_ptr, unsigned* thr_ptr, const int LK, const int chl) {
  constexpr int E = (TILE_CAP + GS - 1) / GS;
  constexpr int QW = 32 / GS;                      // groups per warp; lanes beyond QW * GS idle along
  const int sub = (threadIdx.x & 31) / GS;
  const bool live = sub < QW;
  const int cnt = live ? *cnt_ptr : 0;
  unsigned k[E];
  int x[E];
#pragma unroll
  for (int j = 0; j < E; j++) {
    const int e = chl + j * GS;
    k[j] = (e < cnt) ? pk[e] : 0xffffffffu;
    x[j] = (e < cnt) ? px[e] : 0x7fffffff;
  }
  __syncwarp();
  // this lane's share, ascending by (key, candidate index): odd-even transposition network
#pragma unroll
  for (int r = 0; r < E; r++) {
#pragma unroll
    for (int a = (r & 1); a + 1 < E; a += 2) {
      const bool sw = (k[a] > k[a + 1]) || (k[a] == k[a + 1] && x[a] > x[a + 1]);
      const unsigned tk = sw ? k[a + 1] : k[a], uk = sw ? k[a] : k[a + 1];
      const int tx = sw ? x[a + 1] : x[a], ux = sw ? x[a] : x[a + 1];
      k[a] = tk; k[a + 1] = uk; x[a] = tx; x[a + 1] = ux;
    }
  }
  unsigned last = 0xffffffffu;
  // min over the GS lanes of a sub-group: one warp-wide redux.sync per sub-group (the other sub-groups contribute the
  // neutral element) -- one instruction deep, where a shuffle butterfly is log2(FP) dependent steps
  auto group_min = [&](unsigned v) {
    unsigned r = 0xffffffffu;
#pragma unroll
    for (int gsub = 0; gsub < QW; gsub++) {
      const unsigned mg = __reduce_min_sync(FULL_MASK, (sub == gsub) ? v : 0xffffffffu);
      r = (sub == gsub) ? mg : r;
    }
    return r;
  };
  for (int kk = 0; kk < LK; kk++) {
    const unsigned m = group_min(k[0]);
    const unsigned wi = group_min((k[0] == m) ? (unsigned)x[0] : 0x7fffffffu);
    if (chl == 0 && live) { pk[kk] = m; px[kk] = (int)wi; }
    last = m;
    const bool won = (m != 0xffffffffu) && (k[0] == m) && (x[0] == (int)wi);
#pragma unroll
    for (int j = 0; j + 1 < E; j++) {
      k[j] = won ? k[j + 1] : k[j];
      x[j] = won ? x[j + 1] : x[j];
    }
    if (won) { k[E - 1] = 0xffffffffu; x[E - 1] = 0x7fffffff; }
  }
  if (chl == 0 && live) {
    *cnt_ptr = min(cnt, LK);
    *thr_ptr = (cnt >= LK) ? last : 0xffffffffu;   // admission threshold: the K-th best key once K candidates exist
  }
  __syncwarp();
}

template <int DIST, int PS, int WS, bool HALF, int FP>
__global__ void __launch_bounds__(TILE_THREADS, 2)
search_fwd_tile_kernel(const __grid_constant__ FwdParams P, const __grid_constant__ TileCfg C,
                       const __grid_constant__ CUtensorMap tmap) {
  extern __shared__ __align__(128) unsigned char smem[];
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // RAW: ps == 1 with the product distance -- the bilinear blend commutes with the channel contraction, so a lane owns
  // one COLUMN of the (ws+1)^2 raw grid and does whole channel dot products itself (see the RAW pass below); otherwise
  // lane = channel.  GS lanes work on one query, QW queries per warp (lanes beyond QW * GS idle along).
  constexpr bool RAW = (PS == 1 && DIST == 0 && !HALF);
  constexpr int GS = RAW ? WS + 1 : FP;
  constexpr int QW = 32 / GS;                 // queries per warp
  constexpr int NS = (32 + GS - 1) / GS;      // list entries per lane (NS * GS >= 32 >= K)
  constexpr int NPASS = (TILE_NQ + 8 * QW - 1) / (8 * QW);
  const int sub = lane / GS, chl = lane % GS;
  const unsigned submask = (sub < QW) ? (((GS >= 32) ? 0xffffffffu : ((1u << (GS & 31)) - 1u)) << (sub * GS)) : 0u;
  const float s1 = o.stride1;
  const int s0 = o.stride0, po = o.patch_offset, W_t = P.W_t;
  const bool anchored = o.anchor_self != 0;
  const float invalid = (DIST == 0) ? -INFINITY : INFINITY;

  // ---- which tile ----
  const int ntiles = C.tiles_h * C.tiles_w;
  const int tile = blockIdx.x % ntiles, rr = blockIdx.x / ntiles;
  const int tl = rr % g.Tq, n = rr / g.Tq;
  const int tile_h = tile / C.tiles_w, tile_w = tile - tile_h * C.tiles_w;
  const int b = n / g.HD, hd = n - b * g.HD;
  const int t = tl + g.tq0;
  const int t_max = time_tmax(t, g.T, o.wt);
  const int row0 = tile_h * TILE_TQ * s0, col0 = tile_w * TILE_TQ * s0;   // pixel of the tile's first query

  // ---- shared memory ----
  float* const box0 = reinterpret_cast<float*>(smem);
  float4* const desc = reinterpret_cast<float4*>(smem + C.off_desc);       // [W_t][64]: (bh, bw, masks, -)
  unsigned* const poolk = reinterpret_cast<unsigned*>(smem + C.off_pk);    // [64][TILE_CAP] pool: ordered keys ...
  int* const pooli = reinterpret_cast<int*>(smem + C.off_pi);              // [64][TILE_CAP] ... and candidate indices
  float* const qt = reinterpret_cast<float*>(smem + C.off_qt);             // [QT][QT][FP] query pixels (reflected at the border)
  float* const scr = reinterpret_cast<float*>(smem + C.off_scr) + warp * C.scr_warp + sub * C.q_stride;
  int* const misc = reinterpret_cast<int*>(smem + C.off_misc);
  // misc: [0..3] mbarriers (2 x 8 bytes), then ints
  int* const stats = misc + 4;                           // [W_t][8]: min_h max_h sum_h min_w max_w sum_w count -
  int* const origin = stats + TILE_MAXWT * 8;            // [W_t][2]
  int* const state = origin + TILE_MAXWT * 2;            // [64]: 0 searched here, 1 deferred to direct_query, 2 outside the grid
  int* const a_nls = state + TILE_NQP;                   // [64] anchor slot of the query
  float* const adist = reinterpret_cast<float*>(a_nls + TILE_NQP);   // [64] distance of the anchor candidate
  int* const pcnt = a_nls + 2 * TILE_NQP;                // [64] pool entries in use
  unsigned* const pthr = reinterpret_cast<unsigned*>(pcnt + TILE_NQP);   // [64] admission threshold (ordered key)
  float* const akey = reinterpret_cast<float*>(smem + C.off_akey);   // [W_t][64] (anchored only)
  int* const apos = reinterpret_cast<int*>(smem + C.off_apos);
  const unsigned mbar0 = (unsigned)__cvta_generic_to_shared(misc);
  const int LK = C.LK;

  auto query_ctx = [&](int ql, QueryCtx& c) {
    c.b = b; c.hd = hd; c.n = n; c.t = t; c.tl = tl;
    c.nh = tile_h * TILE_TQ + (ql >> 3);
    c.nw = tile_w * TILE_TQ + (ql & 7);
    c.ah = c.nh * s0; c.aw = c.nw * s0;        // nh < nH  =>  nh*s0 <= kH-1: the "% kH" of the reference is a no-op
    c.rh = c.ah; c.rw = c.aw;
    c.t_max = t_max;
    return c.nh < g.nH && c.nw < g.nW;
  };

  // ---- prologue 1: init, query tile ----
  if (tid < TILE_NQP) {
    QueryCtx c;
    state[tid] = (tid < TILE_NQ && query_ctx(tid, c)) ? 0 : 2;
    adist[tid] = invalid;
    a_nls[tid] = 0;
    pcnt[tid] = 0;
    pthr[tid] = 0xffffffffu;
  }
  if (tid < W_t) {
    int* s = stats + tid * 8;
    s[0] = s[3] = 0x7fffffff; s[1] = s[4] = -0x7fffffff; s[2] = s[5] = s[6] = 0;
  }
  {
    const float* const qfr = P.q + ((int64_t)n * g.Tq + tl) * (int64_t)g.qH * g.qW * FP;
    for (int i = tid; i < C.QT * C.QT * (FP / 4); i += TILE_THREADS) {
      const int c4 = i % (FP / 4), px = i / (FP / 4);
      const int y = px / C.QT, x = px - y * C.QT;
      const int r1 = bnd(row0 + y + po, g.qH), r2 = bnd(col0 + x + po, g.qW);
      float4 v = f4zero();
      if (inb(r1, g.qH) && inb(r2, g.qW)) v = ldg4(qfr + ((int64_t)r1 * g.qW + r2) * FP + c4 * 4);
      reinterpret_cast<float4*>(qt)[i] = v;
    }
  }
  if (tid == 0) {
    mbar_init(mbar0, 1);
    mbar_init(mbar0 + 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // ---- prologue 2: geometry of every (query, frame slot), one thread each ----
  for (int item = tid; item < TILE_NQ * W_t; item += TILE_THREADS) {
    const int st = item >> 6, ql = item & 63;
    QueryCtx c;
    float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
    if (query_ctx(ql, c)) {
      const Group<float> gp = group_params<false, false>(P, c, st);
      const float bh = fmaf(-s1, gp.offh, gp.ah) + (float)po;
      const float bw = fmaf(-s1, gp.offw, gp.aw) + (float)po;
      unsigned hm = 0, wm = 0;
#pragma unroll
      for (int i = 0; i < WS; i++) {
        hm |= (inb(fmaf(s1, (float)i - gp.offh, gp.ah), g.kH) ? 1u : 0u) << i;
        wm |= (inb(fmaf(s1, (float)i - gp.offw, gp.aw), g.kW) ? 1u : 0u) << i;
      }
      d = make_float4(bh, bw, __uint_as_float(hm | (wm << 16)), 0.f);
      if (fabsf(bh) < 1e8f && fabsf(bw) < 1e8f) {
        int* s = stats + st * 8;
        const int dh = (int)floorf(bh) - (row0 + (ql >> 3) * s0), dw = (int)floorf(bw) - (col0 + (ql & 7) * s0);
        atomicMin(s + 0, dh); atomicMax(s + 1, dh); atomicAdd(s + 2, dh);
        atomicMin(s + 3, dw); atomicMax(s + 4, dw); atomicAdd(s + 5, dw);
        atomicAdd(s + 6, 1);
      } else {
        state[ql] = 1;   // non-finite flow: the generic path reproduces the reference's handling
      }
      if (anchored) {   // anchor_self_kernel.cu:62-89: first candidate with offset 0, else the nearest (L1)
        float bkey = INFINITY;
        int bpos = 0x7fffffff;
        const float zero[3] = {0.f, 0.f, 0.f};
        for (int s = 0; s < WS * WS; s++) {
          const int wi = s / WS, wj = s - wi * WS;
          float ind[3];
          ind[0] = (float)gp.t - (float)c.t;
          ind[1] = fmaf(s1, (float)wi - gp.offh, gp.ah) - (float)c.ah;
          ind[2] = fmaf(s1, (float)wj - gp.offw, gp.aw) - (float)c.aw;
          const float key = anchor_key(ind, zero);
          if (key < bkey) { bkey = key; bpos = st * WS * WS + s; }
        }
        akey[item] = bkey;
        apos[item] = bpos;
      }
    }
    desc[item] = d;
  }
  __syncthreads();
  // box origin per frame slot: the smallest flow-displaced window start of the tile when the spread fits the budget,
  // else centred on the mean (the queries that then fall outside are deferred)
  if (tid < W_t) {
    const int* s = stats + tid * 8;
    int oh = 0, ow = 0;
    if (s[6] > 0) {
      oh = (s[1] - s[0] <= C.RB) ? s[0] : (int)floorf((float)s[2] / (float)s[6]) - C.RB / 2;
      ow = (s[4] - s[3] <= C.RB) ? s[3] : (int)floorf((float)s[5] / (float)s[6]) - C.RB / 2;
    }
    origin[tid * 2] = row0 + oh;
    origin[tid * 2 + 1] = col0 + ow;
  }
  if (anchored && tid < TILE_NQ) {
    float bkey = INFINITY;
    int bpos = 0x7fffffff;
    for (int st = 0; st < W_t; st++) {
      const float k2 = akey[st * 64 + tid];
      if (k2 < bkey) { bkey = k2; bpos = apos[st * 64 + tid]; }
    }
    a_nls[tid] = (bpos == 0x7fffffff) ? 0 : bpos;
  }
  __syncthreads();
  for (int item = tid; item < TILE_NQ * W_t; item += TILE_THREADS) {
    const int st = item >> 6, ql = item & 63;
    if (state[ql] != 0) continue;
    const float4 d = desc[item];
    const int r0 = (int)floorf(d.x) - origin[st * 2], c0 = (int)floorf(d.y) - origin[st * 2 + 1];
    if (r0 < 0 || c0 < 0 || r0 + C.NRq > C.BH || c0 + C.NRq > C.BW) state[ql] = 1;
  }
  __syncthreads();

  // ---- frame slots: TMA box -> mirror fill -> search ----
  const unsigned box_sa = (unsigned)__cvta_generic_to_shared(box0);
  auto issue = [&](int st) {   // one thread
    const int t_next = t + st;
    const int tp = (t_next > t_max) ? t_max - st : t_next;
    const int buf = st % C.nbuf;
    fence_proxy_async();        // the generic-proxy reads / mirror writes of this buffer are ordered before the refill
    mbar_expect_tx(mbar0 + 8 * buf, (unsigned)C.box_tx);
    tma_load_box(box_sa + buf * C.box_bytes, &tmap, mbar0 + 8 * buf, 0, origin[st * 2 + 1], origin[st * 2],
                 n * g.Tk + (tp - g.tk0));
  };
  if (tid == 0) {
    for (int st = 0; st < C.nbuf && st < W_t; st++) issue(st);
  }

  const int rowpitch = C.BW * FP;

  for (int st = 0; st < W_t; st++) {
    const int buf = st % C.nbuf;
    const unsigned par = (unsigned)(st / C.nbuf) & 1u;
    for (int spin = 0; !mbar_try_wait(mbar0 + 8 * buf, par); spin++)
      if (spin > (1 << 22)) __trap();   // a TMA that never completes must not hang the device
    float* const box = box0 + buf * (C.box_bytes >> 2);
    const int oh = origin[st * 2], ow = origin[st * 2 + 1];
    if (oh < 0 || ow < 0 || oh + C.BH > g.kH || ow + C.BW > g.kW) {
      // TMA filled the cells outside the frame with zeros: put the mirrored pixels there (reflect once)
      for (int cell = tid; cell < C.BH * C.BW; cell += TILE_THREADS) {
        const int y = cell / C.BW, x = cell - y * C.BW;
        const int gy = oh + y, gx = ow + x;
        if (inb(gy, g.kH) && inb(gx, g.kW)) continue;
        const int ry = bnd(gy, g.kH), rx = bnd(gx, g.kW);
        const int sy = ry - oh, sx = rx - ow;
        if (!inb(ry, g.kH) || !inb(rx, g.kW) || sy < 0 || sy >= C.BH || sx < 0 || sx >= C.BW) continue;
        const float4* src = reinterpret_cast<const float4*>(box + (sy * C.BW + sx) * FP);
        float4* dst = reinterpret_cast<float4*>(box + cell * FP);
#pragma unroll
        for (int c4 = 0; c4 < FP / 4; c4++) dst[c4] = src[c4];
      }
      __syncthreads();
    }
    for (int pass = 0; pass < NPASS; pass++) {
      const int qlr = (pass * 8 + warp) * QW + sub;
      const bool qvalid = sub < QW && qlr < TILE_NQ;
      const int ql = qvalid ? qlr : TILE_NQ;     // per-query state (the dummy entry for lanes without a query)
      const int qls = qvalid ? qlr : 0;          // read-only tables
      const bool active = state[ql] == 0;
      const float4 dsc = desc[st * 64 + qls];
      const unsigned masks = __float_as_uint(dsc.z);
      const unsigned hm = masks & 0xffffu, wm = masks >> 16;
      const int an = a_nls[ql];
      unsigned* const pk = poolk + ql * TILE_CAP;
      int* const px = pooli + ql * TILE_CAP;
      int cnt = pcnt[ql];            // register copies, uniform inside the sub-group
      unsigned thr = pthr[ql];
      if constexpr (RAW) {
        // ---- ps == 1, product: d(i,j) = bilinear blend of the raw dot products D(x,y) = <q, k_raw(r0+x, c0+y)> on the
        // (ws+1)^2 integer grid.  Lane = grid column y, loop over grid rows: 4 LDS.128 + Fp FMAs per grid point, no sum
        // across lanes; the right neighbour's D comes by one shuffle, the previous row's blend stays in a register.
        // Lanes walk the channel quads in an order rotated by their column so that the 16-byte reads of neighbouring
        // columns (one pixel = Fp*4 bytes apart) fall on different banks.
        constexpr int NQD = FP / 4;
        const int y = chl;
        const int rot = (y >> 1) & (NQD - 1);
        float4 qr[NQD];
        {
          const float* qp = qt + (((qls >> 3) * s0) * C.QT + (qls & 7) * s0) * FP;
#pragma unroll
          for (int t2 = 0; t2 < NQD; t2++) qr[t2] = *reinterpret_cast<const float4*>(qp + (((t2 + rot) & (NQD - 1)) << 2));
        }
        const float flh = floorf(dsc.x), flw = floorf(dsc.y);
        const float fh = dsc.x - flh, fw = dsc.y - flw, gh0 = 1.f - fh, gw0 = 1.f - fw;
        const int r0 = active ? (int)flh - oh : 0, c0 = active ? (int)flw - ow : 0;
        const float* bp = box + (r0 * C.BW + c0 + y) * FP;
        float Hprev = 0.f;
#pragma unroll
        for (int x = 0; x <= WS; x++) {
          const float* pp = bp + x * rowpitch;
          float Dq[NQD];   // one short FMA chain per channel quad (a single 4*NQD-deep chain is latency-bound)
#pragma unroll
          for (int t2 = 0; t2 < NQD; t2++) {
            const float4 kv = *reinterpret_cast<const float4*>(pp + (((t2 + rot) & (NQD - 1)) << 2));
            Dq[t2] = fmaf(qr[t2].w, kv.w, fmaf(qr[t2].z, kv.z, fmaf(qr[t2].y, kv.y, qr[t2].x * kv.x)));
          }
          float D = Dq[0];
#pragma unroll
          for (int t2 = 1; t2 < NQD; t2++) D += Dq[t2];
          const float Dn = __shfl_down_sync(FULL_MASK, D, 1);
          const float H = fmaf(fw, Dn, gw0 * D);
          if (x >= 1) {
            float d = fmaf(fh, H, gh0 * Hprev);
            const int wi = x - 1, wj = y;
            unsigned key = 0xffffffffu;
            int eff = 0;
            if (y < WS) {
              if (!((hm >> wi) & (wm >> wj) & 1u)) d = invalid;
              eff = st * WS * WS + wi * WS + wj;
              key = min(f2ord_t(-d + 0.f), 0xfffffffeu);
              if (anchored) {
                if (eff == an) { adist[ql] = d; key = 0xffffffffu; }
                else if (eff == 0) eff = an;
              }
            }
            if (__any_sync(FULL_MASK, cnt + GS > TILE_CAP)) {
              if (chl == 0) pcnt[ql] = cnt;
              __syncwarp();
              tile_merge<GS>(pk, px, pcnt + ql, pthr + ql, LK, chl);
              cnt = pcnt[ql];
              thr = pthr[ql];
            }
            const bool pass_ = active && key <= thr && key != 0xffffffffu;
            const unsigned mine = __ballot_sync(FULL_MASK, pass_) & submask;
            if (pass_) {
              const int slot = cnt + __popc(mine & ((1u << lane) - 1u));
              pk[slot] = key;
              px[slot] = eff;
            }
            cnt += __popc(mine);
          }
          Hprev = H;
        }
        if (chl == 0) pcnt[ql] = cnt;
        continue;
      }
      // query patch, this lane's channel
      float qv[PS * PS];
      {
        const float* qp = qt + (((qls >> 3) * s0) * C.QT + (qls & 7) * s0) * FP + (RAW ? 0 : chl);
#pragma unroll
        for (int pi = 0; pi < PS; pi++)
#pragma unroll
          for (int pj = 0; pj < PS; pj++) qv[pi * PS + pj] = qp[(pi * C.QT + pj) * FP];
      }

      // one integer-spaced (sub-)window: candidates (eh + step*i, ew + step*j) of the ws x ws window
      auto run_window = [&](auto WHc, auto WWc, const int eh, const int ew) {
        constexpr int WH = decltype(WHc)::value, WW = decltype(WWc)::value;
        constexpr int FL = tile_flush_size(WH, WW, FP);
        constexpr int STEP = HALF ? 2 : 1;
        const float ph = HALF ? dsc.x + 0.5f * (float)eh : dsc.x, pw = HALF ? dsc.y + 0.5f * (float)ew : dsc.y;
        const float flh = floorf(ph), flw = floorf(pw);
        const int r0 = active ? (int)flh - oh : 0, c0 = active ? (int)flw - ow : 0;
        const float* bp = box + (r0 * C.BW + c0) * FP + chl;
        auto flush = [&](const int lin0, const int nv, const float* vals) {
#pragma unroll
          for (int v = 0; v < FL; v++)
            if (v < nv) scr[v * C.cand_pitch + chl] = vals[v];
          __syncwarp();
#pragma unroll
          for (int m = 0; m < (FL + FP - 1) / FP; m++) {
            const int v = chl + m * FP;
            unsigned key = 0xffffffffu;
            int eff = 0;
            if (v < nv) {
              const float4* sp = reinterpret_cast<const float4*>(scr + v * C.cand_pitch);
              float d = 0.f;
#pragma unroll
              for (int c4 = 0; c4 < FP / 4; c4++) {
                const float4 x4 = sp[c4];
                d += (x4.x + x4.y) + (x4.z + x4.w);
              }
              const int lin = lin0 + v;
              const int vi = lin / WW, vj = lin - vi * WW;
              const int wi = eh + STEP * vi, wj = ew + STEP * vj;
              if (!((hm >> wi) & (wm >> wj) & 1u)) d = invalid;
              eff = st * WS * WS + wi * WS + wj;
              key = min(f2ord_t(((DIST == 0) ? -d : d) + 0.f), 0xfffffffeu);
              if (anchored) {   // anchor swap (anchor_self_kernel.cu:101-113): the anchor is reported in slot 0 and is not
                                // a member of the list; the candidate it displaces from position 0 takes its index
                if (eff == an) { adist[ql] = d; key = 0xffffffffu; }
                else if (eff == 0) eff = an;
              }
            }
            // room for one entry per lane?  (rare; warp-uniform, each sub-group compacts its own pool)
            if (__any_sync(FULL_MASK, cnt + (FL < FP ? FL : FP) > TILE_CAP)) {   // room for what one round can append
              if (chl == 0) pcnt[ql] = cnt;
              __syncwarp();
              tile_merge<GS>(pk, px, pcnt + ql, pthr + ql, LK, chl);
              cnt = pcnt[ql];
              thr = pthr[ql];
            }
            const bool pass = active && key <= thr && key != 0xffffffffu;
            const unsigned mine = __ballot_sync(FULL_MASK, pass) & submask;
            if (pass) {
              const int slot = cnt + __popc(mine & ((1u << lane) - 1u));
              pk[slot] = key;
              px[slot] = eff;
            }
            cnt += __popc(mine);
          }
          __syncwarp();
        };
        constexpr bool ROLLABLE = !HALF && PS > 1 && WH % PS == 0 && WH >= 2 * PS;
        if constexpr (ROLLABLE) {   // measured with both forms in one binary: c4s 183.8 (unrolled) vs 162.3 ms (rolled)
          tile_core_rolled<WH, WW, PS, DIST, FP, FL>(bp, rowpitch, ph - flh, pw - flw, qv, flush);
        } else {
          tile_core<WH, WW, PS, DIST, FP, FL>(bp, rowpitch, ph - flh, pw - flw, qv, flush);
        }
      };
      if (!HALF) {
        run_window(std::integral_constant<int, WS>(), std::integral_constant<int, WS>(), 0, 0);
      } else {
        constexpr int WE = (WS + 1) / 2, WO = WS / 2;
        run_window(std::integral_constant<int, WE>(), std::integral_constant<int, WE>(), 0, 0);
        run_window(std::integral_constant<int, WE>(), std::integral_constant<int, WO>(), 0, 1);
        run_window(std::integral_constant<int, WO>(), std::integral_constant<int, WE>(), 1, 0);
        run_window(std::integral_constant<int, WO>(), std::integral_constant<int, WO>(), 1, 1);
      }
      if (chl == 0) pcnt[ql] = cnt;
    }
    __syncthreads();
    if (tid == 0 && st + C.nbuf < W_t) issue(st + C.nbuf);
  }

  // ---- outputs: the sorted list -> dists / inds ----
  float* const inds_out = reinterpret_cast<float*>(P.inds);
  for (int pass = 0; pass < NPASS; pass++) {
    const int qlr = (pass * 8 + warp) * QW + sub;
    const int ql = (sub < QW && qlr < TILE_NQ) ? qlr : TILE_NQ;
    const int an = a_nls[ql];
    __syncwarp();
    tile_merge<GS>(poolk + ql * TILE_CAP, pooli + ql * TILE_CAP, pcnt + ql, pthr + ql, LK, chl);
    if (state[ql] != 0) continue;
    QueryCtx c;
    query_ctx(ql, c);
    const int64_t wq = (((int64_t)n * g.Tq + tl) * g.nH + c.nh) * g.nW + c.nw;
    const int64_t obase = wq * P.Kout;
    if (anchored && chl == 0) {
      float ind[3] = {0.f, 0.f, 0.f};
      if (an == 0) cand_inds<float>(P, c, group_params<false, false>(P, c, 0), 0, ind);
      P.dists[obase] = adist[ql];
      inds_out[obase * 3] = ind[0]; inds_out[obase * 3 + 1] = ind[1]; inds_out[obase * 3 + 2] = ind[2];
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
      const int e = s * GS + chl;
      if (e >= LK) continue;
      const unsigned mk = poolk[ql * TILE_CAP + e];
      const int mi = pooli[ql * TILE_CAP + e];
      if (mk == 0xffffffffu) continue;   // cannot happen: K <= number of candidates
      int gq = mi / (WS * WS), sl = mi - gq * (WS * WS);
      if (anchored && mi == an) { gq = 0; sl = 0; }   // the entry displaced from slot 0 by the anchor swap
      float ind[3];
      cand_inds<float>(P, c, group_params<false, false>(P, c, gq), sl, ind);
      const int64_t oi = obase + e + (anchored ? 1 : 0);
      const float kf_ = ord2f_t(mk);
      P.dists[oi] = (DIST == 0) ? -kf_ : kf_;
      inds_out[oi * 3] = ind[0]; inds_out[oi * 3 + 1] = ind[1]; inds_out[oi * 3 + 2] = ind[2];
    }
  }

  // ---- queries this tile could not stage: the generic per-query path, one warp each ----
  for (int ql = warp; ql < TILE_NQ; ql += TILE_THREADS / 32) {
    if (state[ql] != 1) continue;
    QueryCtx c;
    query_ctx(ql, c);
    const int64_t wq = (((int64_t)n * g.Tq + tl) * g.nH + c.nh) * g.nW + c.nw;
    direct_query<false, DIST, false>(P, wq, lane);
  }
}

template <int DIST, int PS, int WS, bool HALF, int FP>
int launch_tile_one(const FwdParams& P, const TileCfg& C, const CUtensorMap& map, cudaStream_t st) {
  auto kern = search_fwd_tile_kernel<DIST, PS, WS, HALF, FP>;
  static int configured = 0;   // per instantiation
  if (configured < C.smem_bytes) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C.smem_bytes) != cudaSuccess) return -1;
    configured = C.smem_bytes;
  }
  const int64_t blocks = (int64_t)C.tiles_h * C.tiles_w * P.g.Tq * P.g.B * P.g.HD;
  kern<<<(unsigned)blocks, TILE_THREADS, C.smem_bytes, st>>>(P, C, map);
  return 1;
}

// (ps, ws, stride1, Fp) menu compiled into each instantiation unit; everything else stays on the table kernel
template <int DIST>
int launch_tile_menu(const FwdParams& P, const TileCfg& C, const CUtensorMap& map, cudaStream_t st) {
  const int ps = P.o.ps, ws = P.o.ws, fp = P.g.Fp;
  const bool half = P.o.stride1 == 0.5f;
#define TILE_CASE(PS_, WS_, HALF_, FP_) \
  if (ps == PS_ && ws == WS_ && half == HALF_ && fp == FP_) return launch_tile_one<DIST, PS_, WS_, HALF_, FP_>(P, C, map, st);
  TILE_CASE(3, 5, true, 8)
  TILE_CASE(3, 5, false, 8)
  TILE_CASE(3, 3, false, 8)
  TILE_CASE(1, 9, false, 16)
  TILE_CASE(3, 9, false, 16)
  TILE_CASE(3, 5, false, 16)
  TILE_CASE(1, 5, false, 16)
  TILE_CASE(3, 5, true, 16)
  TILE_CASE(3, 3, false, 16)
#undef TILE_CASE
  return 0;
}

#endif  // __CUDACC__

}  // namespace stnls_fwd
