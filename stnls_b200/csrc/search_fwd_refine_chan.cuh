// RefineSearch forward, float, channel-parallel (round 2).
//
// A refinement evaluates, for every query, wr x wr candidates around each of its Ks input offsets; the offsets of one
// query point into different frames and places, so there is no key box a tile of queries could share (the NLS tile
// kernel's TMA staging does not apply).  What does carry over from the tile kernel is the lane = CHANNEL core
// (tile_core): per (query, input offset) a lane walks the (wr+ps)^2 raw key pixels of its channel ONCE -- straight
// from global memory, coalesced over the channels of a pixel -- interpolates each of the (wr+ps-1)^2 lattice points
// once and feeds it to the candidates that cover it.  Channel sums go through the per-warp transpose scratch, the K best
// are kept by the append / compact pool of the tile kernel (tile_merge).  The table kernel instead builds an
// interpolated table per offset in shared memory with a whole warp per query.
// Fast path per query: every input offset legal, its window + patch + bilinear taps inside the frame (no reflection),
// the frame resident, the query patch inside the frame.  Any other query is evaluated by direct_query<REFINE>
// (search_fwd_direct.cuh) -- complete semantics (reflection, invalid offsets, window shifts).
// Scope: itype float, stride1 = 1, dilation 1, wr 1|3, ps 1|3, no self_action, K <= 32, 8/16/32 channels per head.
// Reference: lib/csrc/search/refinement_bilin2d_kernel.cu:18-185 (+ lib/stnls/search/impl/refinement.py:56-98).
#pragma once
#include "search_fwd_tile.cuh"

namespace stnls_fwd {

// 1 launched, 0 not applicable, -1 CUDA error
int launch_refine_chan(const FwdParams& P, cudaStream_t st);

#ifdef __CUDACC__

template <int DIST, int PS, int WR, int FP>
__global__ void __launch_bounds__(256, 3) refine_chan_kernel(const __grid_constant__ FwdParams P) {
  constexpr int QW = 32 / FP;            // queries per warp
  constexpr int NQ = 8 * QW;             // queries per CTA
  constexpr int NS = 32 / FP;            // list entries per lane
  constexpr int FL = tile_flush_size(WR, WR, FP);
  constexpr int PITCH = FP + 4;
  __shared__ __align__(16) float scr_all[8 * QW * FL * PITCH];
  __shared__ unsigned poolk[NQ * TILE_CAP];
  __shared__ int pooli[NQ * TILE_CAP];
  __shared__ int pcnt[NQ];
  __shared__ unsigned pthr[NQ];
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / FP, chl = lane % FP;
  const unsigned submask = ((FP >= 32) ? 0xffffffffu : ((1u << (FP & 31)) - 1u)) << (sub * FP);
  const int ql = warp * QW + sub;                               // query slot inside the CTA
  const int64_t wq_raw = (int64_t)blockIdx.x * NQ + ql;
  const bool exists = wq_raw < P.nquery;
  const int64_t wq = exists ? wq_raw : P.nquery - 1;           // lanes without a query idle along on the last one
  float* const scr = scr_all + (warp * QW + sub) * FL * PITCH;
  unsigned* const pk = poolk + ql * TILE_CAP;
  int* const px = pooli + ql * TILE_CAP;
  const int po = o.patch_offset, LK = P.Kout;

  QueryCtx c;
  {
    int64_t r = divmod(wq, g.nW, c.nw);
    r = divmod(r, g.nH, c.nh);
    r = divmod(r, g.Tq, c.tl);
    c.n = (int)r;
    c.b = (int)divmod(r, g.HD, c.hd);
    c.t = c.tl + g.tq0;
    c.ah = (c.nh * o.stride0) % g.kH;
    c.aw = (c.nw * o.stride0) % g.kW;
    c.rh = c.ah; c.rw = c.aw;
    c.t_max = time_tmax(c.t, g.T, o.wt);
  }
  if (chl == 0) { pcnt[ql] = 0; pthr[ql] = 0xffffffffu; }
  // the query patch must lie inside the frame (no reflected / clipped query pixels on this path)
  bool fast = exists && c.rh + po >= 0 && c.rh + po + PS - 1 <= g.qH - 1 && c.rw + po >= 0 && c.rw + po + PS - 1 <= g.qW - 1;
  float qv[PS * PS];
  {
    const float* qf = P.q + ((int64_t)c.n * g.Tq + c.tl) * g.qH * g.qW * FP + chl;
#pragma unroll
    for (int pi = 0; pi < PS; pi++)
#pragma unroll
      for (int pj = 0; pj < PS; pj++)
        qv[pi * PS + pj] = fast ? __ldg(qf + ((int64_t)(c.rh + pi + po) * g.qW + (c.rw + pj + po)) * FP) : 0.f;
  }
  // window descriptions of the Ks input offsets: lane `chl` computes offsets chl, chl + FP, ... (this also records
  // `reflect`, refinement_bilin2d_kernel.cu:121-122); a query leaves the fast path when any offset is illegal, its
  // frame is not resident, or its footprint touches the frame border
  const int Ks = o.Ks;
  constexpr int MAXG = 2;   // offsets per lane held in registers: Ks <= MAXG * FP
  Group<float> mine[MAXG];
#pragma unroll
  for (int j = 0; j < MAXG; j++) {
    const int gi = chl + j * FP;
    mine[j] = group_params<false, true>(P, c, min(gi, Ks - 1));
    if (gi < Ks) {
      const Group<float>& gp = mine[j];
      const float bh = fmaf(-o.stride1, gp.offh, gp.ah) + (float)po, bw = fmaf(-o.stride1, gp.offw, gp.aw) + (float)po;
      const bool ok = gp.ok && resident(gp.t, g) && bh >= 0.f && bw >= 0.f &&
                      floorf(bh) + (float)(WR + PS) <= (float)(g.kH - 1) && floorf(bw) + (float)(WR + PS) <= (float)(g.kW - 1) &&
                      gp.offh == (float)((WR - 1) / 2) && gp.offw == (float)((WR - 1) / 2);
      fast = fast && ok;
    }
  }
  fast = (__ballot_sync(FULL_MASK, !fast) & submask) == 0;   // the whole sub-group agrees
  __syncwarp();

  int cnt = 0;
  unsigned thr = 0xffffffffu;
  const int64_t kfs = (int64_t)g.kH * g.kW * FP;
  const float* const kbase = P.k + (int64_t)c.n * g.Tk * kfs + chl;
  const int rowpitch = g.kW * FP;
  for (int gi = 0; gi < Ks; gi++) {
    // broadcast offset gi's window from the lane that holds it
    const int src = sub * FP + (gi % FP), slot = gi / FP;
    float ah = 0.f, aw = 0.f;
    int gt = 0;
#pragma unroll
    for (int j = 0; j < MAXG; j++)
      if (j == slot) { ah = mine[j].ah; aw = mine[j].aw; gt = mine[j].t; }
    ah = __shfl_sync(FULL_MASK, ah, src);
    aw = __shfl_sync(FULL_MASK, aw, src);
    gt = __shfl_sync(FULL_MASK, gt, src);
    const float half = (float)((WR - 1) / 2);
    const float ph = fast ? fmaf(-o.stride1, half, ah) + (float)po : 0.f, pw = fast ? fmaf(-o.stride1, half, aw) + (float)po : 0.f;
    const float flh = floorf(ph), flw = floorf(pw);
    const float* bp = kbase + (int64_t)(fast ? gt - g.tk0 : 0) * kfs + ((int64_t)flh * g.kW + (int64_t)flw) * FP;
    auto flush = [&](const int lin0, const int nv, const float* vals) {
#pragma unroll
      for (int v = 0; v < FL; v++)
        if (v < nv) scr[v * PITCH + chl] = vals[v];
      __syncwarp();
#pragma unroll
      for (int m = 0; m < (FL + FP - 1) / FP; m++) {
        const int v = chl + m * FP;
        unsigned key = 0xffffffffu;
        int eff = 0;
        if (v < nv) {
          const float4* sp = reinterpret_cast<const float4*>(scr + v * PITCH);
          float d = 0.f;
#pragma unroll
          for (int c4 = 0; c4 < FP / 4; c4++) {
            const float4 x4 = sp[c4];
            d += (x4.x + x4.y) + (x4.z + x4.w);
          }
          eff = gi * WR * WR + lin0 + v;
          key = min(f2ord_t(((DIST == 0) ? -d : d) + 0.f), 0xfffffffeu);
        }
        if (__any_sync(FULL_MASK, cnt + (FL < FP ? FL : FP) > TILE_CAP)) {   // room for what one round can append
          if (chl == 0) pcnt[ql] = cnt;
          __syncwarp();
          tile_merge<FP>(pk, px, pcnt + ql, pthr + ql, LK, chl);
          cnt = pcnt[ql];
          thr = pthr[ql];
        }
        const bool pass_ = fast && key <= thr && key != 0xffffffffu;
        const unsigned mine_m = __ballot_sync(FULL_MASK, pass_) & submask;
        if (pass_) {
          const int slot2 = cnt + __popc(mine_m & ((1u << lane) - 1u));
          pk[slot2] = key;
          px[slot2] = eff;
        }
        cnt += __popc(mine_m);
      }
      __syncwarp();
    };
    tile_core<WR, WR, PS, DIST, FP, FL>(bp, rowpitch, ph - flh, pw - flw, qv, flush);
  }
  if (chl == 0) pcnt[ql] = cnt;
  __syncwarp();
  tile_merge<FP>(pk, px, pcnt + ql, pthr + ql, LK, chl);

  // ---- outputs ----
  if (fast) {
    const int64_t obase = wq * P.Kout;
    float* const inds_out = reinterpret_cast<float*>(P.inds);
#pragma unroll
    for (int s = 0; s < NS; s++) {
      const int e = s * FP + chl;
      if (e >= LK) continue;
      const unsigned mk = pk[e];
      const int mi = px[e];
      if (mk == 0xffffffffu) continue;
      const int gq = mi / (WR * WR), sl = mi - gq * (WR * WR);
      float ind[3];
      cand_inds<float>(P, c, group_params<false, true>(P, c, gq), sl, ind);
      const int64_t oi = obase + e;
      const float kf_ = ord2f_t(mk);
      P.dists[oi] = (DIST == 0) ? -kf_ : kf_;
      inds_out[oi * 3] = ind[0]; inds_out[oi * 3 + 1] = ind[1]; inds_out[oi * 3 + 2] = ind[2];
      if (P.kselect) P.kselect[oi] = gq;
    }
  }
  // ---- the other queries: generic path, one warp per query ----
  for (int s2 = 0; s2 < QW; s2++) {
    const bool slow = __shfl_sync(FULL_MASK, (int)(!fast && exists), s2 * FP) != 0;
    if (!slow) continue;
    const int64_t wq2 = (int64_t)blockIdx.x * NQ + warp * QW + s2;
    direct_query<false, DIST, true>(P, wq2, lane);
  }
}

template <int DIST, int PS, int WR, int FP>
int launch_refine_chan_one(const FwdParams& P, cudaStream_t st) {
  constexpr int NQ = 8 * (32 / FP);
  const int64_t blocks = (P.nquery + NQ - 1) / NQ;
  if (blocks >= 2147483647LL) return 0;
  refine_chan_kernel<DIST, PS, WR, FP><<<(unsigned)blocks, 256, 0, st>>>(P);
  return 1;
}

template <int DIST>
int launch_refine_chan_menu(const FwdParams& P, cudaStream_t st) {
  const int ps = P.o.ps, wr = P.o.wr, fp = P.g.Fp;
#define RC_CASE(PS_, WR_, FP_) \
  if (ps == PS_ && wr == WR_ && fp == FP_) return launch_refine_chan_one<DIST, PS_, WR_, FP_>(P, st);
  RC_CASE(3, 3, 32)
  RC_CASE(3, 3, 16)
  RC_CASE(3, 3, 8)
  RC_CASE(1, 3, 16)
  RC_CASE(3, 1, 16)
  RC_CASE(3, 1, 8)
#undef RC_CASE
  return 0;
}

#endif  // __CUDACC__

}  // namespace stnls_fwd
