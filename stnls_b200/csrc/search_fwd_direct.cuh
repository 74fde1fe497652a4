// Direct (table-free) evaluation of ONE query by one warp: candidate enumeration + patch distance straight from
// global memory + anchor_self + streaming top-K.  Complete reference semantics with no
// restriction on stride1 / patch size / flow range, so it is (a) the kernel of last resort (search_fwd.cu) and
// (b) the in-kernel fallback of the tile kernel for queries whose window leaves the staged key box.
// Reference: lib/csrc/search/non_local_search_{int,bilin2d}_kernel.cu, lib/csrc/search/refinement_*_kernel.cu,
// lib/csrc/nn/anchor_self_kernel.cu:24-116,547-650, lib/stnls/nn/topk.py:65-123.
#pragma once
#include "search_fwd_common.cuh"

namespace stnls_fwd {

template <bool INT, int DIST, bool REFINE>
__device__ __noinline__ void direct_query(const FwdParams& P, const int64_t wq, const int lane) {
  typedef typename CT<INT>::type C;
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;

  // -- decode query (shared_kernel.cu:76-98) --
  QueryCtx c;
  int64_t r = divmod(wq, g.nW, c.nw);
  r = divmod(r, g.nH, c.nh);
  r = divmod(r, g.Tq, c.tl);
  c.n = (int)r;
  c.b = (int)divmod(r, g.HD, c.hd);
  c.t = c.tl + g.tq0;
  c.ah = (c.nh * o.stride0) % g.kH;
  c.aw = (c.nw * o.stride0) % g.kW;
  if (INT) { c.rh = (c.nh * o.strideQ) % g.qH; c.rw = (c.nw * o.strideQ) % g.qW; }
  else { c.rh = c.ah; c.rw = c.aw; }
  c.t_max = time_tmax(c.t, g.T, o.wt);

  const float* qf = P.q + ((int64_t)c.n * g.Tq + c.tl) * g.qH * g.qW * g.Fp;
  const float* kbase = P.k + (int64_t)c.n * g.Tk * g.kH * g.kW * g.Fp;
  const int64_t kfs = (int64_t)g.kH * g.kW * g.Fp;
  const float invalid = (DIST == 0) ? -INFINITY : INFINITY;
  const bool anchored = o.anchor_self != 0;
  const bool keep_all = o.K <= 0;
  const int nchunks = (P.nG + 31) / 32;
  const int64_t obase = wq * P.Kout;
  C* inds_out = reinterpret_cast<C*>(P.inds);

  // -- pass 1 (NLS anchor): candidate whose offset is (0,0,0) / nearest to it (anchor_self_kernel.cu:62-89)
  int a_nls = 0;
  if (anchored && !REFINE) {
    float bkey = INFINITY;
    int bpos = 0x7fffffff;
    const C zero[3] = {(C)0, (C)0, (C)0};
    for (int ch = 0; ch < nchunks; ch++) {
      int gi = ch * 32 + lane;
      Group<C> mine = group_params<INT, REFINE>(P, c, min(gi, P.nG - 1));
      int pend = min(P.L, (ch + 1) * 32 * P.G);
      for (int base = ch * 32 * P.G; base < pend; base += 32) {
        int pos = min(base + lane, pend - 1);
        int gq = pos / P.G;
        Group<C> gp = shfl_group(mine, gq - ch * 32);
        C ind[3];
        cand_inds<C>(P, c, gp, pos - gq * P.G, ind);
        float key = anchor_key(ind, zero);
        if (base + lane < pend && (key < bkey || (key == bkey && pos < bpos)) && key < INFINITY) { bkey = key; bpos = pos; }
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      float ok = __shfl_xor_sync(FULL_MASK, bkey, off);
      int op = __shfl_xor_sync(FULL_MASK, bpos, off);
      if (ok < bkey || (ok == bkey && op < bpos)) { bkey = ok; bpos = op; }
    }
    a_nls = (bpos == 0x7fffffff) ? 0 : bpos;
  }

  // -- pass 2: distances + streaming top-K --
  const int Klist = keep_all ? 0 : (anchored ? P.Kout - 1 : P.Kout);
  TopK tk;
  tk.key = INFINITY;
  tk.pos = 0x7fffffff;
  for (int ch = 0; ch < nchunks; ch++) {
    int gi = ch * 32 + lane;
    Group<C> mine = group_params<INT, REFINE>(P, c, min(gi, P.nG - 1));
    int my_ag = 0;
    if (REFINE && anchored) my_ag = refine_group_anchor<C>(P, c, mine, min(gi, P.nG - 1));
    int pend = min(P.L, (ch + 1) * 32 * P.G);
    for (int base = ch * 32 * P.G; base < pend; base += 32) {
      bool active = base + lane < pend;
      int pos = min(base + lane, pend - 1);
      int gq = pos / P.G, s = pos - gq * P.G;
      Group<C> gp = shfl_group(mine, gq - ch * 32);
      int ag = __shfl_sync(FULL_MASK, my_ag, gq - ch * 32);
      float d = invalid;
      if (active && gp.ok) {
        C chh, cww;
        cand_center(P, gp, s, chh, cww);
        bool valid = resident(gp.t, g) && inb(chh, g.kH) && inb(cww, g.kW);
        if (valid) d = patch_dist<DIST>(P, qf, kbase + (int64_t)(gp.t - g.tk0) * kfs, c, chh, cww);
      }
      // position after the anchor swap (an involution inside the group)
      int eff = pos;
      if (anchored) {
        if (REFINE) { if (s == 0) eff = gq * P.G + ag; else if (s == ag) eff = gq * P.G; }
        else { if (pos == 0) eff = a_nls; else if (pos == a_nls) eff = 0; }
      }
      // the reference swap rewrites slot 0 with the anchor offset and then restores the displaced
      // entry, so nothing changes when the anchor already sits in slot 0 (anchor_self_kernel.cu:101-113)
      bool to_slot0 = anchored && (REFINE ? (eff % P.G == 0 && ag != 0) : (eff == 0 && a_nls != 0));
      if (keep_all || (anchored && eff == 0)) {
        if (active) {
          int64_t oi = obase + (keep_all ? eff : 0);
          C ind[3];
          if (to_slot0) {
            if (REFINE) refine_loc<C>(P, c, eff / P.G, ind);
            else ind[0] = ind[1] = ind[2] = (C)0;
          } else cand_inds<C>(P, c, gp, s, ind);
          P.dists[oi] = d;
          inds_out[oi * 3] = ind[0]; inds_out[oi * 3 + 1] = ind[1]; inds_out[oi * 3 + 2] = ind[2];
          if (P.kselect) P.kselect[oi] = gp.ok ? gq : 1;
        }
        if (keep_all) continue;
      }
      // stream into the list
      float key = (DIST == 0) ? -d : d;
      if (key != key) key = INFINITY;   // NaN distance (non-finite inputs): ranks last instead of never being inserted,
                                        // so every output slot is written (the reference returns a full sorted list)
      bool want = active && !(anchored && eff == 0);
      float wk = __shfl_sync(FULL_MASK, tk.key, max(Klist - 1, 0));
      int wp = __shfl_sync(FULL_MASK, tk.pos, max(Klist - 1, 0));
      want = want && Klist > 0 && ((key < wk) || (key == wk && eff < wp));
      unsigned todo = __ballot_sync(FULL_MASK, want);
      while (todo) {
        int src = __ffs(todo) - 1;
        todo &= todo - 1;
        float nk = __shfl_sync(FULL_MASK, key, src);
        int np = __shfl_sync(FULL_MASK, eff, src);
        topk_insert(tk, nk, np, Klist, lane);
      }
    }
  }
  if (keep_all) return;

  // -- write the list: recompute the offsets of the kept candidates --
  if (lane < Klist && tk.pos != 0x7fffffff) {
    int eff = tk.pos;
    int64_t oi = obase + (anchored ? lane + 1 : lane);
    int gq = eff / P.G, s = eff - gq * P.G;
    Group<C> gp = group_params<INT, REFINE>(P, c, gq);
    C ind[3];
    bool done = false;
    if (anchored) {
      if (REFINE) {
        int ag = refine_group_anchor<C>(P, c, gp, gq);
        if (s == 0 && ag != 0) { refine_loc<C>(P, c, gq, ind); done = true; }
        else if (s == ag) s = 0;
      } else {
        if (eff == a_nls) { gq = 0; s = 0; gp = group_params<INT, REFINE>(P, c, 0); }
      }
    }
    if (!done) cand_inds<C>(P, c, gp, s, ind);
    P.dists[oi] = (DIST == 0) ? -tk.key : tk.key;
    inds_out[oi * 3] = ind[0]; inds_out[oi * 3 + 1] = ind[1]; inds_out[oi * 3 + 2] = ind[2];
    if (P.kselect) P.kselect[oi] = gp.ok ? gq : 1;
  }
}

}  // namespace stnls_fwd
