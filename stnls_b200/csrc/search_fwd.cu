// Fused forward search: candidate enumeration + patch distance + anchor_self + top-K, one warp per
// query, nothing materialised per candidate.  Replaces:
//   lib/csrc/search/non_local_search_{int,bilin2d}_kernel.cu (forward kernels)
//   lib/csrc/search/refinement_{int,bilin2d}_kernel.cu       (forward kernels)
//   lib/csrc/nn/anchor_self_kernel.cu:24-116,547-650
//   lib/stnls/nn/topk.py:65-123,196-220 (argsort + gather)
// Distances accumulate in the reference's order (pi -> pj -> channel, FMA) so the int path is
// bit-identical to the reference kernels; ties in the top-K break to the lower candidate index.
#include "common.cuh"
#include <cstdlib>

#include "search_fwd_direct.cuh"
#include "search_fwd_tab.cuh"
#include "search_fwd_tile.cuh"
#include "search_fwd_refine_chan.cuh"

using namespace stnls_fwd;

namespace {

template <bool INT, int DIST, bool REFINE>
__global__ void __launch_bounds__(256) search_fwd_kernel(const __grid_constant__ FwdParams P) {
  const int64_t wq = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (wq >= P.nquery) return;
  direct_query<INT, DIST, REFINE>(P, wq, threadIdx.x & 31);
}


template <bool INT, int DIST, bool REFINE>
void launch_fwd(const FwdParams& P, cudaStream_t st) {
  const int wpb = 8;
  int64_t blocks = (P.nquery + wpb - 1) / wpb;
  search_fwd_kernel<INT, DIST, REFINE><<<(unsigned)blocks, wpb * 32, 0, st>>>(P);
}

// host: lattice description + smem carve of the table kernel; false when that path does not apply
// (no lattice for this stride1, lattice or key list too large for shared memory)
bool make_tabcfg(const FwdParams& P, TabCfg& C) {
  const stnls_search_opts& o = P.o;
  const int ps = o.ps, wsz = P.wsz;
  if (ps * ps > 32) return false;
  if ((int64_t)P.g.kH * P.g.kW * P.g.Fp >= (1LL << 31) || (int64_t)P.g.qH * P.g.qW * P.g.Fp >= (1LL << 31)) return false;
  int D = 0;
  if (o.itype_int) D = 1;
  else {
    for (int d = 1; d <= 4; d *= 2) {
      float x = o.stride1 * d;
      if (x >= 1.f && x == floorf(x) && x <= 64.f) { D = d; break; }
    }
  }
  if (D == 0) return false;
  C.n1 = o.itype_int ? (int)o.stride1 : (int)(o.stride1 * D);
  C.n2 = o.dilation * D;
  C.u = 1.f / D;
  C.D = D;
  if (C.n1 < 1 || C.n2 < 1) return false;
  C.M = (wsz - 1) * C.n1 + (ps - 1) * C.n2 + 1;
  if (C.M > 32) return false;
  C.rpat = 0;
  for (int p = 0; p < ps; p++) C.rpat |= 1u << (C.n2 * p);
  C.Q4 = P.g.Fp / 4;
  C.q4sh = -1;
  for (int sft = 0; sft < 8; sft++) if ((1 << sft) == C.Q4) C.q4sh = sft;
  C.ax_rounds = (C.M <= 16) ? 1 : 2;
  C.Mp = C.M | 1;
  // small windows (ws < 8) keep the row-major candidate order; a row pitch == ws (mod 8) then puts eight
  // consecutive candidates on eight different 16-byte bank groups.  Only when the padding is modest.
  if (wsz < 8 && (C.n1 & 1) && !getenv("STNLS_FWD_NOPAD")) {
    int mp = C.M;
    while ((mp & 7) != (wsz & 7)) mp++;
    if (mp <= C.M + 4) C.Mp = mp;
  }
  // raw mode (see TabCfg): needs the templated kernel variants that carry the code
  C.raw = 0; C.Mr = C.M + 1; C.Wr = wsz + 1; C.nrr = 0; C.scr_off = 0;
  C.inv_Mr = 1.f / C.Mr; C.inv_Wr = 1.f / C.Wr; C.inv_nPg = 1.f / (C.Wr * C.Wr); C.inv_nTr = 0.f;
  if (!o.itype_int && o.dist_type == 0 && D == 1 && C.n1 == 1 && C.n2 == 1 && wsz <= 60 &&
      ((ps == 1 && (C.Q4 == 2 || C.Q4 == 4)) || (ps == 3 && C.Q4 == 4)) && !getenv("STNLS_FWD_NO_RAW")) {
    C.raw = 1;
    if (ps == 1) { C.Mp = C.Mr; while ((C.Mp & 7) != (C.Wr & 7)) C.Mp++; }
    else C.Mp = C.Mr | 1;
    for (int w0 = 0; w0 < C.Wr && C.nrr < 16; w0 += 4) {
      int cnt = C.Wr - w0 < 4 ? C.Wr - w0 : 4;
      C.rruns[C.nrr++] = (unsigned)(w0 | (cnt << 8));
    }
    C.inv_nTr = 1.f / (C.Wr * C.nrr);
  }
  // run tiling: candidates wj, wj + rs, wj + 2 rs of one row share lattice columns when n1 divides n2
  C.tile = 0; C.rs = 1; C.nruns = 0; C.inv_nTg = 0.f;
  if (ps == 3 && (C.n2 % C.n1) == 0 && wsz <= 64 && !getenv("STNLS_FWD_NO_TILE")) {
    C.rs = C.n2 / C.n1;
    for (int rho = 0; rho < C.rs && rho < wsz; rho++) {
      int members = (wsz - 1 - rho) / C.rs + 1;
      for (int m0 = 0; m0 < members; m0 += 3) {
        if (C.nruns >= 32) break;
        int cnt = members - m0 < 3 ? members - m0 : 3;
        C.runs[C.nruns++] = (unsigned)((rho + m0 * C.rs) | (cnt << 8));
      }
    }
    int covered = 0;
    for (int i = 0; i < C.nruns; i++) covered += C.runs[i] >> 8;
    // enough runs per group to fill a warp's lanes, and the kernel variant that carries the tiled loop
    if (covered == wsz && (C.Mp & 1) && wsz * C.nruns >= 16 && C.Q4 == 4) { C.tile = 1; C.inv_nTg = 1.f / (wsz * C.nruns); }
  }
  int want = (C.Q4 == 2) ? 4 : (C.Q4 == 4) ? 2 : (C.Q4 >= 8 ? 1 : 0);
  C.PS4 = (C.raw ? C.Mr : C.M) * C.Mp;
  if (C.Q4 > 1) while ((C.PS4 & 7) != want) C.PS4++;
  C.GS4 = C.PS4 * C.Q4;
  // groups per batch: best lane utilisation of the candidate loop within a per-warp smem budget
  const char* env_nb = getenv("STNLS_FWD_NB");
  const char* env_bud = getenv("STNLS_FWD_BUDGET");
  const int budget = env_bud ? atoi(env_bud) : 6144;
  int best = 1;
  double best_eff = 0;
  for (int nb = 1; nb <= 8 && nb <= P.nG; nb++) {
    int bytes = nb * C.GS4 * 16;
    if (nb > 1 && bytes > budget) break;
    const int units = C.raw ? (ps == 1 ? C.Wr * C.Wr : C.Wr * C.nrr) : C.tile ? wsz * C.nruns : P.G;   // lanes of work per group
    int rounds = (nb * units + 31) / 32;
    double eff = (double)(nb * units) / (32.0 * rounds);
    if (eff > best_eff + 0.02) { best_eff = eff; best = nb; }
  }
  C.nb = env_nb ? atoi(env_nb) : best;
  if (C.nb < 1) C.nb = 1;
  if (C.nb > 32) C.nb = 32;
  if (C.nb > P.nG) C.nb = P.nG;
  C.dec_f4 = (C.M * C.M * 2 + 15) / 16;
  C.q_off = C.nb * C.GS4;
  C.ax_off = C.q_off + ps * ps * C.Q4;
  C.msk_off = C.ax_off + C.nb * 2 * C.M;
  C.key_off = C.msk_off + (3 * C.nb + 3) / 4;
  const int key_f4 = (o.K > 0) ? (P.L + 3) / 4 : 0;
  if (key_f4 * 16 > 16 * 1024) return false;
  C.scr_off = C.key_off + key_f4;
  C.warp_f4 = C.scr_off + (C.raw ? (C.nb * C.Wr * C.Wr + 3) / 4 : 0);
  const char* env_w = getenv("STNLS_FWD_WARPS");
  C.wpb = env_w ? atoi(env_w) : 8;
  if (C.wpb < 1 || C.wpb > 8) C.wpb = 8;
  // bank-conflict-free candidate order needs an odd lattice step per candidate (and the odd row pitch Mp)
  C.W8 = ((C.n1 & 1) && !getenv("STNLS_FWD_ROWMAJOR")) ? (wsz & ~7) : 0;
  if (wsz < 8 || !(C.Mp & 1)) C.W8 = 0;
  C.nA = wsz * C.W8;
  C.inv_W8 = C.W8 ? 1.f / C.W8 : 0.f;
  C.tA8 = wsz & ~7; C.rA8 = C.Wr & ~7;
  C.inv_tA8 = C.tA8 ? 1.f / C.tA8 : 0.f; C.inv_trem = (wsz - C.tA8) ? 1.f / (wsz - C.tA8) : 0.f;
  C.inv_rA8 = C.rA8 ? 1.f / C.rA8 : 0.f; C.inv_rrem = (C.Wr - C.rA8) ? 1.f / (C.Wr - C.rA8) : 0.f;
  C.inv_G = 1.f / P.G; C.inv_wsz = 1.f / wsz; C.inv_Q4 = 1.f / C.Q4; C.inv_M = 1.f / C.M;
  return (size_t)C.warp_f4 * 16 <= 24 * 1024;
}

}  // namespace

extern "C" int stnls_search_forward(const float* q_cl, const float* k_cl, const void* flows, float* dists,
                                    void* inds, int* kselect, uint8_t* reflect, const stnls_geom* g,
                                    const stnls_search_opts* o, stnls_stream_t stream) {
  STNLS_REQUIRE(q_cl && k_cl && flows && dists && inds && g && o, "search_forward: null argument");
  STNLS_REQUIRE(g->Fp % 4 == 0 && g->Fp >= g->F, "search_forward: Fp must be a multiple of 4 and >= F");
  STNLS_REQUIRE(o->dist_type == 0 || o->dist_type == 1, "Uknown distance type. Must be 0 (product) or 1 (l2)");
  STNLS_REQUIRE(o->ps > 0, "search_forward: patch size is invalid [%d]", o->ps);
  const bool refine = o->wr > 0;
  FwdParams P;
  P.q = q_cl; P.k = k_cl; P.flows = flows; P.dists = dists; P.inds = inds; P.kselect = kselect; P.reflect = reflect;
  P.g = *g; P.o = *o;
  if (refine) {
    STNLS_REQUIRE(o->Ks > 0, "search_forward: refine needs Ks > 0");
    P.G = o->wr * o->wr; P.nG = o->Ks; P.wsz = o->wr; P.W_t = 0;
  } else {
    STNLS_REQUIRE(o->ws > 0 && o->wt >= 0, "search_forward: need ws > 0 and wt >= 0");
    P.W_t = 2 * o->wt + 1;
    STNLS_REQUIRE(g->T >= P.W_t, "Num Frames [%d] must be >= Temporal Window [%d]", g->T, P.W_t);
    STNLS_REQUIRE(o->st_offset == 0 || o->st_offset == 1, "search_forward: flows.size(3) must be W_t or W_t-1");
    P.G = o->ws * o->ws; P.nG = P.W_t; P.wsz = o->ws;
  }
  if (int rc = stnls_check_geom(g, refine ? -1 : o->wt, "search_forward")) return rc;
  P.L = P.G * P.nG;
  P.Kout = (o->K > 0) ? (o->K < P.L ? o->K : P.L) : P.L;
  STNLS_REQUIRE(o->K <= 0 || P.Kout <= 32 + (o->anchor_self ? 1 : 0),
                "search_forward: fused top-K supports K <= 32 (got %d); use K<=0 and sort outside", o->K);
  P.nquery = (int64_t)g->B * g->HD * g->Tq * g->nH * g->nW;
  if (P.nquery == 0) return STNLS_OK;
  STNLS_REQUIRE((P.nquery + 7) / 8 < 2147483647LL, "search_forward: too many queries for one launch");
  cudaStream_t st = (cudaStream_t)stream;
  const int key = (o->itype_int ? 4 : 0) | (o->dist_type ? 2 : 0) | (refine ? 1 : 0);
  {
    // float NonLocalSearch: the tile kernel (TMA-staged key box shared by an 8x8 tile of queries)
    const int rc = launch_tile(P, st);
    STNLS_REQUIRE(rc >= 0, "search_forward: cannot reserve shared memory for the tile kernel");
    if (rc == 1) {
      STNLS_CHECK_LAUNCH("search_forward(tile)");
      return STNLS_OK;
    }
  }
  if (refine) {
    // float RefineSearch: the channel-parallel kernel (lane = channel, raw pixels read once per input offset)
    const int rc = launch_refine_chan(P, st);
    if (rc == 1) {
      STNLS_CHECK_LAUNCH("search_forward(refine, channel-parallel)");
      return STNLS_OK;
    }
  }
  TabCfg C;
  const char* env_v1 = getenv("STNLS_FWD_V1");
  if (!(env_v1 && env_v1[0] == '1') && make_tabcfg(P, C)) {
    int rc = 0;
    if (o->itype_int) rc = refine ? launch_tab_ref_int(P, C, st) : launch_tab_nls_int(P, C, st);
    else rc = refine ? launch_tab_ref_float(P, C, st) : launch_tab_nls_float(P, C, st);
    STNLS_REQUIRE(rc == 0, "search_forward: cannot reserve shared memory for the table kernel");
    STNLS_CHECK_LAUNCH("search_forward(table)");
    return STNLS_OK;
  }
  switch (key) {
    case 0: launch_fwd<false, 0, false>(P, st); break;
    case 1: launch_fwd<false, 0, true>(P, st); break;
    case 2: launch_fwd<false, 1, false>(P, st); break;
    case 3: launch_fwd<false, 1, true>(P, st); break;
    case 4: launch_fwd<true, 0, false>(P, st); break;
    case 5: launch_fwd<true, 0, true>(P, st); break;
    case 6: launch_fwd<true, 1, false>(P, st); break;
    case 7: launch_fwd<true, 1, true>(P, st); break;
  }
  STNLS_CHECK_LAUNCH("search_forward");
  return STNLS_OK;
}
