// Lattice-table forward search: host-visible configuration + launcher declarations.
// The kernel itself is in search_fwd_tab_kernel.cuh and is instantiated in four translation units
// (search_fwd_tab_{nls,ref}_{int,float}.cu) so the template variants compile in parallel.
#pragma once
#include "search_fwd_common.cuh"

namespace stnls_fwd {

struct TabCfg {
  int n1, n2;     // lattice steps of one candidate step / one patch-pixel step
  float u;        // lattice unit
  int M, Mp;      // lattice points per axis (<= 32), row pitch of T (float4 units)
  int PS4, GS4;   // plane (channel quad) stride and group stride of T, float4 units
  int nb;         // groups per batch
  int Q4, q4sh;   // channel quads; log2(Q4) or -1
  int ax_rounds;  // 1: both axes of a group in one 32-lane round (M <= 16), 2: one round per axis
  unsigned rpat;  // bits n2*p, p < ps: lattice rows/cols one patch touches relative to its first
  int dec_f4;     // float4s of the per-CTA item decode table
  int q_off, ax_off, msk_off, key_off, warp_f4;  // per-warp smem carve, float4 units
  int wpb;        // warps per block
  float inv_G, inv_wsz, inv_Q4, inv_M;
  int D;          // lattice denominator (u = 1/D)
  int W8, nA;     // candidate enumeration: columns [0,W8) row-wise in chunks of 8 (nA = wsz*W8 lanes), rest column-wise
  float inv_W8;
  // run tiling of the candidate phase (ps == 3, n1 | n2): a lane owns up to 3 candidates wj = w0 + rs*j of a row
  int tile, rs, nruns;
  float inv_nTg;
  unsigned runs[32];  // w0 | count << 8
  // "raw" mode (float, prod, unit lattice: stride1 == 1, dilation == 1; ps 1 or 3): when a whole window is
  // interior the bilinear interpolation commutes with the patch contraction, so the kernel correlates the
  // query patch with the RAW key pixels on the (ws+1)^2 integer grid and blends the four neighbouring
  // correlations per candidate -- no interpolated lattice table is built at all
  int raw, Mr, Wr, nrr, scr_off;   // Mr = M+1 raw rows/cols, Wr = ws+1 correlation grid, nrr runs per grid row
  float inv_Mr, inv_Wr, inv_nPg, inv_nTr;
  unsigned rruns[16];              // runs of up to 4 grid columns: w0 | count << 8
  // lane -> (run, row) of the tiled loops: rows [0,A8) of every run first (A8 = rows & ~7, so an 8-lane LDS.128
  // wavefront stays inside one run and reads 8 different rows), then the leftover rows
  int tA8, rA8;
  float inv_tA8, inv_trem, inv_rA8, inv_rrem;
};

// one per translation unit; each picks dist_type / ps / channel-quad template variants at run time
int launch_tab_nls_int(const FwdParams& P, const TabCfg& C, cudaStream_t st);
int launch_tab_nls_float(const FwdParams& P, const TabCfg& C, cudaStream_t st);
int launch_tab_ref_int(const FwdParams& P, const TabCfg& C, cudaStream_t st);
int launch_tab_ref_float(const FwdParams& P, const TabCfg& C, cudaStream_t st);

}  // namespace stnls_fwd
