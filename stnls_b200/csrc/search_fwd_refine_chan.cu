// instantiation unit + applicability test of the channel-parallel RefineSearch forward (search_fwd_refine_chan.cuh)
#include <cstdlib>
#include <type_traits>
#include "search_fwd_refine_chan.cuh"

namespace stnls_fwd {

int launch_refine_chan(const FwdParams& P, cudaStream_t st) {
  const stnls_geom& g = P.g;
  const stnls_search_opts& o = P.o;
  const char* env = getenv("STNLS_REFINE_CHAN");
  if (env && env[0] == '0') return 0;
  if (o.itype_int || o.wr <= 0 || o.anchor_self || o.K <= 0 || P.Kout > 32) return 0;
  if (o.dilation != 1 || !o.reflect_bounds || o.stride1 != 1.0f || o.off_Hq != 0 || o.off_Wq != 0) return 0;
  if (g.qH != g.kH || g.qW != g.kW || o.strideQ != o.stride0) return 0;
  if (!(g.Fp == 8 || g.Fp == 16 || g.Fp == 32) || o.Ks > 2 * g.Fp) return 0;
  if (g.kH < 16 || g.kW < 16) return 0;
  if ((int64_t)g.kH * g.kW * g.Fp >= (1LL << 31)) return 0;
  return o.dist_type == 0 ? launch_refine_chan_menu<0>(P, st) : launch_refine_chan_menu<1>(P, st);
}

}  // namespace stnls_fwd
