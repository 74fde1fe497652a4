// instantiation unit of the tile forward kernel: dist_type = l2 (see search_fwd_tile.cuh)
#include <type_traits>
#include "search_fwd_tile.cuh"

namespace stnls_fwd {
int launch_tile_l2(const FwdParams& P, const TileCfg& C, const CUtensorMap& map, cudaStream_t st) {
  return launch_tile_menu<1>(P, C, map, st);
}
}  // namespace stnls_fwd
