// instantiation unit of the tile forward kernel: dist_type = prod (see search_fwd_tile.cuh)
#include <type_traits>
#include "search_fwd_tile.cuh"

namespace stnls_fwd {
int launch_tile_prod(const FwdParams& P, const TileCfg& C, const CUtensorMap& map, cudaStream_t st) {
  return launch_tile_menu<0>(P, C, map, st);
}
}  // namespace stnls_fwd
