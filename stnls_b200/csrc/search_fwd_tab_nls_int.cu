// instantiation unit of the lattice-table forward kernel: INT=true, REFINE=false (see search_fwd_tab_kernel.cuh)
#include "search_fwd_tab_kernel.cuh"

namespace stnls_fwd {
int launch_tab_nls_int(const FwdParams& P, const TabCfg& C, cudaStream_t st) { return launch_tab_all<true, false>(P, C, st); }
}  // namespace stnls_fwd
