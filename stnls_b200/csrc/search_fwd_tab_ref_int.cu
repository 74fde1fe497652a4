// instantiation unit of the lattice-table forward kernel: INT=true, REFINE=true (see search_fwd_tab_kernel.cuh)
#include "search_fwd_tab_kernel.cuh"

namespace stnls_fwd {
int launch_tab_ref_int(const FwdParams& P, const TabCfg& C, cudaStream_t st) { return launch_tab_all<true, true>(P, C, st); }
}  // namespace stnls_fwd
