"""stnls_b200 -- B200-native (sm_100a) shifted non-local search + weighted patch aggregation.

Drop-in for the hot path of gauenk/stnls:
    stnls_b200.search.NonLocalSearch / RefineSearch   (lib/stnls/search/{non_local_search,refinement}.py)
    stnls_b200.agg.WeightedPatchSum / NonLocalGatherAdd (lib/stnls/agg/gather_add.py)
backed by hand-written CUDA kernels behind the C ABI in include/stnls_b200.h.  No CPU fallback.
"""
from . import agg, nn, normz, parallel, search, utils  # noqa: F401
from . import ops  # noqa: F401

__version__ = "0.1.0"
