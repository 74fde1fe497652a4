"""stnls_b200.agg -- mirrors lib/stnls/agg/__init__.py for the weighted patch sum."""
from . import gather, gather_add  # noqa: F401
from .api import extract_config, init  # noqa: F401
from .gather_add import NonLocalGatherAdd, NonLocalGatherAddFunction, WeightedPatchSum  # noqa: F401
from .gather import NonLocalGather, non_local_gather  # noqa: F401
