"""stnls_b200.nn -- the callers either side of the hot path: search_flow (N1), NonLocalAttention (N4)."""
from . import search_flow as _search_flow_mod  # noqa: F401
from .search_flow import run as search_flow  # noqa: F401
from .search_flow import search_flow_th  # noqa: F401
from . import non_local_attn  # noqa: F401
from .non_local_attn import NonLocalAttention  # noqa: F401
