"""stnls_b200.normz -- the normalise step between search and aggregation.
Mirrors lib/stnls/normz/{__init__,api,softmax}.py."""
from .softmax import SoftmaxNormalize, extract_config, init, softmax_scaled  # noqa: F401
