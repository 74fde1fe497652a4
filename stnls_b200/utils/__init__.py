from .config import extract_pairs, AttrDict  # noqa: F401
