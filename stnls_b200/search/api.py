"""Dict-config factory: search = stnls_b200.search.init(cfg)   (lib/stnls/search/api.py:27-60)."""
import copy

from ..utils import extract_pairs
from . import non_local_search, refinement

MENU = {"exact": non_local_search, "nls": non_local_search, "nl": non_local_search,
        "non_local_search": non_local_search, "refine": refinement, "refinement": refinement}


def _module(name):
    if name not in MENU:
        raise NotImplementedError(f"stnls_b200: search [{name}] is outside the hot path (SURVEY.md 2.1 row 1b)")
    return MENU[name]


def extract_config(_cfg, restrict=True):
    _cfg = copy.deepcopy(dict(_cfg))
    name = extract_pairs(_cfg, {"search_name": "nls"}, restrict=False)["search_name"]
    cfg = _module(name).extract_config(_cfg)
    cfg.search_name = name
    return cfg


def init(cfg):
    cfg = copy.deepcopy(dict(cfg))
    name = extract_pairs(cfg, {"search_name": "nls"}, restrict=False)["search_name"]
    return _module(name).init(cfg)
