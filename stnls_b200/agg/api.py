"""Dict-config factory: agg = stnls_b200.agg.init(cfg)   (lib/stnls/agg/api.py:20-48).
"wpsum" (the README name) resolves to gather_add, its successor in the reference."""
from ..utils import extract_pairs
from . import gather, gather_add

MENU = {"wpsum": gather_add, "gather_add": gather_add, "gather": gather, "nlstack": gather}


def _module(name):
    if name not in MENU:
        raise NotImplementedError(f"stnls_b200: aggregator [{name}] is outside the hot path (SURVEY.md 2.1 row 3b)")
    return MENU[name]


def extract_config(_cfg, restrict=True):
    name = extract_pairs(_cfg, {"agg_name": "wpsum"}, restrict=False)["agg_name"]
    cfg = _module(name).extract_config(_cfg)
    cfg.agg_name = name
    return cfg


def init(cfg):
    name = extract_pairs(cfg, {"agg_name": "wpsum"}, restrict=False)["agg_name"]
    return _module(name).init(cfg)
