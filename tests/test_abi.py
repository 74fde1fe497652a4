"""CPU: the C-ABI library loads and exports every symbol include/stnls_b200.h declares; argument
validation works without a GPU (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "stnls_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(stnls_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported():
    from stnls_b200 import _C
    lib = _C.lib()
    syms = declared_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/stnls_b200.h but not exported"
    assert sorted(_C.EXPORTS) == syms
    assert lib.stnls_version() == 100


def test_overlap_counts_host_helper():
    from stnls_b200 import _C
    lib = _C.lib()
    buf = (ctypes.c_int * 8)()
    mn = lib.stnls_overlap_counts(buf, 8, 8, 3, 1, 1, -1)
    assert list(buf) == [2, 3, 3, 3, 3, 3, 3, 2] and mn == 2
    mn = lib.stnls_overlap_counts(buf, 8, 4, 1, 2, 1, 0)  # stride 2, ps 1: holes -> min 0
    assert list(buf) == [1, 0, 1, 0, 1, 0, 1, 0] and mn == 0


def test_argument_validation_without_gpu():
    from stnls_b200 import _C
    lib = _C.lib()
    g = _C.Geom(1, 1, 3, 0, 3, 0, 3, 3, 3, 8, 8, 8, 8, 8, 8)  # Fp=3 is not a multiple of 4
    o = _C.SearchOpts(3, 1, 0, 0, 1, 1, 1, 1, 1, 1.0, 1, 1, 0, 0, 0, 0, 1, 0, 1, 0)
    one = ctypes.c_void_p(16)
    rc = lib.stnls_search_forward(one, one, one, one, one, None, None, ctypes.byref(g), ctypes.byref(o), None)
    assert rc == -1 and b"multiple of 4" in lib.stnls_last_error()
    with pytest.raises(ValueError):
        _C.check(rc)
    g.Fp = 4
    o.dist_type = 7
    rc = lib.stnls_search_forward(one, one, one, one, one, None, None, ctypes.byref(g), ctypes.byref(o), None)
    assert rc == -1 and b"distance type" in lib.stnls_last_error()
    o.dist_type = 0
    g.T = 2  # T < W_t
    rc = lib.stnls_search_forward(one, one, one, one, one, None, None, ctypes.byref(g), ctypes.byref(o), None)
    assert rc == -1 and b"Temporal Window" in lib.stnls_last_error()


def test_front_end_refuses_cpu_tensors():
    import torch as th
    import stnls_b200
    x = th.randn(1, 3, 4, 8, 8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        stnls_b200.search.NonLocalSearch(3, 1, 1, 2)(x, x)
    with pytest.raises(RuntimeError, match="no CPU path"):
        stnls_b200.agg.WeightedPatchSum(3, 1)(x, th.rand(1, 1, 3, 8, 8, 2), th.zeros(1, 1, 3, 8, 8, 2, 3))


def test_config_factories():
    import stnls_b200
    s = stnls_b200.search.init({"ws": 5, "wt": 2, "ps": 3, "k": 10, "nheads": 2, "stride1": 0.5, "itype": "float",
                                "self_action": "anchor"})
    assert (s.ws, s.wt, s.ps, s.k, s.nheads, s.stride1, s.dist_type) == (5, 2, 3, 10, 2, 0.5, "l2")
    r = stnls_b200.search.init({"search_name": "refine", "ws": 3, "wt": 1, "wr": 3, "k": 9, "kr": -1, "ps": 3})
    assert isinstance(r, stnls_b200.search.RefineSearch) and r.wr == 3
    a = stnls_b200.agg.init({"agg_name": "wpsum", "ps": 3, "strideIn": 2, "strideOut": 2, "itype": "int"})
    assert isinstance(a, stnls_b200.agg.NonLocalGatherAdd) and a.strideIn == 2 and a.itype == "int"
    cfg = stnls_b200.search.non_local_search.extract_config({"ws": 7, "foo": 1})
    assert dict(cfg) == {"ws": 7}
    w = stnls_b200.agg.WeightedPatchSum(ps=3, stride0=1, itype="float")
    assert w.strideIn == 1 and w.strideOut == 1
