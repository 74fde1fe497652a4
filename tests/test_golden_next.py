"""CPU: replay the golden vectors of the next-row components (search_flow, NonLocalGather) -- outputs of the
REFERENCE's own kernels on a B200 (tests/golden/make_golden_next.py) -- through the CPU oracle."""
import glob
import os

import numpy as np
import pytest
import torch as th

from util import assert_close

HERE = os.path.dirname(os.path.abspath(__file__))
FLOW = sorted(glob.glob(os.path.join(HERE, "golden", "next_search_flow_*.npz")))
GATHER = sorted(glob.glob(os.path.join(HERE, "golden", "next_gather_*.npz")))


def load(path):
    z = np.load(path)
    return {k: th.from_numpy(np.asarray(z[k])) for k in z.files}


def test_next_golden_files_present():
    assert len(FLOW) == 2 and len(GATHER) == 2


@pytest.mark.parametrize("path", FLOW, ids=[os.path.basename(f)[:-4] for f in FLOW])
def test_oracle_search_flow_reproduces_reference(path, oracle):
    z = load(path)
    wt, s0 = int(z["wt"]), int(z["stride0"])
    fl = oracle.search_flow_forward(z["fflow"], z["bflow"], wt, s0)
    assert_close(fl, z["flows"], 1e-5, 1e-5, "accumulated flows")
    gf, gb = oracle.search_flow_backward(z["gflows"], z["fflow"], z["bflow"], z["flows"], wt, s0)
    assert_close(gf, z["gfflow"], 1e-3, 1e-3, "grad_fflow")
    assert_close(gb, z["gbflow"], 1e-3, 1e-3, "grad_bflow")


@pytest.mark.parametrize("path", GATHER, ids=[os.path.basename(f)[:-4] for f in GATHER])
def test_oracle_gather_reproduces_reference(path, oracle):
    z = load(path)
    itype = "int" if "int" in os.path.basename(path) else "float"
    ps, s0 = int(z["ps"]), int(z["stride0"])
    st, counts = oracle.gather_forward(z["v"], z["w"], z["inds"], ps, s0, itype=itype)
    assert th.equal(counts, z["counts"])
    assert_close(st, z["stack"], 1e-5, 1e-5, "stack")
    gv, gw, gi = oracle.gather_backward(z["gs"], z["v"], z["w"], z["inds"], counts, ps, s0, itype=itype)
    assert_close(gv, z["gv"], 1e-3, 1e-3, "grad_v")
    assert_close(gw, z["gw"], 1e-3, 1e-3, "grad_w")
    if itype == "float":
        assert_close(gi, z["gi"], 2e-3, 2e-3, "grad_inds")
