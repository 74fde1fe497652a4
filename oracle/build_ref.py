"""
oracle/build_ref.py -- TEST INFRASTRUCTURE.  Builds the REFERENCE's own CUDA kernels for the hot path
into oracle/_ref/stnls_cuda_ref*.so so the GPU box can run the reference itself as the parity anchor
(SURVEY.md 8c "O1").

* Sources are read where they lie under /root/reference/lib/csrc; patched copies exist only in a
  scratch build dir under /tmp (never in the repo).  The one mechanical patch is
  `.type()` -> `.scalar_type()` (torch 2.11 dropped the implicit DeprecatedTypeProperties ->
  ScalarType conversion).
* Only the hot-path files are built: search/non_local_search*, search/refinement*, agg/gather_add*,
  nn/anchor_self*, and the next rows nn/search_flow*, agg/gather{,_int,_bilin2d}* (+ their header-style .cu
  includes) and a 10-line pybind file.
* Output: oracle/_ref/ (git-ignored, NOT gpurun-ignored).  Run in this container (no GPU needed);
  /root/reference does not exist on the GPU box.
"""
import glob
import os
import re
import shutil
import sys

REF = "/root/reference/lib/csrc"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
SCRATCH = "/tmp/stnls_ref_build"

FILES = ["shared_kernel.cu",
         "search/nls_int.cu", "search/nls_bilin2d.cu",
         "search/non_local_search.cpp", "search/non_local_search_int_kernel.cu",
         "search/non_local_search_bilin2d_kernel.cu",
         "search/refinement.cpp", "search/refinement_int_kernel.cu", "search/refinement_bilin2d_kernel.cu",
         "agg/gather_add.cpp", "agg/gather_add_kernel.cu",
         "nn/anchor_self.cpp", "nn/anchor_self_kernel.cu",
         # next rows: N1 search_flow, N3 NonLocalGather
         "nn/shared_flows.cu", "nn/search_flow.cpp", "nn/search_flow_kernel.cu",
         "agg/gather_int.cu", "agg/gather_bilin2d.cu", "agg/gather.cpp",
         "agg/gather_int_kernel.cu", "agg/gather_bilin2d_kernel.cu"]
HEADER_ONLY = {"shared_kernel.cu", "search/nls_int.cu", "search/nls_bilin2d.cu", "nn/shared_flows.cu",
               "agg/gather_int.cu", "agg/gather_bilin2d.cu"}

PYBIND = """
#include <torch/extension.h>
void init_non_local_search(py::module &);
void init_refinement(py::module &);
void init_anchor_self(py::module &);
void init_gather_add(py::module &);
void init_gather(py::module &);
void init_search_flow(py::module &);
PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  init_non_local_search(m);
  init_refinement(m);
  init_anchor_self(m);
  init_gather_add(m);
  init_gather(m);
  init_search_flow(m);
}
"""


def existing():
    hits = glob.glob(os.path.join(OUT, "stnls_cuda_ref*.so"))
    return hits[0] if hits else None


def build(force=False):
    if existing() and not force:
        return existing()
    if not os.path.isdir(REF):
        return None
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    from torch.utils import cpp_extension
    shutil.rmtree(SCRATCH, ignore_errors=True)
    srcdir = os.path.join(SCRATCH, "src")
    for f in FILES:
        dst = os.path.join(srcdir, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        txt = open(os.path.join(REF, f)).read()
        txt = re.sub(r"\.type\(\)", ".scalar_type()", txt)
        if f == "shared_kernel.cu":
            # get_type(): `my_tensor.type()` -> DeprecatedTypeProperties no longer converts; same intent
            txt = txt.replace("const auto& the_type = my_tensor.scalar_type();\n  at::ScalarType _st = ::detail::scalar_type(the_type);",
                              "at::ScalarType _st = my_tensor.scalar_type();")
        open(dst, "w").write(txt)
    open(os.path.join(srcdir, "pybind_ref.cpp"), "w").write(PYBIND)
    sources = [os.path.join(srcdir, f) for f in FILES if f not in HEADER_ONLY]
    sources.append(os.path.join(srcdir, "pybind_ref.cpp"))
    bdir = os.path.join(SCRATCH, "build")
    os.makedirs(bdir, exist_ok=True)
    cpp_extension.load(name="stnls_cuda_ref", sources=sources, build_directory=bdir,
                       extra_cuda_cflags=["-O2", "-w", "-gencode", "arch=compute_100a,code=sm_100a"],
                       extra_cflags=["-O2", "-w"], verbose=True, is_python_module=False)
    os.makedirs(OUT, exist_ok=True)
    so = glob.glob(os.path.join(bdir, "stnls_cuda_ref*.so"))[0]
    shutil.copy(so, os.path.join(OUT, os.path.basename(so)))
    return existing()


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
