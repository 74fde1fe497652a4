"""Build libstnls_b200.so (C ABI, sm_100a) in-tree with nvcc.  No torch headers involved.

The library is rebuilt only when the SOURCES changed: a hash of csrc/ + include/ is stored beside the .so
(file times do not survive a copy of the tree to another machine).  Builds are serialised with a file lock and the
library is moved into place atomically, so several ranks starting at once cannot load a half-written file."""
import fcntl
import hashlib
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB = os.path.join(_HERE, "libstnls_b200.so")
STAMP = LIB + ".srchash"
SOURCES = ["errors.cu", "layout.cu", "search_fwd.cu", "search_fwd_tab_nls_int.cu", "search_fwd_tab_nls_float.cu",
           "search_fwd_tab_ref_int.cu", "search_fwd_tab_ref_float.cu", "search_fwd_tile_host.cu", "search_fwd_tile_prod.cu",
           "search_fwd_tile_l2.cu", "search_fwd_refine_chan.cu", "search_bwd.cu", "wpsum.cu", "search_flow.cu", "normz.cu", "anchor.cu", "peak.cu", "noalloc.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def source_hash():
    h = hashlib.sha256()
    files = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h")))
    files.append(os.path.join(_HERE, "..", "include", "stnls_b200.h"))
    for f in files:
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS + SOURCES).encode())
    return h.hexdigest()


def needs_build():
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    with open(STAMP) as f:
        return f.read().strip() != source_hash()


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    os.makedirs(os.path.join(_HERE, "_obj"), exist_ok=True)
    with open(os.path.join(_HERE, "_obj", ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)          # one builder at a time (torchrun starts N ranks at once)
        try:
            if not force and not needs_build():   # another rank built it while we waited
                return LIB
            want = source_hash()
            objs, procs = [], []
            for src in SOURCES:
                obj = os.path.join(_HERE, "_obj", src.replace(".cu", ".o"))
                cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
                      ["-c", os.path.join(CSRC, src), "-o", obj]
                procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
                objs.append(obj)
            for cmd, p in procs:
                out = p.communicate()[0].decode()
                if verbose or p.returncode != 0:
                    sys.stderr.write(out)
                if p.returncode != 0:
                    raise RuntimeError("nvcc failed: " + " ".join(cmd))
            tmp = LIB + f".tmp{os.getpid()}"
            subprocess.check_call([_nvcc(), "-shared", "-o", tmp] + objs + ["-lcudart"])
            os.replace(tmp, LIB)
            with open(STAMP + ".tmp", "w") as f:
                f.write(want)
            os.replace(STAMP + ".tmp", STAMP)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
