"""ctypes binding of libstnls_b200.so (the C ABI declared in include/stnls_b200.h).

There is no CPU fallback: if the shared library cannot be loaded (and cannot be built because
nvcc is absent) importing the ops raises.  Kernels launch on the current torch CUDA stream.
"""
import ctypes
import os

from . import build as _build

c_int, c_i64, c_f32, c_vp = ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_void_p


class Geom(ctypes.Structure):
    """stnls_geom (include/stnls_b200.h)"""
    _fields_ = [(n, c_int) for n in ["B", "HD", "T", "tq0", "Tq", "tk0", "Tk", "F", "Fp",
                                     "qH", "qW", "kH", "kW", "nH", "nW"]]


class SearchOpts(ctypes.Structure):
    """stnls_search_opts (include/stnls_b200.h)"""
    _fields_ = [("ws", c_int), ("wt", c_int), ("Ks", c_int), ("wr", c_int), ("ps", c_int), ("K", c_int),
                ("stride0", c_int), ("strideQ", c_int), ("dilation", c_int), ("stride1", c_f32),
                ("reflect_bounds", c_int), ("full_ws", c_int), ("patch_offset", c_int),
                ("off_Hq", c_int), ("off_Wq", c_int), ("dist_type", c_int), ("itype_int", c_int),
                ("anchor_self", c_int), ("HD_f", c_int), ("st_offset", c_int)]


class AggOpts(ctypes.Structure):
    """stnls_agg_opts (include/stnls_b200.h)"""
    _fields_ = [(n, c_int) for n in ["ps", "K", "stride0", "dilation", "reflect_bounds",
                                     "patch_offset", "itype_int"]]


EXPORTS = ["stnls_last_error", "stnls_version", "stnls_pack_cl", "stnls_unpack_cl",
           "stnls_search_forward", "stnls_search_backward", "stnls_wpsum_forward",
           "stnls_wpsum_backward", "stnls_gather_forward", "stnls_gather_backward",
           "stnls_search_flow_forward", "stnls_search_flow_backward",
           "stnls_softmax_forward", "stnls_softmax_backward", "stnls_anchor_self_time",
           "stnls_overlap_counts", "stnls_add_inplace", "stnls_ffma_peak", "stnls_seg_copy",
           "stnls_search_forward_planar_noalloc", "stnls_pack_cl_multi", "stnls_unpack_cl_multi",
           "stnls_search_backward_sm", "stnls_wpsum_forward_sm", "stnls_seg_copy2d"]

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if _build.needs_build():
        # sources differ from the ones the library was built from (or there is no library): build, and fail LOUDLY if
        # that is impossible -- a stale library would run old kernels behind a newer header
        try:
            _build.build()
        except Exception as exc:
            raise ImportError("stnls_b200: libstnls_b200.so is missing or older than its sources and could not be "
                              f"built ({exc}); run `python -m stnls_b200.build`") from exc
    L = ctypes.CDLL(path)
    L.stnls_last_error.restype = ctypes.c_char_p
    P = ctypes.POINTER
    L.stnls_pack_cl.argtypes = [c_vp, c_vp] + [c_int] * 8 + [c_i64] * 4 + [c_vp, c_vp, c_vp]
    L.stnls_unpack_cl.argtypes = [c_vp, c_vp] + [c_int] * 8 + [c_i64] * 4 + [c_int, c_vp]
    L.stnls_search_forward.argtypes = [c_vp] * 7 + [P(Geom), P(SearchOpts), c_vp]
    L.stnls_search_forward_planar_noalloc.argtypes = [c_vp] * 5 + [P(Geom), P(SearchOpts), c_vp]
    L.stnls_pack_cl_multi.argtypes = [c_int] + [c_vp] * 14 + [c_vp]
    L.stnls_unpack_cl_multi.argtypes = [c_int] + [c_vp] * 13 + [c_vp]
    L.stnls_seg_copy2d.argtypes = [c_vp] * 6 + [c_int, c_int, c_vp]
    L.stnls_search_backward_sm.argtypes = [c_vp] * 8 + [ctypes.c_float] + [c_vp] * 4 + [c_int, P(Geom), P(SearchOpts), c_vp]
    L.stnls_wpsum_forward_sm.argtypes = [c_vp] * 4 + [c_int, ctypes.c_float] + [c_vp] * 3 + [P(Geom), P(AggOpts), c_vp, c_vp]
    L.stnls_search_backward.argtypes = [c_vp] * 11 + [c_int, P(Geom), P(SearchOpts), c_vp]
    L.stnls_wpsum_forward.argtypes = [c_vp] * 6 + [P(Geom), P(AggOpts), c_vp]
    L.stnls_wpsum_backward.argtypes = [c_vp] * 7 + [P(Geom), P(AggOpts), c_vp]
    L.stnls_gather_forward.argtypes = [c_vp] * 6 + [P(Geom), P(AggOpts), c_vp]
    L.stnls_gather_backward.argtypes = [c_vp] * 7 + [P(Geom), P(AggOpts), c_vp]
    L.stnls_search_flow_forward.argtypes = [c_vp] * 3 + [c_int] * 6 + [c_vp]
    L.stnls_search_flow_backward.argtypes = [c_vp] * 6 + [c_int] * 6 + [c_vp]
    L.stnls_softmax_forward.argtypes = [c_vp, c_vp, c_i64, c_int, c_int, c_f32, c_vp]
    L.stnls_softmax_backward.argtypes = [c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_f32, c_vp]
    L.stnls_anchor_self_time.argtypes = [c_vp, c_vp, c_vp] + [c_int] * 14 + [c_vp]
    L.stnls_overlap_counts.argtypes = [c_vp] + [c_int] * 6
    L.stnls_add_inplace.argtypes = [c_vp, c_vp, c_i64, c_vp]
    L.stnls_ffma_peak.argtypes = [c_vp, c_int, c_int, c_vp]
    L.stnls_seg_copy.argtypes = [c_vp, c_vp, c_vp, c_int, c_int, c_vp]
    for name in EXPORTS:
        getattr(L, name)  # raises AttributeError if a declared symbol is missing
    _lib = L
    return L


def check(rc):
    if rc != 0:
        msg = lib().stnls_last_error().decode()
        if rc == -1:
            raise ValueError(msg)
        raise RuntimeError(msg)
