"""ctypes binding of liborbfe.so (include/orbfe_c.h).  The library is the product; this module only loads it.

There is no fallback: if the shared library is missing the import of any compute entry point fails loudly, and if
no CUDA device is present the library itself returns ORBFE_ERR_CUDA from every constructor.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("ORBFE_SO") or os.path.join(_HERE, "liborbfe.so")    # ORBFE_SO: kernel-tuning builds
CSRC = os.path.join(_HERE, "csrc")

OK, ERR_ARG, ERR_CAPACITY, ERR_CUDA, ERR_NCCL, ERR_INTERNAL = 0, -1, -2, -3, -4, -5
IN_DEVICE, OUT_DEVICE = 1, 2
KNN2_STATIC_MAX_NQ = 16384    # ORBFE_KNN2_STATIC_MAX_NQ (orbfe_c.h)
MAX_LEVELS = 16

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32)]


class OrbfeError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__("liborbfe error %d: %s" % (code, msg))
        self.code = code


def build(force=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a for every kernel (csrc/Makefile); cross-compiles without a GPU."""
    args = ["make", "-C", CSRC, "-s", "-j8"]
    if force:
        args.append("-B")
    subprocess.check_call(args)
    return SO_PATH


# every symbol include/orbfe_c.h declares: (name, restype, argtypes)
_vp, _i, _u, _i64, _sz, _f32p = C.c_void_p, C.c_int, C.c_uint, C.c_int64, C.c_size_t, C.c_void_p
SYMBOLS = {
    "orbfe_version": (C.c_char_p, []),
    "orbfe_device_count": (_i, []),
    "orbfe_kernel_launches": (C.c_uint64, []),
    "orbfe_extractor_create": (_i, [_i, C.POINTER(Params), C.POINTER(_vp)]),
    "orbfe_extractor_destroy": (_i, [_vp]),
    "orbfe_last_error": (C.c_char_p, [_vp]),
    "orbfe_extractor_set_stream": (_i, [_vp, _vp]),
    "orbfe_matcher_set_stream": (_i, [_vp, _vp]),
    "orbfe_extractor_set_workspace_limit": (_i, [_vp, _sz]),
    "orbfe_extractor_tables": (_i, [_vp] + [_vp] * 6),
    "orbfe_extractor_max_keypoints": (_i, [_vp]),
    "orbfe_extract_batch": (_i, [_vp, _vp, _i, _i, _i, _i, _sz, _i, _i, _vp, _vp, _i, _vp, _vp, _u]),
    "orbfe_extract": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp, _vp]),
    "orbfe_extract_stereo_batch": (_i, [_vp, _vp, _i, _i, _i, _i, _sz, _vp, _vp, _i, _vp, C.c_float, C.c_float, _vp,
                                        _vp, _u]),
    "orbfe_extract_match_batch": (_i, [_vp, _vp, _i, _i, _i, _i, _sz, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _u]),
    "orbfe_extractor_stereo_ms": (_i, [_vp, _vp]),
    "orbfe_level_size": (_i, [_vp, _i, _vp, _vp]),
    "orbfe_level_read": (_i, [_vp, _i, _i, _i, _vp, _i]),
    "orbfe_level_candidates": (_i, [_vp, _i, _i, _vp, _i]),
    "orbfe_level_keypoints": (_i, [_vp, _i, _i, _vp, _i]),
    "orbfe_extractor_device_outputs": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "orbfe_extractor_stage_ms": (_i, [_vp, _vp]),
    "orbfe_extractor_last_launches": (_i, [_vp]),
    "orbfe_matcher_create": (_i, [_i, C.POINTER(_vp)]),
    "orbfe_matcher_destroy": (_i, [_vp]),
    "orbfe_matcher_last_error": (C.c_char_p, [_vp]),
    "orbfe_knn2": (_i, [_vp, _vp, _i64, _vp, _i64, _vp, _vp, _u]),
    "orbfe_knn2_batch": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _u]),
    "orbfe_ratio_test": (_i, [_vp, _vp, _i64, C.c_double, _vp]),
    "orbfe_matcher_last_ms": (_i, [_vp, _vp, _vp]),
    "orbfe_search_by_projection_stereo_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp,
                                                     _i, _i, C.c_float, _i, C.c_float, C.c_float, _vp, _vp, _vp, _vp,
                                                     _u]),
    "orbfe_search_by_bow_batch": (_i, [_vp] * 16 + [_i, _i, _i, C.c_float, _i, _vp, _vp, _u]),
    "orbfe_search_by_projection_frame_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, C.c_float, _vp,
                                                    _vp, _i, _i, C.c_float, _i, _vp, _vp, _vp, _vp, _u]),
    "orbfe_search_by_projection_keyframe_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _i, C.c_float,
                                                       _i, _i, _vp, _vp, _vp, _u]),
    "orbfe_search_by_projection_sim3_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _i, C.c_float,
                                                   C.c_float, _vp, _vp, _vp, _u]),
    "orbfe_fuse_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, C.c_float, _vp, _vp, _i, _i, C.c_float, _i, _vp,
                              _vp, _vp, _u]),
    "orbfe_search_by_sim3_batch": (_i, [_vp] * 7 + [_i, _vp, _vp, _vp, _i, _vp, _vp, _i, C.c_float, _vp, _vp, _vp, _vp, _u]),
    "orbfe_search_by_bow_kf_batch": (_i, [_vp] * 15 + [_i, _i, _i, C.c_float, _i, _vp, _vp, _u]),
    "orbfe_search_for_triangulation_batch": (_i, [_vp] * 17 + [_i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _u]),
    "orbfe_distinctive_descriptors": (_i, [_vp, _vp, _vp, _i, C.c_int64, _i, _vp, _vp, _vp, _u]),
    "orbfe_knn2_static_scratch": (_i, [_vp, C.c_int64, _vp, C.c_int64, _vp, _vp, _vp, _vp, C.c_size_t]),
    "orbfe_search_for_triangulation_kb8_batch": (_i, [_vp] * 17 + [_i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _u]),
    "orbfe_search_for_initialization_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _i, C.c_float, _i, _vp,
                                                   _vp, _vp, _u]),
    "orbfe_stereo_fisheye_matches_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp,
                                                _vp, _vp, _vp, _u]),
    "orbfe_undistort_keypoints": (_i, [_vp, _vp, _i64, C.c_float, C.c_float, C.c_float, C.c_float, _vp, _i, _vp, _u]),
    "orbfe_search_by_projection": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, C.c_float, _i, C.c_float,
                                        C.c_float, _vp, _vp, _vp]),
    "orbfe_search_by_projection_batch": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _i,
                                              C.c_float, _i, C.c_float, C.c_float, _vp, _vp, _vp, _u]),
    "orbfe_shard_range": (_i, [_i64, _i, _i, C.POINTER(_i64), C.POINTER(_i64)]),
    "orbfe_nccl_unique_id": (_i, [_vp]),
    "orbfe_nccl_comm_create": (_i, [_vp, _i, _i, _i, C.POINTER(_vp)]),
    "orbfe_nccl_comm_destroy": (_i, [_vp]),
    "orbfe_nccl_last_error": (C.c_char_p, []),
    "orbfe_knn2_sharded": (_i, [_vp, _vp, _i, _i, _vp, _i64, _vp, _i64, _i, _vp, _vp, _vp]),
    "orbfe_knn2_sharded_times": (_i, [_vp, _vp]),
    "orbfe_ubench_pipe": (_i, [_i, _i, C.POINTER(C.c_double)]),
    "orbslam3_open": (_i, [C.c_char_p, C.POINTER(C.c_uint64)]),
    "orbslam3_close": (_i, [C.c_uint64]),
    "orbslam3_configure": (_i, [C.c_uint64, C.POINTER(Params)]),
    "orbslam3_extractFeatures": (_i, [C.c_uint64, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i,
                                      _vp, _vp, _i, _vp, _i, _vp, _i, _vp, _i, _vp, _i,
                                      _vp, _vp, _i, _vp, _i, _vp, _i, _vp, _i, _vp, _i,
                                      _vp, _vp, _vp, _i, _vp, _i, _vp, _i]),
    "orbslam3_bfMatchStereo": (_i, [C.c_uint64, _vp, _i, _vp, _i, _vp, _i, _vp, _i, _vp, _i]),
}

_lib = None


def lib():
    """Load liborbfe.so (raises if it has not been built: there is no other implementation to fall back to)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise OrbfeError(ERR_INTERNAL, "%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                           "(nvcc, sm_100a); there is no CPU fallback" % SO_PATH)
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def ptr(a):
    """void* of a numpy array, a torch tensor (host or device), an int address or None."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError("unsupported buffer type %r" % type(a))
