"""CPU (not gpu): the C-ABI shared library loads, exports every symbol include/orbfe_c.h declares, and fails loudly
(no CPU fallback) when there is no CUDA device.  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from orbslam3lib_b200 import BFMatcher, ORBextractor, OrbfeError, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "orbfe_c.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = src[src.index('extern "C"'):]
    names = re.findall(r"\b((?:orbfe|orbslam3)_[A-Za-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_symbols_exported():
    L = C.CDLL(_lib.SO_PATH)
    names = _declared_functions()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), "liborbfe.so does not export %s" % n


def test_binding_table_matches_header():
    assert sorted(_lib.SYMBOLS) == _declared_functions()


def test_version_and_device_count():
    L = _lib.lib()
    assert b"sm_100a" in L.orbfe_version()
    assert L.orbfe_device_count() >= 0


def test_no_cpu_fallback_without_device():
    L = _lib.lib()
    if L.orbfe_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(OrbfeError) as e:
        ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.code == _lib.ERR_CUDA
    with pytest.raises(OrbfeError) as e:
        BFMatcher()
    assert e.value.code == _lib.ERR_CUDA
    h = C.c_uint64(123)
    assert L.orbslam3_open(b"orbslam3", C.byref(h)) == _lib.ERR_CUDA and h.value == 0


def test_bad_arguments_are_reported():
    L = _lib.lib()
    h = C.c_void_p()
    assert L.orbfe_extractor_create(0, None, C.byref(h)) == _lib.ERR_ARG
    p = _lib.Params(1000, 1.0, 8, 20, 7)    # scaleFactor must be > 1
    assert L.orbfe_extractor_create(0, C.byref(p), C.byref(h)) == _lib.ERR_ARG
    p = _lib.Params(1000, 1.2, _lib.MAX_LEVELS + 1, 20, 7)
    assert L.orbfe_extractor_create(0, C.byref(p), C.byref(h)) == _lib.ERR_ARG
    assert L.orbfe_extractor_destroy(None) == _lib.ERR_ARG
    assert L.orbfe_matcher_destroy(None) == _lib.ERR_ARG
    assert L.orbslam3_close(0) == _lib.ERR_ARG
    assert L.orbfe_extract_batch(None, None, 1, 1, 1, 1, 1, 0, 0, None, None, 0, None, None, 0) == _lib.ERR_ARG
    assert L.orbfe_knn2(None, None, 1, None, 1, None, None, 0) == _lib.ERR_ARG


def test_ratio_test_host():
    # Frame.cc:1230: accept iff distance0 < distance1 * 0.7 (float distances, double product), strict
    idx = np.array([[3, 9], [4, -1], [5, 6], [7, 8], [-1, -1]], np.int32)
    dist = np.array([[10, 20], [5, -1], [70, 100], [69, 100], [-1, -1]], np.int32)
    assert BFMatcher.ratio_test(idx, dist, 0.7).tolist() == [3, -1, -1, 7, -1]
    assert BFMatcher.ratio_test(np.zeros((0, 2), np.int32), np.zeros((0, 2), np.int32)).tolist() == []


def test_keypoint_layout_is_cv_keypoint():
    # cv::KeyPoint: Point2f pt, float size, angle, response, int octave, class_id = 28 bytes
    assert _lib.KP_DTYPE.itemsize == 28
    assert [_lib.KP_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == \
        [0, 4, 8, 12, 16, 20, 24]
