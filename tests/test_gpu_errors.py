"""-m gpu: error behaviour of the C ABI at the extractor boundary (SURVEY.md 8b "Error convention" / "Ownership"):
int status codes, true counts reported when a capacity is too small, empty images return -1 like
ORBextractor_old.cc:1091-1092, nothing exits or throws across the boundary."""
import ctypes as C

import numpy as np
import pytest

from orbslam3lib_b200 import ORBextractor, _lib, scenes
from orbslam3lib_b200._lib import KP_DTYPE, ptr

pytestmark = pytest.mark.gpu


def test_capacity_too_small_reports_true_count():
    img = scenes.make("texture", 320, 240, 9300)
    ex = ORBextractor(400, 1.2, 6, 20, 7)
    mono, kps, desc = ex(img)
    n = len(kps)
    assert n > 100
    cap = 50
    k2 = np.zeros(cap, KP_DTYPE)
    d2 = np.zeros((cap, 32), np.uint8)
    cnt, mn = C.c_int32(0), C.c_int32(0)
    rc = _lib.lib().orbfe_extract(ex._h, ptr(img), 320, 240, 320, 0, 0, ptr(k2), ptr(d2), cap, C.byref(cnt), C.byref(mn))
    assert rc == _lib.ERR_CAPACITY
    assert cnt.value == n and b"capacity" in _lib.lib().orbfe_last_error(ex._h)
    # the handle stays usable
    mono3, kps3, desc3 = ex(img)
    np.testing.assert_array_equal(desc3, desc)


def test_empty_and_bad_images():
    ex = ORBextractor(400, 1.2, 6, 20, 7)
    L = _lib.lib()
    cap = ex.max_keypoints
    k = np.zeros(cap, KP_DTYPE)
    d = np.zeros((cap, 32), np.uint8)
    cnt, mn = C.c_int32(7), C.c_int32(7)
    img = np.zeros((10, 10), np.uint8)
    # empty image (ORBextractor_old.cc:1091-1092: returns -1, no keypoints)
    rc = L.orbfe_extract(ex._h, ptr(img), 0, 0, 0, 0, 0, ptr(k), ptr(d), cap, C.byref(cnt), C.byref(mn))
    assert rc == _lib.OK and cnt.value == 0 and mn.value == -1
    # null image pointer / stride smaller than the width: argument errors, not crashes
    assert L.orbfe_extract(ex._h, None, 64, 64, 64, 0, 0, ptr(k), ptr(d), cap, C.byref(cnt), C.byref(mn)) == _lib.ERR_ARG
    big = np.zeros((64, 64), np.uint8)
    assert L.orbfe_extract(ex._h, ptr(big), 64, 64, 32, 0, 0, ptr(k), ptr(d), cap, C.byref(cnt), C.byref(mn)) == _lib.ERR_ARG
    assert L.orbfe_extract(None, ptr(big), 64, 64, 64, 0, 0, ptr(k), ptr(d), cap, C.byref(cnt), C.byref(mn)) == _lib.ERR_ARG
    # an image smaller than the FAST border (no cell at any level): zero keypoints, mono 0
    tiny = scenes.make("noise", 40, 36, 9301)
    mono, kps, desc = ex(tiny)
    assert len(kps) == 0 and mono == 0


def test_constant_image_and_strided_rows(oracle):
    ex = ORBextractor(300, 1.2, 5, 20, 7)
    flat = np.full((200, 260), 93, np.uint8)
    mono, kps, desc = ex(flat)
    assert len(kps) == 0
    # rows with padding (stride > width): the extractor must honour the stride (the DSP ignores it, DSP:978-981)
    wide = np.zeros((173, 256), np.uint8)
    img = scenes.make("texture", 211, 173, 9302)
    wide[:, :211] = img
    wide[:, 211:] = 255
    cap = ex.max_keypoints
    k = np.zeros(cap, KP_DTYPE)
    d = np.zeros((cap, 32), np.uint8)
    cnt, mn = C.c_int32(0), C.c_int32(0)
    rc = _lib.lib().orbfe_extract(ex._h, ptr(wide), 211, 173, 256, 0, 0, ptr(k), ptr(d), cap, C.byref(cnt), C.byref(mn))
    assert rc == _lib.OK
    rm, rk, rd = oracle.OracleExtractor(300, 1.2, 5, 20, 7)(img)
    assert cnt.value == len(rk) and mn.value == rm
    np.testing.assert_array_equal(d[:cnt.value], rd)
    np.testing.assert_array_equal(k[:cnt.value]["x"], rk["x"])


def test_bad_constructor_arguments():
    from orbslam3lib_b200._lib import OrbfeError
    for args in ((-1, 1.2, 8, 20, 7), (500, 1.2, 0, 20, 7), (500, 1.2, 17, 20, 7), (500, 0.9, 8, 20, 7)):
        with pytest.raises((OrbfeError, ValueError)):
            ORBextractor(*args)


@pytest.mark.gpu
def test_failed_create_does_not_poison_the_next_call(oracle):
    """a CUDA error reported by one call (device 99 does not exist) must not come back as the launch error of the next, valid
    call of the thread: the library clears the runtime's last-error state when it reports it."""
    from orbslam3lib_b200._lib import OrbfeError
    from orbslam3lib_b200 import BFMatcher, scenes
    with pytest.raises(OrbfeError):
        ORBextractor(500, 1.2, 8, 20, 7, device=99)
    with pytest.raises(OrbfeError):
        BFMatcher(device=99)
    img = scenes.make("texture", 320, 240, 77)
    mono, kps, desc = ORBextractor(500, 1.2, 8, 20, 7)(img)
    rmono, rk, rd = oracle.OracleExtractor(500, 1.2, 8, 20, 7)(img)
    np.testing.assert_array_equal(desc, rd)
    q, t = scenes.map_descriptors(300, 700, 78)
    idx, dist = BFMatcher().knnMatch(q, t)
    ri, rdist = oracle.knn2_hamming(q, t)
    np.testing.assert_array_equal(idx, ri)


@pytest.mark.gpu
def test_two_devices_in_one_process(oracle):
    """The opt-in for > 48 KiB of dynamic shared memory is a per-device property: extractors and matchers on
    device 0 AND device 1 of one process, with the configuration whose kernels need the opt-in (nfeatures = 5000: ~87 KiB
    quadtree; batch of 32: 96 KiB resize tiles).  Needs two GPUs."""
    from orbslam3lib_b200 import BFMatcher, ORBextractor, _lib, scenes
    if _lib.lib().orbfe_device_count() < 2:
        pytest.skip("needs two GPUs")
    img = scenes.make("texture", 640, 480, 9900)
    batch = np.stack([img] * 32)
    ref = oracle.OracleExtractor(5000, 1.2, 8, 20, 7)
    _, rk, rd = ref(img)
    for dev in (0, 1, 0):
        ex = ORBextractor(5000, 1.2, 8, 20, 7, device=dev)
        counts, monos, kps, desc = ex.extract_batch(batch)
        assert int(counts[31]) == len(rk)
        np.testing.assert_array_equal(desc[31, :len(rk)], rd)
        idx, dist = BFMatcher(device=dev).knnMatch(rd, rd[::-1].copy())
        ri, rdist = oracle.knn2_hamming(rd, rd[::-1].copy())
        np.testing.assert_array_equal(idx, ri)
