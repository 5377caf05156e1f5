"""ctypes binding of oracle/_ref/libext_old.so: the REFERENCE'S OWN ORB_SLAM3::ORBextractor (ORBextractor_old.cc),
compiled from /root/reference by oracle/ref_ext/make_ref_ext.py.  Test infrastructure only.

The library exists where /root/reference existed when `make -C oracle ref_ext` ran (the build container; the built
.so travels to the GPU box with the snapshot).  available() says whether it can be used."""
import ctypes as C
import os

import numpy as np

from oracle_lib import KP_DTYPE, _p, _u8

_ORACLE_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ORACLE_DIR, "_ref", "libext_old.so")
_lib = None


def available():
    return os.path.exists(_SO) and os.path.exists(os.path.join(_ORACLE_DIR, "_build", "liborb_oracle.so"))


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_SO)
        L.refext_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.refext_create.restype = C.c_void_p
        L.refext_destroy.argtypes = [C.c_void_p]
        L.refext_tables.argtypes = [C.c_void_p] * 7
        L.refext_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                     C.c_void_p, C.c_int, C.c_void_p]
        L.refext_extract.restype = C.c_int
        L.refext_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.refext_level_pixels.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.refext_level_padded.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.refext_level_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.refext_level_keypoints.restype = C.c_int
        L.refext_distribute.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_int]
        L.refext_distribute.restype = C.c_int
        _lib = L
    return _lib


class RefExtractor:
    """The reference class itself: ORB_SLAM3::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nfeatures, self.nlevels = int(nfeatures), int(nlevels)
        self._h = lib().refext_create(int(nfeatures), float(scale_factor), int(nlevels), int(ini_th), int(min_th))
        n = self.nlevels
        self.scale = np.empty(n, np.float32)
        self.inv_scale = np.empty(n, np.float32)
        self.sigma2 = np.empty(n, np.float32)
        self.inv_sigma2 = np.empty(n, np.float32)
        self.features_per_level = np.empty(n, np.int32)
        self.umax = np.empty(16, np.int32)
        lib().refext_tables(self._h, _p(self.scale), _p(self.inv_scale), _p(self.sigma2), _p(self.inv_sigma2),
                            _p(self.features_per_level), _p(self.umax))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().refext_destroy(self._h)
            self._h = None

    def __call__(self, img, lapping=(0, 0)):
        """operator()(image, cv::Mat(), keypoints, descriptors, vLappingArea) -> (monoIndex, keypoints, descriptors)"""
        img = _u8(img)
        cap = self.nfeatures + 4 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        cnt = C.c_int(0)
        mono = lib().refext_extract(self._h, _p(img), img.shape[1], img.shape[0], img.strides[0], int(lapping[0]),
                                    int(lapping[1]), _p(kps), _p(desc), cap, C.byref(cnt))
        assert cnt.value <= cap
        return mono, kps[:cnt.value].copy(), desc[:cnt.value].copy()

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        lib().refext_level_size(self._h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_pixels(self, level):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        lib().refext_level_pixels(self._h, level, _p(out), w)
        return out

    def level_padded(self, level):
        w, h = self.level_size(level)
        out = np.empty((h + 38, w + 38), np.uint8)
        lib().refext_level_padded(self._h, level, _p(out), w + 38)
        return out

    def level_keypoints(self, level):
        """ComputeKeyPointsOctTree on the pyramid of the last call: the level's keypoints, level coordinates."""
        cap = self.nfeatures + 64
        out = np.zeros(cap, KP_DTYPE)
        n = lib().refext_level_keypoints(self._h, level, _p(out), cap)
        return out[:n].copy()

    def distribute(self, keys, min_x, max_x, min_y, max_y, n_wanted, level=0):
        """ORBextractor::DistributeOctTree on caller-made keys."""
        k = np.ascontiguousarray(keys, KP_DTYPE)
        out = np.zeros(max(len(k), 1), KP_DTYPE)
        n = lib().refext_distribute(self._h, _p(k), len(k), int(min_x), int(max_x), int(min_y), int(max_y),
                                    int(n_wanted), int(level), _p(out), len(out))
        return out[:n].copy()
