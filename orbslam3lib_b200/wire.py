"""Host-side mirror of ORB_SLAM3::LynxHardwareAccelerator (reference: app/src/main/cpp/include/
LynxHardwareAcceleration/LynxHardwareAccelerator.h:31-93, .cpp:40-213) on top of the four wire entry points of
liborbfe.so (orbslam3_open / configure / extractFeatures / close = the FastRPC skel of dsp/src/orbslam_dsp.cpp:866,
931,1017-1021; include/orbfe_c.h).  Same method names, argument meaning and the frame-id -> 3-slot matching cache
(MATCHING_CACHE_SIZE, LHA.h:31, LHA.cpp:137,207).  Used by the tests; the C++ form is include/LynxHardwareAccelerator.h.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbfeError, Params, ptr

MATCHING_CACHE_SIZE = 3
MAX_POINTS = 20000


class LynxHardwareAccelerator:
    def __init__(self, width=640, height=400, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7,
                 uri=b"orbslam3"):
        self.width, self.height = int(width), int(height)
        self._h = C.c_uint64(0)
        rc = _lib.lib().orbslam3_open(uri, C.byref(self._h))
        if rc != _lib.OK:
            raise OrbfeError(rc, "orbslam3_open")
        p = Params(int(nfeatures), float(scale_factor), int(nlevels), int(ini_th), int(min_th))
        rc = _lib.lib().orbslam3_configure(self._h, C.byref(p))
        if rc != _lib.OK:
            raise OrbfeError(rc, "orbslam3_configure")
        self.threshold = int(ini_th)
        self.inputImage = np.zeros((self.height, 2 * self.width), np.uint8)     # inputImageRPCMem (LHA.cpp:70)
        mk = lambda dt, n=MAX_POINTS: np.zeros(n, dt)
        self.left = [mk(np.int32), mk(np.int32), mk(np.int32), mk(np.int32), np.zeros((MAX_POINTS, 32), np.uint8)]
        self.right = [mk(np.int32), mk(np.int32), mk(np.int32), mk(np.int32), np.zeros((MAX_POINTS, 32), np.uint8)]
        self.cache = [[mk(np.int16), mk(np.int16), mk(np.int16)] for _ in range(MATCHING_CACHE_SIZE)]
        self.mFrameCounter = 0

    def close(self):
        if self._h.value:
            _lib.lib().orbslam3_close(self._h)
            self._h = C.c_uint64(0)

    __del__ = close

    def StoreInputBuffer(self, frame, stride=None):
        """LHA.cpp:121-123: the side-by-side frame (left eye | right eye)."""
        frame = np.ascontiguousarray(frame, np.uint8)
        assert frame.shape[0] == self.height and frame.shape[1] >= 2 * self.width
        self.inputImage = frame

    def extract_raw(self, lapL, lapR, slot):
        """orbslam3_extractFeatures exactly as LHA.cpp:146-150 calls it -> (rc, totalL, totalR, monoL, monoR)."""
        tl, tr, ml, mr = C.c_int32(0), C.c_int32(0), C.c_int32(0), C.c_int32(0)
        img = self.inputImage
        L, R, M = self.left, self.right, self.cache[slot]
        n = MAX_POINTS
        rc = _lib.lib().orbslam3_extractFeatures(
            self._h, ptr(img), img.size, self.width, self.height, img.strides[0], self.threshold,
            int(lapL[0]), int(lapL[1]), int(lapR[0]), int(lapR[1]),
            C.byref(tl), ptr(L[0]), n, ptr(L[1]), n, ptr(L[2]), n, ptr(L[3]), n, ptr(L[4]), n * 32,
            C.byref(tr), ptr(R[0]), n, ptr(R[1]), n, ptr(R[2]), n, ptr(R[3]), n, ptr(R[4]), n * 32,
            C.byref(ml), C.byref(mr), ptr(M[0]), n, ptr(M[1]), n, ptr(M[2]), n)
        return rc, tl.value, tr.value, ml.value, mr.value

    @staticmethod
    def _keypoints(X, Y, A, Lv, n):
        """LHA.cpp:166-201: integer position, level, and the angle rebuilt from the packed int8 sin/cos word."""
        k = np.zeros(n, KP_DTYPE)
        k["x"], k["y"], k["octave"] = X[:n], Y[:n], Lv[:n]
        cos8 = (A[:n] & 0xFF).astype(np.int8).astype(np.float32)
        sin8 = ((A[:n] >> 8) & 0xFF).astype(np.int8).astype(np.float32)
        a = np.arctan2(sin8 / np.float32(64.0), cos8 / np.float32(64.0)).astype(np.float32) * np.float32(180.0)
        k["angle"] = (a.astype(np.float64) / np.pi).astype(np.float32)    # float * 180.0f / M_PI (double), LHA.cpp:177
        k["class_id"] = -1
        return k

    def ExtractORB(self, vLappingLeft=(0, 0), vLappingRight=(0, 0)):
        """LHA.cpp:133-204 -> (frame id, left keypoints, right keypoints, left descriptors, right descriptors,
        monoLeft, monoRight).  Counts are the true counts (the DSP reports count + 1 and LHA.cpp:155-159 subtracts)."""
        self.mFrameCounter += 1
        slot = self.mFrameCounter % MATCHING_CACHE_SIZE
        rc, tl, tr, ml, mr = self.extract_raw(vLappingLeft, vLappingRight, slot)
        if rc != _lib.OK:
            raise OrbfeError(rc, "orbslam3_extractFeatures")
        nl, nr = min(10000, tl), min(10000, tr)
        kl = self._keypoints(*self.left[:4], nl)
        kr = self._keypoints(*self.right[:4], nr)
        return self.mFrameCounter, kl, kr, self.left[4][:nl].copy(), self.right[4][:nr].copy(), ml, mr

    def BFMatchORB(self, nIdMatchingData, nrows):
        """LHA.cpp:206-213: the cached stereo matches of frame nIdMatchingData (left stereo row -> right stereo row),
        as uint16 vectors.  nrows = rows of the stereo-area LEFT descriptors (Frame.cc:1164 passes them as the
        second matrix)."""
        slot = (nIdMatchingData + 3) % MATCHING_CACHE_SIZE
        i, d1, d2 = self.cache[slot]
        return i[:nrows].view(np.uint16).copy(), d1[:nrows].view(np.uint16).copy(), d2[:nrows].view(np.uint16).copy()

    def bfMatchStereo(self, left_desc, right_desc, cap=None):
        """orbslam3_bfMatchStereo (dsp/src/orbslam_dsp.cpp:567) -> (rc, indices, distances1, distances2) int16."""
        l = np.ascontiguousarray(left_desc, np.uint8).reshape(-1, 32)
        r = np.ascontiguousarray(right_desc, np.uint8).reshape(-1, 32)
        n = l.shape[0] if cap is None else int(cap)
        i, d1, d2 = np.full(max(n, 1), -7, np.int16), np.full(max(n, 1), -7, np.int16), np.full(max(n, 1), -7, np.int16)
        rc = _lib.lib().orbslam3_bfMatchStereo(self._h, ptr(l), l.size, ptr(r), r.size, ptr(i), n, ptr(d1), n, ptr(d2), n)
        return rc, i[:n], d1[:n], d2[:n]
