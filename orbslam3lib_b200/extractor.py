"""Python mirror of ORB_SLAM3::ORBextractor (include/ORBextractor.h:33-108 / include/ORBextractor_old.h:46-117 of the
reference) on top of the C ABI.  Same constructor arguments, getters and call semantics; the C++ drop-in class is
include/ORBextractor.h of this repository, this module exists for the test-suite and the benchmark.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbfeError, Params, ptr


class ORBextractor:
    """ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) — ORBextractor_old.cc:411-471."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, device=0):
        self._h = None
        L = _lib.lib()
        p = Params(int(nfeatures), float(scaleFactor), int(nlevels), int(iniThFAST), int(minThFAST))
        h = C.c_void_p()
        rc = L.orbfe_extractor_create(int(device), C.byref(p), C.byref(h))
        if rc != _lib.OK:
            raise OrbfeError(rc, "orbfe_extractor_create (no CUDA device?)" if rc == _lib.ERR_CUDA else "bad parameters")
        self._h = h
        self.nfeatures, self.nlevels, self.device = int(nfeatures), int(nlevels), int(device)
        self.scaleFactor = float(np.float32(scaleFactor))
        n = self.nlevels
        self._scale = np.empty(n, np.float32)
        self._inv_scale = np.empty(n, np.float32)
        self._sigma2 = np.empty(n, np.float32)
        self._inv_sigma2 = np.empty(n, np.float32)
        self.features_per_level = np.empty(n, np.int32)
        self.umax = np.empty(16, np.int32)
        L.orbfe_extractor_tables(h, ptr(self._scale), ptr(self._inv_scale), ptr(self._sigma2), ptr(self._inv_sigma2),
                                 ptr(self.features_per_level), ptr(self.umax))
        self.max_keypoints = L.orbfe_extractor_max_keypoints(h)

    def close(self):
        if self._h is not None:
            _lib.lib().orbfe_extractor_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # getters of include/ORBextractor.h:61-82
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactor(self):
        return self.scaleFactor

    def GetScaleFactors(self):
        return self._scale.copy()

    def GetInverseScaleFactors(self):
        return self._inv_scale.copy()

    def GetScaleSigmaSquares(self):
        return self._sigma2.copy()

    def GetInverseScaleSigmaSquares(self):
        return self._inv_sigma2.copy()

    def _check(self, rc):
        if rc != _lib.OK:
            raise OrbfeError(rc, _lib.lib().orbfe_last_error(self._h).decode())

    def set_stream(self, cuda_stream):
        """run on the caller's CUDA stream (int handle, e.g. torch.cuda.current_stream().cuda_stream); None = private."""
        self._check(_lib.lib().orbfe_extractor_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def set_workspace_limit(self, nbytes):
        self._check(_lib.lib().orbfe_extractor_set_workspace_limit(self._h, int(nbytes)))

    def __call__(self, image, vLappingArea=(0, 0)):
        """operator()(image, mask, keypoints, descriptors, vLappingArea) -> (monoIndex, keypoints, descriptors)."""
        image = np.ascontiguousarray(image, np.uint8)
        if image.ndim != 2:
            raise ValueError("CV_8UC1 image expected")
        counts, monos, kps, desc = self.extract_batch(image[None], vLappingArea)
        n = int(counts[0])
        return int(monos[0]), kps[0, :n].copy(), desc[0, :n].copy()

    def extract_batch(self, images, vLappingArea=(0, 0), cap=None):
        """images uint8 (n, h, w) host array -> (counts (n,), monos (n,), kps (n, cap) KP_DTYPE, desc (n, cap, 32))."""
        images = np.ascontiguousarray(images, np.uint8)
        n, h, w = images.shape
        cap = self.max_keypoints if cap is None else int(cap)
        kps = np.zeros((n, cap), KP_DTYPE)
        desc = np.zeros((n, cap, 32), np.uint8)
        counts = np.zeros(n, np.int32)
        monos = np.zeros(n, np.int32)
        self._check(_lib.lib().orbfe_extract_batch(self._h, ptr(images), n, w, h, w, w * h, int(vLappingArea[0]),
                                                   int(vLappingArea[1]), ptr(kps), ptr(desc), cap, ptr(counts),
                                                   ptr(monos), 0))
        return counts, monos, kps, desc

    def extract_batch_raw(self, images, n, w, h, kps, desc, cap, counts, monos, vLappingArea=(0, 0), flags=0):
        """Thin call with caller-owned buffers (numpy arrays, torch tensors — pinned host or device — or addresses)."""
        self._check(_lib.lib().orbfe_extract_batch(self._h, ptr(images), int(n), int(w), int(h), int(w), int(w) * int(h),
                                                   int(vLappingArea[0]), int(vLappingArea[1]), ptr(kps), ptr(desc),
                                                   int(cap), ptr(counts), ptr(monos), int(flags)))

    def extract_stereo_batch(self, images, mb, mbf, cap=None):
        """Frame::Frame stereo ctor path (Frame.cc:142-146) + Frame::ComputeStereoMatches (Frame.cc:827-997).
        images uint8 (npairs, 2, h, w): [p, 0] = left, [p, 1] = right ->
        (counts (npairs, 2), kps (npairs, 2, cap), desc (npairs, 2, cap, 32), uRight (npairs, cap), depth (npairs, cap));
        uRight / depth are mvuRight / mvDepth of the left keypoints (-1 = no match)."""
        images = np.ascontiguousarray(images, np.uint8)
        npairs, two, h, w = images.shape
        if two != 2:
            raise ValueError("images must be (npairs, 2, h, w)")
        cap = self.max_keypoints if cap is None else int(cap)
        kps = np.zeros((npairs, 2, cap), KP_DTYPE)
        desc = np.zeros((npairs, 2, cap, 32), np.uint8)
        counts = np.zeros((npairs, 2), np.int32)
        u_right = np.full((npairs, cap), -1, np.float32)
        depth = np.full((npairs, cap), -1, np.float32)
        self._check(_lib.lib().orbfe_extract_stereo_batch(self._h, ptr(images), npairs, w, h, w, w * h, ptr(kps), ptr(desc),
                                                          cap, ptr(counts), float(mb), float(mbf), ptr(u_right),
                                                          ptr(depth), 0))
        return counts, kps, desc, u_right, depth

    def extract_stereo_batch_raw(self, images, npairs, w, h, kps, desc, cap, counts, mb, mbf, u_right, depth, flags=0):
        """Thin call with caller-owned buffers (numpy arrays, torch tensors — pinned host or device — or addresses)."""
        self._check(_lib.lib().orbfe_extract_stereo_batch(self._h, ptr(images), int(npairs), int(w), int(h), int(w),
                                                          int(w) * int(h), ptr(kps), ptr(desc), int(cap), ptr(counts),
                                                          float(mb), float(mbf), ptr(u_right), ptr(depth), int(flags)))

    def extract_match_batch(self, pairs, lapping=(0, 0)):
        """pairs uint8 (npairs, 2, h, w) -> (counts (npairs, 2), monos (npairs, 2), kps (npairs, 2, cap), desc (npairs, 2, cap,
        32), knn_idx (npairs, cap, 2), knn_dist (npairs, cap, 2)): extraction of both eyes + left -> right kNN per pair in
        one call (orbfe_extract_match_batch)."""
        pairs = np.ascontiguousarray(pairs, np.uint8)
        P, two, h, w = pairs.shape
        assert two == 2
        cap = self.max_keypoints
        kps = np.zeros((P, 2, cap), KP_DTYPE)
        desc = np.zeros((P, 2, cap, 32), np.uint8)
        counts, monos = np.zeros((P, 2), np.int32), np.zeros((P, 2), np.int32)
        idx, dist = np.full((P, cap, 2), -1, np.int32), np.full((P, cap, 2), -1, np.int32)
        self.extract_match_batch_raw(pairs, P, w, h, kps, desc, cap, counts, monos, idx, dist, lapping=lapping)
        return counts, monos, kps, desc, idx, dist

    def extract_match_batch_raw(self, images, npairs, w, h, kps, desc, cap, counts, monos, knn_idx, knn_dist, lapping=(0, 0),
                                flags=0):
        self._check(_lib.lib().orbfe_extract_match_batch(self._h, ptr(images), int(npairs), int(w), int(h), int(w),
                                                         int(w) * int(h), int(lapping[0]), int(lapping[1]), ptr(kps), ptr(desc),
                                                         int(cap), ptr(counts), ptr(monos), ptr(knn_idx), ptr(knn_dist), int(flags)))

    def stereo_ms(self):
        ms = C.c_float(0)
        _lib.lib().orbfe_extractor_stereo_ms(self._h, C.byref(ms))
        return float(ms.value)

    # ---- stage outputs of the last call (parity tests; mvImagePyramid) ----
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        rc = _lib.lib().orbfe_level_size(self._h, level, C.byref(w), C.byref(h))
        if rc != _lib.OK:
            raise OrbfeError(rc, "orbfe_level_size")
        return w.value, h.value

    def _read(self, index, level, which, pad):
        w, h = self.level_size(level)
        out = np.empty((h + 2 * pad, w + 2 * pad), np.uint8)
        self._check(_lib.lib().orbfe_level_read(self._h, index, level, which, ptr(out), out.strides[0]))
        return out

    def level_pixels(self, level, index=0):
        return self._read(index, level, 0, 0)

    def level_padded(self, level, index=0):
        return self._read(index, level, 1, 19)

    def level_blurred(self, level, index=0):
        return self._read(index, level, 2, 0)

    def level_candidates(self, level, index=0):
        n = _lib.lib().orbfe_level_candidates(self._h, index, level, None, 0)
        if n < 0:
            self._check(n)
        out = np.empty((max(n, 1), 3), np.int32)
        n2 = _lib.lib().orbfe_level_candidates(self._h, index, level, ptr(out), n)
        assert n2 == n
        return out[:n].copy()

    def level_keypoints(self, level, index=0):
        cap = self.max_keypoints
        out = np.zeros(cap, KP_DTYPE)
        n = _lib.lib().orbfe_level_keypoints(self._h, index, level, ptr(out), cap)
        if n < 0:
            self._check(n)
        return out[:n].copy()

    def device_outputs(self):
        """-> (kps_ptr, desc_ptr, counts_ptr, n, cap): device addresses of the outputs of the last call's last chunk."""
        k, d, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        n, cap = C.c_int(), C.c_int()
        self._check(_lib.lib().orbfe_extractor_device_outputs(self._h, C.byref(k), C.byref(d), C.byref(c), C.byref(n),
                                                              C.byref(cap)))
        return k.value, d.value, c.value, n.value, cap.value

    def stage_ms(self):
        out = np.zeros(7, np.float32)
        _lib.lib().orbfe_extractor_stage_ms(self._h, ptr(out))
        return dict(zip(("h2d", "pyramid", "fast", "octree", "blur", "describe", "d2h"), out.tolist()))

    def last_launches(self):
        return int(_lib.lib().orbfe_extractor_last_launches(self._h))
