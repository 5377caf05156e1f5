"""ctypes binding of the CPU ORACLE (oracle/_build/liborb_oracle.so).

Test infrastructure only: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
Nothing under orbslam3lib_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_DIR = os.path.dirname(_HERE)
_SO = os.path.join(_ORACLE_DIR, "_build", "liborb_oracle.so")

TRACK_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_xr", "<f4"), ("view_cos", "<f4"),
                        ("track_depth", "<f4"), ("level", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])
assert TRACK_DTYPE.itemsize == 60

TRACK_STEREO_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_xr", "<f4"), ("proj_yr", "<f4"),
                               ("view_cos", "<f4"), ("view_cos_r", "<f4"), ("track_depth", "<f4"), ("level", "<i4"),
                               ("level_r", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])
assert TRACK_STEREO_DTYPE.itemsize == 72

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build(force=False):
    """Compile the oracle with g++ (oracle/Makefile). Building the checker is not using it."""
    if force or not os.path.exists(_SO) or any(
            os.path.getmtime(os.path.join(_ORACLE_DIR, "cpp", f)) > os.path.getmtime(_SO)
            for f in ("orb_oracle.cpp", "orb_oracle.h")):
        subprocess.check_call(["make", "-C", _ORACLE_DIR, "-s"] + (["-B"] if force else []))
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        u8p, i32p, f32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float)
        L.orc_resize_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orc_border_reflect101_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.orc_gaussian7x7_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_fast9_nms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_fast9_nms.restype = C.c_int
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_knn2_hamming.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_extractor_create.restype = C.c_void_p
        L.orc_extractor_destroy.argtypes = [C.c_void_p]
        L.orc_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orc_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_int, C.c_void_p]
        L.orc_extract.restype = C.c_int
        L.orc_level_size.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        for name in ("orc_level_pixels", "orc_level_padded", "orc_level_blurred"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orc_level_candidates.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orc_level_candidates.restype = C.c_int
        L.orc_level_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orc_level_keypoints.restype = C.c_int
        L.orc_extractor_stage_seconds.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_extractor_sort_ties.argtypes = [C.c_void_p]
        L.orc_extractor_sort_ties.restype = C.c_int64
        L.orc_distribute_octtree.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 6 + [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_distribute_octtree.restype = C.c_int
        L.orc_extract_batch.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                        C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                        C.c_void_p, C.c_void_p, C.c_int]
        L.orc_extract_batch.restype = C.c_int
        L.orc_hardware_threads.restype = C.c_int
        L.orc_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                    C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_int,
                                                    C.c_float, C.c_int, C.c_void_p]
        L.orc_search_for_initialization.restype = C.c_int
        L.orc_search_by_projection_stereo.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                                      C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p,
                                                      C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_float, C.c_float,
                                                      C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_search_by_projection_stereo.restype = C.c_int
        L.orc_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p,
                                              C.c_int, C.c_void_p]
        L.orc_search_by_projection.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float,
                                               C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int,
                                               C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_search_by_projection.restype = C.c_int
        L.orc_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_stereo_matches.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _u8(img):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    assert img.ndim == 2
    return img


def resize_linear(img, dw, dh):
    img = _u8(img)
    out = np.empty((dh, dw), np.uint8)
    lib().orc_resize_linear_u8(_p(img), img.shape[1], img.shape[0], img.strides[0], _p(out), dw, dh, dw)
    return out


def border_reflect101(img, border=19):
    img = _u8(img)
    h, w = img.shape
    out = np.empty((h + 2 * border, w + 2 * border), np.uint8)
    lib().orc_border_reflect101_u8(_p(img), w, h, img.strides[0], _p(out), out.strides[0], border)
    return out


def gaussian7x7(img):
    img = _u8(img)
    out = np.empty_like(img)
    lib().orc_gaussian7x7_u8(_p(img), img.shape[1], img.shape[0], img.strides[0], _p(out), out.strides[0])
    return out


def fast9_nms(img, threshold):
    """-> int32 array (n, 3) of (x, y, response) in OpenCV order."""
    img = _u8(img)
    cap = max(16, img.size // 4 + 16)
    out = np.empty((cap, 3), np.int32)
    n = lib().orc_fast9_nms(_p(img), img.shape[1], img.shape[0], img.strides[0], int(threshold), _p(out), cap)
    return out[:n].copy()


def fast_atan2(y, x):
    return float(lib().orc_fast_atan2(float(y), float(x)))


def knn2_hamming(q, t, threads=1):
    """-> (idx int32 (nq,2), dist int32 (nq,2)); OpenCV BFMatcher(NORM_HAMMING).knnMatch(k=2) semantics."""
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
    t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    idx = np.empty((q.shape[0], 2), np.int32)
    dist = np.empty((q.shape[0], 2), np.int32)
    lib().orc_knn2_hamming(_p(q), q.shape[0], _p(t), t.shape[0], _p(idx), _p(dist), int(threads))
    return idx, dist


class OracleExtractor:
    """ORB_SLAM3::ORBextractor (ORBextractor_old.cc) restated on the CPU."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.args = (int(nfeatures), float(scale_factor), int(nlevels), int(ini_th), int(min_th))
        self.nfeatures, self.nlevels = int(nfeatures), int(nlevels)
        self._h = lib().orc_extractor_create(*self.args)
        if not self._h:
            raise ValueError("bad extractor parameters")
        n = self.nlevels
        self.scale = np.empty(n, np.float32)
        self.inv_scale = np.empty(n, np.float32)
        self.sigma2 = np.empty(n, np.float32)
        self.inv_sigma2 = np.empty(n, np.float32)
        self.features_per_level = np.empty(n, np.int32)
        self.umax = np.empty(16, np.int32)
        lib().orc_extractor_tables(self._h, _p(self.scale), _p(self.inv_scale), _p(self.sigma2), _p(self.inv_sigma2),
                                   _p(self.features_per_level), _p(self.umax))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_extractor_destroy(self._h)
            self._h = None

    def __call__(self, img, lapping=(0, 0)):
        """-> (mono_index, keypoints[KP_DTYPE], descriptors uint8 (n,32))"""
        img = _u8(img)
        cap = self.nfeatures + 4 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        cnt = C.c_int(0)
        mono = lib().orc_extract(self._h, _p(img), img.shape[1], img.shape[0], img.strides[0], int(lapping[0]),
                                 int(lapping[1]), _p(kps), _p(desc), cap, C.byref(cnt))
        n = cnt.value
        assert n <= cap
        return mono, kps[:n].copy(), desc[:n].copy()

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        lib().orc_level_size(self._h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_pixels(self, level):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        lib().orc_level_pixels(self._h, level, _p(out), w)
        return out

    def level_padded(self, level):
        w, h = self.level_size(level)
        out = np.empty((h + 38, w + 38), np.uint8)
        lib().orc_level_padded(self._h, level, _p(out), w + 38)
        return out

    def level_blurred(self, level):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        lib().orc_level_blurred(self._h, level, _p(out), w)
        return out

    def level_candidates(self, level):
        w, h = self.level_size(level)
        cap = w * h // 4 + 64
        out = np.empty((cap, 3), np.int32)
        n = lib().orc_level_candidates(self._h, level, _p(out), cap)
        return out[:n].copy()

    def level_keypoints(self, level):
        cap = self.nfeatures + 64
        out = np.zeros(cap, KP_DTYPE)
        n = lib().orc_level_keypoints(self._h, level, _p(out), cap)
        return out[:n].copy()

    def stage_seconds(self):
        out = np.zeros(7, np.float64)
        lib().orc_extractor_stage_seconds(self._h, _p(out))
        return dict(zip(("pyramid", "fast", "octree", "angle", "blur", "desc", "pack"), out.tolist()))

    def sort_ties(self):
        return int(lib().orc_extractor_sort_ties(self._h))

    def distribute(self, keys, min_x, max_x, min_y, max_y, n_wanted):
        """ORBextractor::DistributeOctTree on caller-made keys -> (kept keys, compareNodes ties seen by std::sort)."""
        k = np.ascontiguousarray(keys, KP_DTYPE)
        out = np.zeros(max(len(k), 1), KP_DTYPE)
        ties = C.c_int64(0)
        n = lib().orc_distribute_octtree(self._h, _p(k), len(k), int(min_x), int(max_x), int(min_y), int(max_y),
                                         int(n_wanted), _p(out), len(out), C.byref(ties))
        return out[:n].copy(), int(ties.value)


def extract_batch(imgs, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, lapping=(0, 0), threads=1,
                  want_outputs=True):
    """imgs uint8 (n,h,w) -> (counts, monos, kps (n,cap), desc (n,cap,32)); frames spread over std::threads."""
    imgs = np.ascontiguousarray(imgs, np.uint8)
    n, h, w = imgs.shape
    cap = int(nfeatures) + 4 * int(nlevels) + 64
    counts = np.zeros(n, np.int32)
    monos = np.zeros(n, np.int32)
    kps = np.zeros((n, cap), KP_DTYPE) if want_outputs else None
    desc = np.zeros((n, cap, 32), np.uint8) if want_outputs else None
    lib().orc_extract_batch(int(nfeatures), float(scale_factor), int(nlevels), int(ini_th), int(min_th), _p(imgs), n, w,
                            h, w, int(lapping[0]), int(lapping[1]), _p(kps) if want_outputs else None,
                            _p(desc) if want_outputs else None, cap, _p(counts), _p(monos), int(threads))
    return counts, monos, kps, desc


def stereo_matches(ex_left, ex_right, kps_l, desc_l, kps_r, desc_r, mb, mbf):
    """Frame::ComputeStereoMatches (Frame.cc:827-997).  ex_left / ex_right: OracleExtractor objects whose LAST call
    extracted the left / right image (they hold mvImagePyramid).  -> (uRight f32, depth f32, sad i32, faults)."""
    kl = np.ascontiguousarray(kps_l, KP_DTYPE)
    kr = np.ascontiguousarray(kps_r, KP_DTYPE)
    dl = np.ascontiguousarray(desc_l, np.uint8).reshape(-1, 32)
    dr = np.ascontiguousarray(desc_r, np.uint8).reshape(-1, 32)
    n = kl.shape[0]
    u = np.empty(n, np.float32)
    z = np.empty(n, np.float32)
    sad = np.empty(n, np.int32)
    faults = lib().orc_stereo_matches(ex_left._h, ex_right._h, _p(kl), _p(dl), n, _p(kr), _p(dr), kr.shape[0], float(mb),
                                      float(mbf), _p(u), _p(z), _p(sad))
    return u, z, sad, int(faults)


def search_by_projection(keys_un, desc, u_right, occupied, bounds, scale_factors, pts, th=1.0, far_points=False,
                         th_far=0.0, nn_ratio=0.8):
    """Frame grid + ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) for Nleft == -1
    (ORBmatcher.cc:44-143).  bounds = (mnMinX, mnMinY, mnMaxX, mnMaxY); pts: TRACK_DTYPE array.
    -> (match int32 (npts,), owner int32 (n,), nmatches)"""
    k = np.ascontiguousarray(keys_un, KP_DTYPE)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = k.shape[0]
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    p = np.ascontiguousarray(pts, TRACK_DTYPE)
    match = np.empty(p.shape[0], np.int32)
    owner = np.empty(n, np.int32)
    nm = lib().orc_search_by_projection(_p(k), _p(d), None if ur is None else _p(ur), None if oc is None else _p(oc), n,
                                        float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]), _p(sf),
                                        _p(p), p.shape[0], float(th), int(bool(far_points)), float(th_far),
                                        float(nn_ratio), _p(match), _p(owner))
    return match, owner, int(nm)


def search_by_projection_stereo(keys, desc, occupied, nleft, left_to_right, right_to_left, bounds, scale_factors, pts,
                                th=1.0, far_points=False, th_far=0.0, nn_ratio=0.8):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, ...) for stereo-camera frames (Nleft != -1, ORBmatcher.cc:44-214).
    keys / desc: the N features, left ones first; pts: TRACK_STEREO_DTYPE.
    -> (match_left (npts,), match_right (npts,), owner (N,), nmatches)"""
    k = np.ascontiguousarray(keys, KP_DTYPE)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = k.shape[0]
    oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    ltr = np.ascontiguousarray(left_to_right, np.int32)
    rtl = np.ascontiguousarray(right_to_left, np.int32)
    assert ltr.shape[0] == nleft and rtl.shape[0] == n - nleft
    sf = np.ascontiguousarray(scale_factors, np.float32)
    p = np.ascontiguousarray(pts, TRACK_STEREO_DTYPE)
    ml = np.empty(p.shape[0], np.int32)
    mr = np.empty(p.shape[0], np.int32)
    owner = np.empty(n, np.int32)
    nm = lib().orc_search_by_projection_stereo(_p(k), _p(d), None if oc is None else _p(oc), n, int(nleft), _p(ltr),
                                               _p(rtl), float(bounds[0]), float(bounds[1]), float(bounds[2]),
                                               float(bounds[3]), _p(sf), _p(p), p.shape[0], float(th),
                                               int(bool(far_points)), float(th_far), float(nn_ratio), _p(ml), _p(mr),
                                               _p(owner))
    return ml, mr, owner, int(nm)


def search_for_initialization(keys1, desc1, keys2, desc2, bounds, prev_matched, window_size=100, nn_ratio=0.9,
                              check_orientation=True):
    """ORBmatcher::SearchForInitialization (ORBmatcher.cc:649-764).  prev_matched: (n1, 2) float32 (vbPrevMatched).
    -> (vnMatches12 int32 (n1,), updated prev_matched (n1, 2), nmatches)"""
    k1 = np.ascontiguousarray(keys1, KP_DTYPE)
    k2 = np.ascontiguousarray(keys2, KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    pm = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.empty(k1.shape[0], np.int32)
    nm = lib().orc_search_for_initialization(_p(k1), _p(d1), k1.shape[0], _p(k2), _p(d2), k2.shape[0], float(bounds[0]),
                                             float(bounds[1]), float(bounds[2]), float(bounds[3]), _p(pm),
                                             int(window_size), float(nn_ratio), int(bool(check_orientation)), _p(m12))
    return m12, pm, int(nm)


FRAME_POINT_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("ur", "<f4"), ("vr", "<f4"), ("invz", "<f4"), ("angle", "<f4"),
                              ("level", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])
assert FRAME_POINT_DTYPE.itemsize == 64


def search_by_projection_frame(keys, desc, u_right, occupied, nleft, bounds, scale_factors, mbf, pts, th=15.0, direction=0,
                               check_orientation=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1683-1921).  keys / desc: the N
    slots of the current frame, left first (nleft == N: Nleft == -1).  -> (matchL, matchR, owner, nmatches)"""
    k = np.ascontiguousarray(keys, KP_DTYPE)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = k.shape[0]
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    p = np.ascontiguousarray(pts, FRAME_POINT_DTYPE)
    ml, mr, owner = np.empty(p.shape[0], np.int32), np.empty(p.shape[0], np.int32), np.empty(max(n, 1), np.int32)
    L = lib()
    L.orc_search_by_projection_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                                 C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_float, C.c_void_p, C.c_int,
                                                 C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_search_by_projection_frame.restype = C.c_int
    nm = L.orc_search_by_projection_frame(_p(k), _p(d), None if ur is None else _p(ur), None if oc is None else _p(oc), n,
                                          int(nleft), float(bounds[0]), float(bounds[1]), float(bounds[2]), float(bounds[3]),
                                          _p(sf), float(mbf), _p(p), p.shape[0], float(th), int(direction),
                                          int(bool(check_orientation)), _p(ml), _p(mr), _p(owner))
    return ml, mr, owner[:n], int(nm)


def search_by_projection_keyframe(keys_un, desc, occupied, bounds, scale_factors, pts, th=10.0, orb_dist=100,
                                  check_orientation=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (ORBmatcher.cc:1923-2059).
    -> (match, owner, nmatches)"""
    k = np.ascontiguousarray(keys_un, KP_DTYPE)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = k.shape[0]
    oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    p = np.ascontiguousarray(pts, FRAME_POINT_DTYPE)
    m, owner = np.empty(p.shape[0], np.int32), np.empty(max(n, 1), np.int32)
    L = lib()
    L.orc_search_by_projection_keyframe.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float,
                                                    C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_float,
                                                    C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.orc_search_by_projection_keyframe.restype = C.c_int
    nm = L.orc_search_by_projection_keyframe(_p(k), _p(d), None if oc is None else _p(oc), n, float(bounds[0]),
                                             float(bounds[1]), float(bounds[2]), float(bounds[3]), _p(sf), _p(p), p.shape[0],
                                             float(th), int(orb_dist), int(bool(check_orientation)), _p(m), _p(owner))
    return m, owner[:n], int(nm)


def search_by_bow_kf(desc1, angle1, valid1, fv1, desc2, angle2, valid2, fv2, nn_ratio=0.75, check_orientation=True):
    """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (ORBmatcher.cc:766-906).  -> (matches12 int32 (n1,), nmatches)"""
    d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    a1, a2 = np.ascontiguousarray(angle1, np.float32), np.ascontiguousarray(angle2, np.float32)
    v1, v2 = np.ascontiguousarray(valid1, np.uint8), np.ascontiguousarray(valid2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]
    b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    m = np.empty(max(len(d1), 1), np.int32)
    L = lib()
    L.orc_search_by_bow_kf.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_float, C.c_int, C.c_void_p]
    L.orc_search_by_bow_kf.restype = C.c_int
    nm = L.orc_search_by_bow_kf(_p(d1), _p(a1), _p(v1), len(d1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(d2), _p(a2), _p(v2),
                                _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), float(nn_ratio), int(bool(check_orientation)), _p(m))
    return m[:len(d1)].copy(), int(nm)


def search_for_triangulation(keys1, desc1, free1, stereo1, fv1, keys2, desc2, free2, stereo2, fv2, f12, ep, scale_factors,
                             level_sigma2, coarse=False, check_orientation=True):
    """ORBmatcher::SearchForTriangulation (ORBmatcher.cc:908-1153), one camera per key frame.  -> (vMatches12, nmatches)"""
    k1, k2 = np.ascontiguousarray(keys1, KP_DTYPE), np.ascontiguousarray(keys2, KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    f1, f2 = np.ascontiguousarray(free1, np.uint8), np.ascontiguousarray(free2, np.uint8)
    s1 = None if stereo1 is None else np.ascontiguousarray(stereo1, np.uint8)
    s2 = None if stereo2 is None else np.ascontiguousarray(stereo2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]
    b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    F = np.ascontiguousarray(f12, np.float32).reshape(9)
    e = np.ascontiguousarray(ep, np.float32).reshape(2)
    sf, sg = np.ascontiguousarray(scale_factors, np.float32), np.ascontiguousarray(level_sigma2, np.float32)
    m = np.empty(max(len(k1), 1), np.int32)
    L = lib()
    L.orc_search_for_triangulation.argtypes = [C.c_void_p] * 4 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 7 + \
        [C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_int, C.c_void_p]
    L.orc_search_for_triangulation.restype = C.c_int
    nm = L.orc_search_for_triangulation(_p(k1), _p(d1), _p(f1), None if s1 is None else _p(s1), len(k1), _p(a[0]), _p(a[1]),
                                        _p(a[2]), len(a[0]), _p(k2), _p(d2), _p(f2), None if s2 is None else _p(s2), _p(b[0]),
                                        _p(b[1]), _p(b[2]), len(b[0]), _p(F), _p(e), _p(sf), _p(sg), int(bool(coarse)),
                                        int(bool(check_orientation)), _p(m))
    return m[:len(k1)].copy(), int(nm)


def distinctive_descriptors(desc, offsets):
    """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401) per block of rows.  -> (best_idx, best_median); -1 for
    a point without rows."""
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    off = np.asarray(offsets, np.int64)
    L = lib()
    L.orc_distinctive_descriptor.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.orc_distinctive_descriptor.restype = C.c_int
    bi, bm = np.full(len(off) - 1, -1, np.int32), np.full(len(off) - 1, -1, np.int32)
    med = C.c_int32(0)
    for p in range(len(off) - 1):
        n = int(off[p + 1] - off[p])
        if n <= 0:
            continue
        blk = np.ascontiguousarray(d[off[p]:off[p + 1]])
        bi[p] = L.orc_distinctive_descriptor(_p(blk), n, C.byref(med))
        bm[p] = med.value
    return bi, bm


def search_by_projection_sim3(keys_un, desc, occupied, bounds, scale_factors, pts, th=10.0, ratio_hamming=1.0):
    """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) (ORBmatcher.cc:428-533, :535-647).
    -> (match, owner, nmatches)"""
    k = np.ascontiguousarray(keys_un, KP_DTYPE)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = k.shape[0]
    oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    p = np.ascontiguousarray(pts, FRAME_POINT_DTYPE)
    m, owner = np.empty(max(p.shape[0], 1), np.int32), np.empty(max(n, 1), np.int32)
    L = lib()
    L.orc_search_by_projection_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float,
                                                C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                                C.c_void_p]
    L.orc_search_by_projection_sim3.restype = C.c_int
    nm = L.orc_search_by_projection_sim3(_p(k), _p(d), None if oc is None else _p(oc), n, float(bounds[0]), float(bounds[1]),
                                         float(bounds[2]), float(bounds[3]), _p(sf), _p(p), p.shape[0], float(th),
                                         float(ratio_hamming), _p(m), _p(owner))
    return m[:p.shape[0]], owner[:n], int(nm)


def fuse(keys, desc, u_right, bounds, scale_factors, inv_level_sigma2, bf, pts, th=3.0, check_reprojection=True):
    """The search of ORBmatcher::Fuse (ORBmatcher.cc:1155-1345 / :1347-1462).  -> (best_idx, best_dist, count)"""
    k = np.ascontiguousarray(keys, KP_DTYPE)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = k.shape[0]
    ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    iv = np.ascontiguousarray(inv_level_sigma2, np.float32)
    p = np.ascontiguousarray(pts, FRAME_POINT_DTYPE)
    bi, bd = np.empty(max(p.shape[0], 1), np.int32), np.empty(max(p.shape[0], 1), np.int32)
    L = lib()
    L.orc_fuse.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p,
                           C.c_void_p, C.c_float, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p, C.c_void_p]
    L.orc_fuse.restype = C.c_int
    nf = L.orc_fuse(_p(k), _p(d), None if ur is None else _p(ur), n, float(bounds[0]), float(bounds[1]), float(bounds[2]),
                    float(bounds[3]), _p(sf), _p(iv), float(bf), _p(p), p.shape[0], float(th), int(bool(check_reprojection)),
                    _p(bi), _p(bd))
    return bi[:p.shape[0]], bd[:p.shape[0]], int(nf)


def search_by_sim3(keys_un1, desc1, keys_un2, desc2, bounds1, bounds2, scale_factors, pts1, pts2, th=7.5):
    """ORBmatcher::SearchBySim3 (ORBmatcher.cc:1464-1681).  -> (matches12, vnMatch1, vnMatch2, nFound)"""
    k1, k2 = np.ascontiguousarray(keys_un1, KP_DTYPE), np.ascontiguousarray(keys_un2, KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    n1, n2 = k1.shape[0], k2.shape[0]
    b1, b2 = np.ascontiguousarray(bounds1, np.float32), np.ascontiguousarray(bounds2, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    p1, p2 = np.ascontiguousarray(pts1, FRAME_POINT_DTYPE), np.ascontiguousarray(pts2, FRAME_POINT_DTYPE)
    assert p1.shape[0] == n1 and p2.shape[0] == n2
    m12, v1, v2 = np.empty(max(n1, 1), np.int32), np.empty(max(n1, 1), np.int32), np.empty(max(n2, 1), np.int32)
    L = lib()
    L.orc_search_by_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_search_by_sim3.restype = C.c_int
    nf = L.orc_search_by_sim3(_p(k1), _p(d1), n1, _p(k2), _p(d2), n2, _p(b1), _p(b2), _p(sf), _p(p1), _p(p2), float(th),
                              _p(m12), _p(v1), _p(v2))
    return m12[:n1], v1[:n1], v2[:n2], int(nf)


def search_by_bow(desc_kf, angle_kf, valid_kf, fv_kf, desc_f, angle_f, nleft_f, fv_f, nn_ratio=0.7, check_orientation=True):
    """ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches) (ORBmatcher.cc:224-425); fv_* = (node_ids, start, idx) CSR.
    -> (matches int32 (nF,), nmatches)"""
    dk = np.ascontiguousarray(desc_kf, np.uint8).reshape(-1, 32)
    df = np.ascontiguousarray(desc_f, np.uint8).reshape(-1, 32)
    ak, af = np.ascontiguousarray(angle_kf, np.float32), np.ascontiguousarray(angle_f, np.float32)
    vk = np.ascontiguousarray(valid_kf, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv_kf]
    b = [np.ascontiguousarray(x, np.int32) for x in fv_f]
    m = np.empty(max(len(df), 1), np.int32)
    L = lib()
    L.orc_search_by_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                    C.c_float, C.c_int, C.c_void_p]
    L.orc_search_by_bow.restype = C.c_int
    nm = L.orc_search_by_bow(_p(dk), _p(ak), _p(vk), len(dk), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]), _p(df), _p(af), len(df),
                             int(nleft_f), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), float(nn_ratio),
                             int(bool(check_orientation)), _p(m))
    return m[:len(df)].copy(), int(nm)


FISHEYE_RIG_DTYPE = np.dtype([("cam1", "<f4", 8), ("cam2", "<f4", 8), ("Rlr", "<f4", 9), ("tlr", "<f4", 3),
                              ("precision", "<f4"), ("nlevels", "<i4"), ("level_sigma2", "<f4", 16)])
_EIGEN_HOOK = None


def use_eigen_svd(on=True):
    """Route the 4x4 null-vector solve of the fisheye oracle through the reference's vendored Eigen 3.4.0
    (oracle/_ref/libeigen_svd.so, `make -C oracle ref`; exists only where /root/reference does).  Fixture generation
    and the restatement-vs-Eigen test use it; returns False when the helper is not built."""
    global _EIGEN_HOOK
    L = lib()
    L.orc_set_null_vector4_hook.argtypes = [C.c_void_p]
    if not on:
        L.orc_set_null_vector4_hook(None)
        return True
    so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_ref", "libeigen_svd.so")
    if not os.path.exists(so):
        return False
    if _EIGEN_HOOK is None:
        _EIGEN_HOOK = C.CDLL(so)
    L.orc_set_null_vector4_hook(C.cast(_EIGEN_HOOK.ref_null_vector4, C.c_void_p))
    return True


def stereo_fisheye_matches(keys_left, mono_left, keys_right, mono_right, knn_idx, knn_dist, rig):
    """Frame::ComputeStereoFishEyeMatches after BFMatchORB (Frame.cc:1164-1204).
    -> (mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, mvStereo3Dpoints (nL, 3), status, nMatches, slack)"""
    kl = np.ascontiguousarray(keys_left, KP_DTYPE)
    kr = np.ascontiguousarray(keys_right, KP_DTYPE)
    ki = np.ascontiguousarray(knn_idx, np.int32)
    kd = np.ascontiguousarray(knn_dist, np.int32)
    rg = np.ascontiguousarray(rig, FISHEYE_RIG_DTYPE)
    nl, nr = kl.shape[0], kr.shape[0]
    ltr, rtl = np.empty(nl, np.int32), np.empty(nr, np.int32)
    depth, p3d, status = np.empty(nl, np.float32), np.empty((nl, 3), np.float32), np.empty(nl, np.int32)
    slack = np.empty(nl, np.float32)
    L = lib()
    L.orc_stereo_fisheye_matches.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 9
    L.orc_stereo_fisheye_matches.restype = C.c_int
    n = L.orc_stereo_fisheye_matches(_p(kl), nl, int(mono_left), _p(kr), nr, int(mono_right), _p(ki), _p(kd), _p(rg),
                                     _p(ltr), _p(rtl), _p(depth), _p(p3d), _p(status), _p(slack))
    return ltr, rtl, depth, p3d, status, int(n), slack


def search_for_triangulation_kb8(keys1, desc1, free1, nleft1, fv1, keys2, desc2, free2, nleft2, fv2, rig, t12, level_sigma2,
                                 coarse=False, check_orientation=True):
    """ORBmatcher::SearchForTriangulation for two-camera (KannalaBrandt8) key frames (ORBmatcher.cc:908-1153, :1014-1043).
    -> (vMatches12, nmatches, slack)"""
    k1, k2 = np.ascontiguousarray(keys1, KP_DTYPE), np.ascontiguousarray(keys2, KP_DTYPE)
    d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    f1, f2 = np.ascontiguousarray(free1, np.uint8), np.ascontiguousarray(free2, np.uint8)
    a = [np.ascontiguousarray(x, np.int32) for x in fv1]
    b = [np.ascontiguousarray(x, np.int32) for x in fv2]
    rg = np.ascontiguousarray(rig, FISHEYE_RIG_DTYPE)
    T = np.ascontiguousarray(t12, np.float32).reshape(48)
    sg = np.ascontiguousarray(level_sigma2, np.float32)
    m, sl = np.empty(max(len(k1), 1), np.int32), np.empty(max(len(k1), 1), np.float32)
    L = lib()
    L.orc_search_for_triangulation_kb8.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_int] + [C.c_void_p] * 3 + [C.c_int] + \
        [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.orc_search_for_triangulation_kb8.restype = C.c_int
    nm = L.orc_search_for_triangulation_kb8(_p(k1), _p(d1), _p(f1), len(k1), int(nleft1), _p(a[0]), _p(a[1]), _p(a[2]), len(a[0]),
                                            _p(k2), _p(d2), _p(f2), int(nleft2), _p(b[0]), _p(b[1]), _p(b[2]), len(b[0]), _p(rg),
                                            _p(T), _p(sg), int(bool(coarse)), int(bool(check_orientation)), _p(m), _p(sl))
    return m[:len(k1)].copy(), int(nm), sl[:len(k1)].copy()


def kb_triangulate_matches(rig, x1, y1, x2, y2, sigma_level, unc):
    """KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:300-363) of rig.cam1 against rig.cam2 with (Rlr, tlr): z1 or
    the negative reason."""
    rg = np.ascontiguousarray(rig, FISHEYE_RIG_DTYPE)
    L = lib()
    L.orc_kb_triangulate_matches.argtypes = [C.c_void_p] + [C.c_float] * 6
    L.orc_kb_triangulate_matches.restype = C.c_float
    return float(L.orc_kb_triangulate_matches(_p(rg), float(x1), float(y1), float(x2), float(y2), float(sigma_level), float(unc)))


def undistort_keypoints(kps, fx, fy, cx, cy, dist):
    """Frame::UndistortKeyPoints (Frame.cc:763-796) -> mvKeysUn."""
    k = np.ascontiguousarray(kps, KP_DTYPE)
    d = np.ascontiguousarray(dist, np.float32)
    out = np.empty_like(k)
    lib().orc_undistort_keypoints(_p(k), k.shape[0], float(fx), float(fy), float(cx), float(cy), _p(d), d.shape[0], _p(out))
    return out


def hardware_threads():
    return int(lib().orc_hardware_threads())
