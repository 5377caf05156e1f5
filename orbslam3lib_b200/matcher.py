"""Python mirror of the matcher call sites of the reference: cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, 2)
(Frame.cc:45,1227-1231) and LynxHardwareAccelerator::BFMatchORB (LynxHardwareAccelerator.cpp:206-213), on top of
the C ABI (orbfe_knn2 / orbfe_knn2_batch)."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbfeError, ptr
from .scenes import TRACK_DTYPE, TRACK_STEREO_DTYPE


class BFMatcher:
    """BFMatcher(NORM_HAMMING) restricted to what the reference uses: knnMatch with k = 2 on 32-byte rows."""

    def __init__(self, device=0):
        self._h = None
        h = C.c_void_p()
        rc = _lib.lib().orbfe_matcher_create(int(device), C.byref(h))
        if rc != _lib.OK:
            raise OrbfeError(rc, "orbfe_matcher_create (no CUDA device?)")
        self._h = h
        self.device = int(device)

    def close(self):
        if self._h is not None:
            _lib.lib().orbfe_matcher_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != _lib.OK:
            raise OrbfeError(rc, _lib.lib().orbfe_matcher_last_error(self._h).decode())

    def set_stream(self, cuda_stream):
        self._check(_lib.lib().orbfe_matcher_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def knnMatch(self, query, train, k=2):
        """-> (idx int32 (nq, 2), dist int32 (nq, 2)); ties keep the lower train index first; missing = -1."""
        if k != 2:
            raise ValueError("the reference path uses k = 2 only (Frame.cc:1227)")
        q = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
        idx = np.full((q.shape[0], 2), -1, np.int32)
        dist = np.full((q.shape[0], 2), -1, np.int32)
        self._check(_lib.lib().orbfe_knn2(self._h, ptr(q), q.shape[0], ptr(t), t.shape[0], ptr(idx), ptr(dist), 0))
        return idx, dist

    def knn2_raw(self, q, nq, t, nt, idx, dist, flags=0):
        """caller-owned buffers (numpy / torch host or device tensors / addresses)."""
        self._check(_lib.lib().orbfe_knn2(self._h, ptr(q), int(nq), ptr(t), int(nt), ptr(idx), ptr(dist), int(flags)))

    def knnMatchBatch(self, query, nq, train, nt):
        """query (p, q_cap, 32), train (p, t_cap, 32), per-pair row counts nq / nt -> idx, dist (p, q_cap, 2)."""
        q = np.ascontiguousarray(query, np.uint8)
        t = np.ascontiguousarray(train, np.uint8)
        p, qc, tc = q.shape[0], q.shape[1], t.shape[1]
        nq = np.ascontiguousarray(nq, np.int32)
        nt = np.ascontiguousarray(nt, np.int32)
        idx = np.full((p, qc, 2), -1, np.int32)
        dist = np.full((p, qc, 2), -1, np.int32)
        self._check(_lib.lib().orbfe_knn2_batch(self._h, ptr(q), ptr(nq), qc, ptr(t), ptr(nt), tc, p, ptr(idx),
                                                ptr(dist), 0))
        return idx, dist

    def knn2_batch_raw(self, q, nq, q_cap, t, nt, t_cap, npairs, idx, dist, flags=0):
        self._check(_lib.lib().orbfe_knn2_batch(self._h, ptr(q), ptr(nq), int(q_cap), ptr(t), ptr(nt), int(t_cap),
                                                int(npairs), ptr(idx), ptr(dist), int(flags)))

    def last_ms(self):
        ms, n = C.c_float(), C.c_int()
        _lib.lib().orbfe_matcher_last_ms(self._h, C.byref(ms), C.byref(n))
        return ms.value, n.value

    @staticmethod
    def ratio_test(idx, dist, ratio=0.7):
        """Lowe ratio of Frame.cc:1230 -> int32 (nq,) best train index or -1."""
        idx = np.ascontiguousarray(idx, np.int32)
        dist = np.ascontiguousarray(dist, np.int32)
        out = np.empty(idx.shape[0], np.int32)
        _lib.lib().orbfe_ratio_test(ptr(idx), ptr(dist), idx.shape[0], float(ratio), ptr(out))
        return out

    def BFMatchORB(self, left_desc, right_desc):
        """LynxHardwareAccelerator::BFMatchORB (LHA:206-213) as consumed at Frame.cc:1164-1182: per LEFT row the
        best right index and the two smallest distances, uint16 like the reference vectors."""
        idx, dist = self.knnMatch(left_desc, right_desc)
        return idx[:, 0].astype(np.uint16), dist[:, 0].astype(np.uint16), dist[:, 1].astype(np.uint16)


class ORBmatcher:
    """ORB_SLAM3::ORBmatcher restricted to the guided search built so far: SearchByProjection(F, vpMapPoints, th,
    bFarPoints, thFarPoints) (ORBmatcher.cc:44-143) on frames with Nleft == -1.  A frame is passed as its arrays
    (mvKeysUn, mDescriptors, mvuRight, image bounds, mvScaleFactors), map points as a TRACK_DTYPE array
    (scenes.TRACK_DTYPE == orbfe_track_point)."""

    TH_HIGH, TH_LOW = 100, 50

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self._bf = BFMatcher(device)

    def SearchByProjection(self, keys_un, desc, u_right, bounds, scale_factors, track_points, th=1.0, bFarPoints=False,
                           thFarPoints=50.0, occupied=None):
        """-> (nmatches, match int32 (npts,), owner int32 (n,)): match[i] = feature given to point i
        (F.mvpMapPoints[match[i]] = vpMapPoints[i]) or -1; owner[j] = last point assigned to feature j or -1."""
        k = np.ascontiguousarray(keys_un, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        p = np.ascontiguousarray(track_points, TRACK_DTYPE)
        match = np.full(p.shape[0], -1, np.int32)
        owner = np.full(k.shape[0], -1, np.int32)
        nm = C.c_int32(0)
        self._bf._check(_lib.lib().orbfe_search_by_projection(
            self._bf._h, ptr(k), ptr(d), ptr(ur), ptr(oc), k.shape[0], ptr(b), ptr(sf), sf.shape[0], ptr(p), p.shape[0],
            float(th), int(bool(bFarPoints)), float(thFarPoints), self.mfNNratio, ptr(match), ptr(owner), C.byref(nm)))
        return int(nm.value), match, owner

    def search_by_projection_batch_raw(self, keys_un, desc, u_right, occupied, n, cap, bounds, scale_factors, pts, npts,
                                       pts_cap, nframes, th, match, owner, nmatches, bFarPoints=False, thFarPoints=50.0,
                                       flags=0):
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        self._bf._check(_lib.lib().orbfe_search_by_projection_batch(
            self._bf._h, ptr(keys_un), ptr(desc), ptr(u_right), ptr(occupied), ptr(n), int(cap), ptr(b), ptr(sf),
            sf.shape[0], ptr(pts), ptr(npts), int(pts_cap), int(nframes), float(th), int(bool(bFarPoints)),
            float(thFarPoints), self.mfNNratio, ptr(match), ptr(owner), ptr(nmatches), int(flags)))

    def SearchByProjectionStereo(self, keys, desc, nleft, left_to_right, right_to_left, bounds, scale_factors,
                                 track_points, th=1.0, bFarPoints=False, thFarPoints=50.0, occupied=None):
        """The same call on a stereo-camera frame (Nleft != -1, ORBmatcher.cc:44-214): keys / desc = mvKeys followed by
        mvKeysRight, track_points: TRACK_STEREO_DTYPE -> (nmatches, match_left, match_right, owner (N,))."""
        k = np.ascontiguousarray(keys, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = k.shape[0]
        oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        ltr = np.full(max(n, 1), -1, np.int32)
        rtl = np.full(max(n, 1), -1, np.int32)
        ltr[:nleft] = left_to_right
        rtl[:n - nleft] = right_to_left
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        p = np.ascontiguousarray(track_points, TRACK_STEREO_DTYPE)
        ml = np.full(p.shape[0], -1, np.int32)
        mr = np.full(p.shape[0], -1, np.int32)
        owner = np.full(n, -1, np.int32)
        nn, nl, npp, nm = (np.array([v], np.int32) for v in (n, nleft, p.shape[0], 0))
        self._bf._check(_lib.lib().orbfe_search_by_projection_stereo_batch(
            self._bf._h, ptr(k), ptr(d), ptr(oc), ptr(nn), ptr(nl), ptr(ltr), ptr(rtl), n, ptr(b), ptr(sf), sf.shape[0],
            ptr(p), ptr(npp), p.shape[0], 1, float(th), int(bool(bFarPoints)), float(thFarPoints), self.mfNNratio, ptr(ml),
            ptr(mr), ptr(owner), ptr(nm), 0))
        return int(nm[0]), ml, mr, owner

    def SearchByProjectionFrame(self, keys, desc, nleft, bounds, scale_factors, frame_points, th, mbf=0.0, u_right=None,
                                occupied=None, direction=0):
        """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1683-1921), the per-frame
        tracking call.  keys / desc: the N slots of CurrentFrame, left first (nleft == N: Nleft == -1, keys = mvKeysUn,
        u_right = mvuRight); frame_points: scenes.FRAME_POINT_DTYPE (the MapPoints of LastFrame projected by the caller);
        direction: 0, 1 = bForward, 2 = bBackward.  -> (nmatches, match_left, match_right, owner (N,))"""
        from .scenes import FRAME_POINT_DTYPE
        k = np.ascontiguousarray(keys, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = k.shape[0]
        ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        p = np.ascontiguousarray(frame_points, FRAME_POINT_DTYPE)
        ml, mr = np.full(p.shape[0], -1, np.int32), np.full(p.shape[0], -1, np.int32)
        owner = np.full(max(n, 1), -1, np.int32)
        nn, nl, npp, nm, dr = (np.array([v], np.int32) for v in (n, nleft, p.shape[0], 0, direction))
        self.search_frame_batch_raw(k, d, ur, oc, nn, nl, dr, n, bounds, scale_factors, mbf, p, npp, p.shape[0], 1, th, ml, mr,
                                    owner, nm)
        return int(nm[0]), ml, mr, owner[:n]

    def search_frame_batch_raw(self, keys, desc, u_right, occupied, n, nleft, direction, cap, bounds, scale_factors, mbf, pts,
                               npts, pts_cap, nframes, th, match_left, match_right, owner, nmatches, flags=0):
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        self._bf._check(_lib.lib().orbfe_search_by_projection_frame_batch(
            self._bf._h, ptr(keys), ptr(desc), ptr(u_right), ptr(occupied), ptr(n), ptr(nleft), ptr(direction), int(cap),
            ptr(b), ptr(sf), sf.shape[0], float(mbf), ptr(pts), ptr(npts), int(pts_cap), int(nframes), float(th),
            int(self.mbCheckOrientation), ptr(match_left), ptr(match_right), ptr(owner), ptr(nmatches), int(flags)))

    def SearchByProjectionKeyFrame(self, keys_un, desc, bounds, scale_factors, frame_points, th, ORBdist, occupied=None):
        """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (ORBmatcher.cc:1923-2059).
        occupied[i] = CurrentFrame.mvpMapPoints[i] != NULL.  -> (nmatches, match, owner)"""
        from .scenes import FRAME_POINT_DTYPE
        k = np.ascontiguousarray(keys_un, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = k.shape[0]
        oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        p = np.ascontiguousarray(frame_points, FRAME_POINT_DTYPE)
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        m, owner = np.full(p.shape[0], -1, np.int32), np.full(max(n, 1), -1, np.int32)
        nn, npp, nm = (np.array([v], np.int32) for v in (n, p.shape[0], 0))
        self._bf._check(_lib.lib().orbfe_search_by_projection_keyframe_batch(
            self._bf._h, ptr(k), ptr(d), ptr(oc), ptr(nn), n, ptr(b), ptr(sf), sf.shape[0], ptr(p), ptr(npp), p.shape[0], 1,
            float(th), int(ORBdist), int(self.mbCheckOrientation), ptr(m), ptr(owner), ptr(nm), 0))
        return int(nm[0]), m, owner[:n]

    def SearchByProjectionSim3(self, keys_un, desc, bounds, scale_factors, frame_points, th, ratioHamming=1.0, occupied=None):
        """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) (ORBmatcher.cc:428-533; :535-647 fills
        vpMatchedKF from the same result).  occupied[i] = vpMatched[i] != NULL.  -> (nmatches, match, owner)"""
        from .scenes import FRAME_POINT_DTYPE
        k = np.ascontiguousarray(keys_un, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = k.shape[0]
        oc = None if occupied is None else np.ascontiguousarray(occupied, np.uint8)
        p = np.ascontiguousarray(frame_points, FRAME_POINT_DTYPE)
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        m, owner = np.full(max(p.shape[0], 1), -1, np.int32), np.full(max(n, 1), -1, np.int32)
        nn, npp, nm = (np.array([v], np.int32) for v in (n, p.shape[0], 0))
        self._bf._check(_lib.lib().orbfe_search_by_projection_sim3_batch(
            self._bf._h, ptr(k), ptr(d), ptr(oc), ptr(nn), n, ptr(b), ptr(sf), sf.shape[0], ptr(p), ptr(npp), p.shape[0], 1,
            float(th), float(ratioHamming), ptr(m), ptr(owner), ptr(nm), 0))
        return int(nm[0]), m[:p.shape[0]], owner[:n]

    def Fuse(self, keys, desc, bounds, scale_factors, inv_level_sigma2, frame_points, th=3.0, u_right=None, bf=0.0,
             check_reprojection=True):
        """The search of ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) (ORBmatcher.cc:1155-1345; check_reprojection=True) or of
        ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1347-1462; False).  The caller applies the map mutations
        in point order.  -> (count, best_idx, best_dist)"""
        from .scenes import FRAME_POINT_DTYPE
        k = np.ascontiguousarray(keys, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = k.shape[0]
        ur = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        p = np.ascontiguousarray(frame_points, FRAME_POINT_DTYPE)
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        iv = np.ascontiguousarray(inv_level_sigma2, np.float32)
        bi, bd = np.full(max(p.shape[0], 1), -1, np.int32), np.full(max(p.shape[0], 1), -1, np.int32)
        nn, npp, nm = (np.array([v], np.int32) for v in (n, p.shape[0], 0))
        self._bf._check(_lib.lib().orbfe_fuse_batch(
            self._bf._h, ptr(k), ptr(d), ptr(ur), ptr(nn), n, ptr(b), ptr(sf), ptr(iv), sf.shape[0], float(bf), ptr(p), ptr(npp),
            p.shape[0], 1, float(th), int(bool(check_reprojection)), ptr(bi), ptr(bd), ptr(nm), 0))
        return int(nm[0]), bi[:p.shape[0]], bd[:p.shape[0]]

    def SearchBySim3(self, keys_un1, desc1, keys_un2, desc2, bounds1, bounds2, scale_factors, frame_points1, frame_points2, th):
        """ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, S12, th) (ORBmatcher.cc:1464-1681).  frame_points1[i1] = KF1's
        MapPoint i1 projected into KF2, frame_points2[i2] the reverse.  -> (nFound, matches12, vnMatch1, vnMatch2)"""
        from .scenes import FRAME_POINT_DTYPE
        k1, k2 = np.ascontiguousarray(keys_un1, KP_DTYPE), np.ascontiguousarray(keys_un2, KP_DTYPE)
        n1, n2 = k1.shape[0], k2.shape[0]
        cap = max(n1, n2, 1)
        K = np.zeros((2, cap), KP_DTYPE)
        D = np.zeros((2, cap, 32), np.uint8)
        P = np.zeros((2, cap), FRAME_POINT_DTYPE)
        K[0, :n1], K[1, :n2] = k1, k2
        D[0, :n1] = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        D[1, :n2] = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        P[0, :n1], P[1, :n2] = frame_points1, frame_points2
        b1, b2 = np.ascontiguousarray(bounds1, np.float32), np.ascontiguousarray(bounds2, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        m12, v1, v2 = (np.full(cap, -1, np.int32) for _ in range(3))
        nn1, nn2, nf = (np.array([v], np.int32) for v in (n1, n2, 0))
        self._bf._check(_lib.lib().orbfe_search_by_sim3_batch(
            self._bf._h, ptr(K[0]), ptr(D[0]), ptr(nn1), ptr(K[1]), ptr(D[1]), ptr(nn2), cap, ptr(b1), ptr(b2), ptr(sf),
            sf.shape[0], ptr(P[0]), ptr(P[1]), 1, float(th), ptr(m12), ptr(v1), ptr(v2), ptr(nf), 0))
        return int(nf[0]), m12[:n1], v1[:n1], v2[:n2]

    def search_sim3_projection_batch_raw(self, keys_un, desc, occupied, n, cap, bounds, scale_factors, pts, npts, pts_cap, nframes,
                                         th, ratio_hamming, match, owner, nmatches, flags=0):
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        self._bf._check(_lib.lib().orbfe_search_by_projection_sim3_batch(
            self._bf._h, ptr(keys_un), ptr(desc), ptr(occupied), ptr(n), int(cap), ptr(b), ptr(sf), sf.shape[0], ptr(pts),
            ptr(npts), int(pts_cap), int(nframes), float(th), float(ratio_hamming), ptr(match), ptr(owner), ptr(nmatches),
            int(flags)))

    def fuse_batch_raw(self, keys, desc, u_right, n, cap, bounds, scale_factors, inv_level_sigma2, bf, pts, npts, pts_cap, nframes,
                       th, check_reprojection, best_idx, best_dist, ncandidates, flags=0):
        b = np.ascontiguousarray(bounds, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        iv = np.ascontiguousarray(inv_level_sigma2, np.float32)
        self._bf._check(_lib.lib().orbfe_fuse_batch(
            self._bf._h, ptr(keys), ptr(desc), ptr(u_right), ptr(n), int(cap), ptr(b), ptr(sf), ptr(iv), sf.shape[0], float(bf),
            ptr(pts), ptr(npts), int(pts_cap), int(nframes), float(th), int(bool(check_reprojection)), ptr(best_idx),
            ptr(best_dist), ptr(ncandidates), int(flags)))

    def search_by_sim3_batch_raw(self, keys1, desc1, n1, keys2, desc2, n2, cap, bounds1, bounds2, scale_factors, pts1, pts2,
                                 nframes, th, matches12, match1, match2, nfound, flags=0):
        b1, b2 = np.ascontiguousarray(bounds1, np.float32), np.ascontiguousarray(bounds2, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        self._bf._check(_lib.lib().orbfe_search_by_sim3_batch(
            self._bf._h, ptr(keys1), ptr(desc1), ptr(n1), ptr(keys2), ptr(desc2), ptr(n2), int(cap), ptr(b1), ptr(b2), ptr(sf),
            sf.shape[0], ptr(pts1), ptr(pts2), int(nframes), float(th), ptr(matches12), ptr(match1), ptr(match2), ptr(nfound),
            int(flags)))

    def SearchByBoW(self, desc_kf, angle_kf, valid_kf, fv_kf, desc_f, angle_f, fv_f, nleft_f=-1):
        """ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vpMapPointMatches) (ORBmatcher.cc:224-425).  fv_* = the DBoW2
        feature vector as CSR (node_ids, start, idx); valid_kf[i] = the key frame's MapPoint i exists and is not bad.
        -> (nmatches, matches int32 (nF,): key-frame feature behind vpMapPointMatches[i] or -1)"""
        dk = np.ascontiguousarray(desc_kf, np.uint8).reshape(-1, 32)
        df = np.ascontiguousarray(desc_f, np.uint8).reshape(-1, 32)
        nk, nf = len(dk), len(df)
        cap = max(nk, nf, 1)
        ncap = max(len(fv_kf[0]), len(fv_f[0]), 1)

        def pad(a, n, dt, fill=0):
            out = np.full((n,) + np.asarray(a).shape[1:], fill, dt)
            out[:len(a)] = a
            return out
        args = [pad(dk, cap, np.uint8), pad(angle_kf, cap, np.float32), pad(valid_kf, cap, np.uint8),
                pad(fv_kf[0], ncap, np.int32), pad(fv_kf[1], ncap + 1, np.int32), pad(fv_kf[2], cap, np.int32),
                np.array([len(fv_kf[0])], np.int32), pad(df, cap, np.uint8), pad(angle_f, cap, np.float32),
                np.array([nf], np.int32), np.array([nleft_f], np.int32), pad(fv_f[0], ncap, np.int32),
                pad(fv_f[1], ncap + 1, np.int32), pad(fv_f[2], cap, np.int32), np.array([len(fv_f[0])], np.int32)]
        m, nm = np.full(cap, -1, np.int32), np.zeros(1, np.int32)
        self.search_by_bow_batch_raw(*args, cap, ncap, 1, m, nm)
        return int(nm[0]), m[:nf]

    def search_by_bow_batch_raw(self, desc_kf, angle_kf, valid_kf, kf_ids, kf_start, kf_idx, kf_nnodes, desc_f, angle_f, n_f,
                                nleft_f, f_ids, f_start, f_idx, f_nnodes, cap, node_cap, npairs, matches, nmatches, flags=0):
        self._bf._check(_lib.lib().orbfe_search_by_bow_batch(
            self._bf._h, ptr(desc_kf), ptr(angle_kf), ptr(valid_kf), ptr(kf_ids), ptr(kf_start), ptr(kf_idx), ptr(kf_nnodes),
            ptr(desc_f), ptr(angle_f), ptr(n_f), ptr(nleft_f), ptr(f_ids), ptr(f_start), ptr(f_idx), ptr(f_nnodes), int(cap),
            int(node_cap), int(npairs), self.mfNNratio, int(self.mbCheckOrientation), ptr(matches), ptr(nmatches), int(flags)))

    @staticmethod
    def _pad(a, n, dt, fill=0):
        a = np.asarray(a)
        out = np.full((n,) + a.shape[1:], fill, dt)
        out[:len(a)] = a
        return out

    def SearchByBoWKeyFrames(self, desc1, angle1, valid1, fv1, desc2, angle2, valid2, fv2):
        """ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vpMatches12) (ORBmatcher.cc:766-906).  validN[i] = MapPoint i
        exists, is not bad (and i < mvKeysUn.size() for stereo-camera key frames).  -> (nmatches, matches12 int32 (n1,))"""
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        n1 = len(d1)
        cap = max(n1, len(d2), 1)
        ncap = max(len(fv1[0]), len(fv2[0]), 1)
        pad = self._pad
        args = [pad(d1, cap, np.uint8), pad(angle1, cap, np.float32), pad(valid1, cap, np.uint8), pad(fv1[0], ncap, np.int32),
                pad(fv1[1], ncap + 1, np.int32), pad(fv1[2], cap, np.int32), np.array([len(fv1[0])], np.int32),
                pad(d2, cap, np.uint8), pad(angle2, cap, np.float32), pad(valid2, cap, np.uint8), pad(fv2[0], ncap, np.int32),
                pad(fv2[1], ncap + 1, np.int32), pad(fv2[2], cap, np.int32), np.array([len(fv2[0])], np.int32)]
        m, nm = np.full(cap, -1, np.int32), np.zeros(1, np.int32)
        self.search_by_bow_kf_batch_raw(*args, cap, ncap, 1, m, nm)
        return int(nm[0]), m[:n1]

    def search_by_bow_kf_batch_raw(self, desc1, angle1, valid1, ids1, start1, idx1, nnodes1, desc2, angle2, valid2, ids2, start2,
                                   idx2, nnodes2, cap, node_cap, npairs, matches12, nmatches, flags=0):
        self._bf._check(_lib.lib().orbfe_search_by_bow_kf_batch(
            self._bf._h, ptr(desc1), ptr(angle1), ptr(valid1), ptr(ids1), ptr(start1), ptr(idx1), ptr(nnodes1), ptr(desc2),
            ptr(angle2), ptr(valid2), ptr(ids2), ptr(start2), ptr(idx2), ptr(nnodes2), int(cap), int(node_cap), int(npairs),
            self.mfNNratio, int(self.mbCheckOrientation), ptr(matches12), ptr(nmatches), int(flags)))

    def SearchForTriangulation(self, keys1, desc1, free1, fv1, keys2, desc2, free2, fv2, f12, epipole, scale_factors,
                               level_sigma2, stereo1=None, stereo2=None, bCoarse=False):
        """ORBmatcher::SearchForTriangulation(pKF1, pKF2, vMatchedPairs, bOnlyStereo, bCoarse) (ORBmatcher.cc:908-1153) for key
        frames with one camera.  freeN[i] = no MapPoint yet (and a stereo measurement when bOnlyStereo); f12 = F12 (3 x 3).
        -> (nmatches, vMatchedPairs int32 (k, 2), vMatches12 int32 (n1,))"""
        k1, k2 = np.ascontiguousarray(keys1, KP_DTYPE), np.ascontiguousarray(keys2, KP_DTYPE)
        n1 = len(k1)
        cap = max(n1, len(k2), 1)
        ncap = max(len(fv1[0]), len(fv2[0]), 1)
        pad = self._pad
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        args = [pad(k1, cap, KP_DTYPE), pad(d1, cap, np.uint8), pad(free1, cap, np.uint8),
                None if stereo1 is None else pad(stereo1, cap, np.uint8), pad(fv1[0], ncap, np.int32),
                pad(fv1[1], ncap + 1, np.int32), pad(fv1[2], cap, np.int32), np.array([len(fv1[0])], np.int32),
                pad(k2, cap, KP_DTYPE), pad(d2, cap, np.uint8), pad(free2, cap, np.uint8),
                None if stereo2 is None else pad(stereo2, cap, np.uint8), pad(fv2[0], ncap, np.int32),
                pad(fv2[1], ncap + 1, np.int32), pad(fv2[2], cap, np.int32), np.array([len(fv2[0])], np.int32)]
        m, nm = np.full(cap, -1, np.int32), np.zeros(1, np.int32)
        self.search_for_triangulation_batch_raw(*args, cap, ncap, 1, np.ascontiguousarray(f12, np.float32).reshape(1, 9),
                                                np.ascontiguousarray(epipole, np.float32).reshape(1, 2), scale_factors,
                                                level_sigma2, bCoarse, m, nm)
        i1 = np.nonzero(m[:n1] >= 0)[0]
        return int(nm[0]), np.stack([i1, m[i1]], axis=1).astype(np.int32), m[:n1]

    def search_for_triangulation_batch_raw(self, keys1, desc1, free1, stereo1, ids1, start1, idx1, nnodes1, keys2, desc2, free2,
                                           stereo2, ids2, start2, idx2, nnodes2, cap, node_cap, npairs, f12, epipole,
                                           scale_factors, level_sigma2, coarse, matches12, nmatches, flags=0):
        sf = np.ascontiguousarray(scale_factors, np.float32)
        sg = np.ascontiguousarray(level_sigma2, np.float32)
        self._bf._check(_lib.lib().orbfe_search_for_triangulation_batch(
            self._bf._h, ptr(keys1), ptr(desc1), ptr(free1), ptr(stereo1), ptr(ids1), ptr(start1), ptr(idx1), ptr(nnodes1),
            ptr(keys2), ptr(desc2), ptr(free2), ptr(stereo2), ptr(ids2), ptr(start2), ptr(idx2), ptr(nnodes2), int(cap),
            int(node_cap), int(npairs), ptr(f12), ptr(epipole), ptr(sf), ptr(sg), sf.shape[0], int(bool(coarse)),
            int(self.mbCheckOrientation), ptr(matches12), ptr(nmatches), int(flags)))

    def ComputeDistinctiveDescriptors(self, desc, offsets, max_n=None, out_desc=None, flags=0):
        """MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401) for a batch of MapPoints: point p owns the rows
        [offsets[p], offsets[p + 1]) of desc (its observations' descriptors in the reference's gathering order).
        -> (best_idx int32 (npoints,), best_median int32 (npoints,), descriptors uint8 (npoints, 32))"""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        off = np.ascontiguousarray(offsets, np.int32)
        npts = len(off) - 1
        if max_n is None:
            max_n = max(int(np.diff(off).max()) if npts > 0 else 1, 1)
        bi, bm = np.full(max(npts, 1), -1, np.int32), np.full(max(npts, 1), -1, np.int32)
        od = np.zeros((max(npts, 1), 32), np.uint8) if out_desc is None else out_desc
        self._bf._check(_lib.lib().orbfe_distinctive_descriptors(self._bf._h, ptr(d), ptr(off), npts, len(d), int(max_n), ptr(bi),
                                                                 ptr(bm), ptr(od), int(flags)))
        return bi[:npts], bm[:npts], od[:npts]

    def SearchForTriangulationKB8(self, keys1, desc1, free1, nleft1, fv1, keys2, desc2, free2, nleft2, fv2, rig, t12,
                                  level_sigma2, bCoarse=False):
        """ORBmatcher::SearchForTriangulation for key frames of a two-camera (KannalaBrandt8) rig (ORBmatcher.cc:908-1153 with
        :1014-1043): keysN / descN = left then right features, nleftN = NLeft, rig.cam1 / cam2 = mpCamera / mpCamera2, t12 =
        (4, 12): R12 | t12 of Tll, Tlr, Trl, Trr.  -> (nmatches, vMatchedPairs int32 (k, 2), vMatches12 int32 (n1,))"""
        from .scenes import FISHEYE_RIG_DTYPE
        k1, k2 = np.ascontiguousarray(keys1, KP_DTYPE), np.ascontiguousarray(keys2, KP_DTYPE)
        n1 = len(k1)
        cap = max(n1, len(k2), 1)
        ncap = max(len(fv1[0]), len(fv2[0]), 1)
        pad = self._pad
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        rg = np.ascontiguousarray(rig, FISHEYE_RIG_DTYPE)
        T = np.ascontiguousarray(t12, np.float32).reshape(1, 48)
        sg = np.ascontiguousarray(level_sigma2, np.float32)
        m, nm = np.full(cap, -1, np.int32), np.zeros(1, np.int32)
        self._bf._check(_lib.lib().orbfe_search_for_triangulation_kb8_batch(
            self._bf._h, ptr(pad(k1, cap, KP_DTYPE)), ptr(pad(d1, cap, np.uint8)), ptr(pad(free1, cap, np.uint8)),
            ptr(np.array([nleft1], np.int32)), ptr(pad(fv1[0], ncap, np.int32)), ptr(pad(fv1[1], ncap + 1, np.int32)),
            ptr(pad(fv1[2], cap, np.int32)), ptr(np.array([len(fv1[0])], np.int32)), ptr(pad(k2, cap, KP_DTYPE)),
            ptr(pad(d2, cap, np.uint8)), ptr(pad(free2, cap, np.uint8)), ptr(np.array([nleft2], np.int32)),
            ptr(pad(fv2[0], ncap, np.int32)), ptr(pad(fv2[1], ncap + 1, np.int32)), ptr(pad(fv2[2], cap, np.int32)),
            ptr(np.array([len(fv2[0])], np.int32)), cap, ncap, 1, ptr(rg), ptr(T), ptr(sg), sg.shape[0], int(bool(bCoarse)),
            int(self.mbCheckOrientation), ptr(m), ptr(nm), 0))
        i1 = np.nonzero(m[:n1] >= 0)[0]
        return int(nm[0]), np.stack([i1, m[i1]], axis=1).astype(np.int32), m[:n1]

    def SearchForInitialization(self, keys1_un, desc1, keys2_un, desc2, bounds, vbPrevMatched, windowSize=10):
        """ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (ORBmatcher.cc:649-764) on
        monocular frames given as arrays; bounds = F2's image bounds.  vbPrevMatched: (n1, 2) float32.
        -> (nmatches, vnMatches12 int32 (n1,), updated vbPrevMatched (n1, 2))"""
        k1 = np.ascontiguousarray(keys1_un, KP_DTYPE)
        k2 = np.ascontiguousarray(keys2_un, KP_DTYPE)
        n1, n2 = k1.shape[0], k2.shape[0]
        cap = max(n1, n2, 1)
        K = np.zeros((2, cap), KP_DTYPE)
        D = np.zeros((2, cap, 32), np.uint8)
        K[0, :n1], K[1, :n2] = k1, k2
        D[0, :n1] = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        D[1, :n2] = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        pm = np.zeros((cap, 2), np.float32)
        pm[:n1] = np.asarray(vbPrevMatched, np.float32).reshape(-1, 2)
        m12 = np.full(cap, -1, np.int32)
        nn1, nn2, nm = (np.array([v], np.int32) for v in (n1, n2, 0))
        self.search_for_initialization_batch_raw(K[0], D[0], nn1, K[1], D[1], nn2, cap, 1, bounds, windowSize, pm, m12, nm)
        return int(nm[0]), m12[:n1], pm[:n1]

    def search_for_initialization_batch_raw(self, keys1, desc1, n1, keys2, desc2, n2, cap, nframes, bounds, windowSize,
                                            prev_matched, matches12, nmatches, flags=0):
        b = np.ascontiguousarray(bounds, np.float32)
        self._bf._check(_lib.lib().orbfe_search_for_initialization_batch(
            self._bf._h, ptr(keys1), ptr(desc1), ptr(n1), ptr(keys2), ptr(desc2), ptr(n2), int(cap), int(nframes), ptr(b),
            int(windowSize), self.mfNNratio, int(self.mbCheckOrientation), ptr(prev_matched), ptr(matches12), ptr(nmatches),
            int(flags)))

    def ComputeStereoFishEyeMatches(self, keys_left, mono_left, keys_right, mono_right, knn_idx, knn_dist, rig):
        """Frame::ComputeStereoFishEyeMatches after BFMatchORB (Frame.cc:1164-1204): knn_idx / knn_dist = knnMatch of the
        stereo-area left descriptors against the stereo-area right ones; rig: scenes.FISHEYE_RIG_DTYPE.
        -> (nMatches, mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, mvStereo3Dpoints (Nleft, 3), status)"""
        kl = np.ascontiguousarray(keys_left, KP_DTYPE)
        kr = np.ascontiguousarray(keys_right, KP_DTYPE)
        nl, nr = kl.shape[0], kr.shape[0]
        cap = max(nl, nr, 1)
        K = np.zeros((2, cap), KP_DTYPE)
        K[0, :nl], K[1, :nr] = kl, kr
        ki = np.ascontiguousarray(knn_idx, np.int32).reshape(-1, 2)
        kd = np.ascontiguousarray(knn_dist, np.int32).reshape(-1, 2)
        a = [np.array([v], np.int32) for v in (nl, nr, mono_left, mono_right, 0)]
        ltr, rtl, st = (np.empty(cap, np.int32) for _ in range(3))
        depth, p3d = np.empty(cap, np.float32), np.empty((cap, 3), np.float32)
        self.stereo_fisheye_matches_batch_raw(K[0], K[1], a[0], a[1], a[2], a[3], cap, ki, kd, ki.shape[0], 1, rig, ltr, rtl,
                                              depth, p3d, st, a[4])
        return int(a[4][0]), ltr[:nl], rtl[:nr], depth[:nl], p3d[:nl], st[:nl]

    def stereo_fisheye_matches_batch_raw(self, keys_left, keys_right, n_left, n_right, mono_left, mono_right, cap, knn_idx,
                                         knn_dist, knn_cap, nframes, rig, ltr, rtl, depth, p3d, status, nmatches, flags=0):
        from .scenes import FISHEYE_RIG_DTYPE
        rg = np.ascontiguousarray(rig, FISHEYE_RIG_DTYPE)
        self._bf._check(_lib.lib().orbfe_stereo_fisheye_matches_batch(
            self._bf._h, ptr(keys_left), ptr(keys_right), ptr(n_left), ptr(n_right), ptr(mono_left), ptr(mono_right),
            int(cap), ptr(knn_idx), ptr(knn_dist), int(knn_cap), int(nframes), ptr(rg), ptr(ltr), ptr(rtl), ptr(depth),
            ptr(p3d), ptr(status), ptr(nmatches), int(flags)))

    def UndistortKeyPoints(self, keys, K, dist_coef):
        """Frame::UndistortKeyPoints (Frame.cc:763-796): keys (KP_DTYPE), K = 3x3 mK or (fx, fy, cx, cy), mDistCoef ->
        mvKeysUn."""
        k = np.ascontiguousarray(keys, KP_DTYPE)
        K = np.asarray(K, np.float32)
        fx, fy, cx, cy = (K[0, 0], K[1, 1], K[0, 2], K[1, 2]) if K.ndim == 2 else K
        d = np.ascontiguousarray(dist_coef, np.float32).reshape(-1)
        out = np.empty_like(k)
        self._bf._check(_lib.lib().orbfe_undistort_keypoints(self._bf._h, ptr(k), k.shape[0], float(fx), float(fy),
                                                             float(cx), float(cy), ptr(d), d.shape[0], ptr(out), 0))
        return out

    def last_ms(self):
        return self._bf.last_ms()
