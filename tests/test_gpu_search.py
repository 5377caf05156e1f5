"""-m gpu: the frame grid + ORBmatcher::SearchByProjection (ORBmatcher.cc:44-143, Nleft == -1) on the GPU
(orbfe_search_by_projection) against the CPU oracle: the feature assigned to every map point, the final owner of every
feature and the match count, including the order-dependent cases (ties, runner-up level, a feature taken by an earlier
point).  Integer / index work: bit-exact."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBextractor, ORBmatcher, _lib, scenes

pytestmark = pytest.mark.gpu


def _frame(oracle, kind, w, h, seed, nf):
    img = scenes.make(kind, w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kps, desc = ex(img)
    u_right = np.where(scenes.rand_u32(len(kps), seed) % 3 == 0, -1.0, kps["x"] - 9.25).astype(np.float32)
    return kps, desc, u_right, ex.scale


@pytest.mark.parametrize("kind,w,h,seed,nf,npts,th,ratio", [
    ("texture", 752, 480, 8000, 1200, 2000, 1.0, 0.8),
    ("texture", 752, 480, 8001, 1200, 3000, 3.0, 0.8),
    ("texture", 640, 480, 8002, 1000, 1500, 15.0, 0.9),     # relocalisation-sized windows: hundreds of candidates
    ("checker", 640, 480, 8003, 1000, 2000, 1.0, 0.6),       # repeated structure: distance ties
    ("noise", 320, 240, 8004, 500, 700, 2.0, 0.8),
    ("flat_blobs", 752, 480, 8005, 1000, 300, 1.0, 0.8),
])
def test_search_by_projection_matches_oracle(oracle, kind, w, h, seed, nf, npts, th, ratio):
    kps, desc, u_right, scale = _frame(oracle, kind, w, h, seed, nf)
    pts = scenes.track_points(kps, desc, u_right, 8, npts, seed, dup_frac=0.3)
    occupied = (scenes.rand_u32(len(kps), seed + 9) % 10 == 0).astype(np.uint8)
    bounds = (0.0, 0.0, float(w), float(h))
    rm, ro, rn = oracle.search_by_projection(kps, desc, u_right, occupied, bounds, scale, pts, th=th, nn_ratio=ratio)
    m = ORBmatcher(ratio)
    n, match, owner = m.SearchByProjection(kps, desc, u_right, bounds, scale, pts, th=th, occupied=occupied)
    np.testing.assert_array_equal(match, rm)
    np.testing.assert_array_equal(owner, ro)
    assert n == rn
    if kind == "texture":
        assert n > npts // 4


def test_search_monocular_far_points_and_undistorted_bounds(oracle):
    """no mvuRight (monocular), bFarPoints gate, image bounds that do not start at 0 (undistorted corners)."""
    kps, desc, _, scale = _frame(oracle, "texture", 640, 480, 8010, 1000)
    pts = scenes.track_points(kps, desc, None, 8, 1500, 8010)
    bounds = (-12.5, -7.25, 655.0, 489.5)
    rm, ro, rn = oracle.search_by_projection(kps, desc, None, None, bounds, scale, pts, th=1.0, far_points=True,
                                             th_far=20.0, nn_ratio=0.8)
    n, match, owner = ORBmatcher(0.8).SearchByProjection(kps, desc, None, bounds, scale, pts, th=1.0, bFarPoints=True,
                                                         thFarPoints=20.0)
    np.testing.assert_array_equal(match, rm)
    np.testing.assert_array_equal(owner, ro)
    assert n == rn and (match[pts["track_depth"] > 20.0] == -1).all()


def test_search_contended_feature_is_sequential(oracle):
    """many points aimed at the SAME few features: every round of the commit loop has conflicts."""
    kps, desc, u_right, scale = _frame(oracle, "texture", 320, 240, 8020, 400)
    base = scenes.track_points(kps, desc, u_right, 8, 40, 8020, dup_frac=0.0)
    pts = np.repeat(base, 16)                       # 16 consecutive copies of each point
    pts["flags"] = np.where(np.arange(len(pts)) % 5 == 0, 1, 5)    # every 5th copy has no observations (does not block)
    bounds = (0.0, 0.0, 320.0, 240.0)
    rm, ro, rn = oracle.search_by_projection(kps, desc, u_right, None, bounds, scale, pts, th=4.0, nn_ratio=0.9)
    n, match, owner = ORBmatcher(0.9).SearchByProjection(kps, desc, u_right, bounds, scale, pts, th=4.0)
    np.testing.assert_array_equal(match, rm)
    np.testing.assert_array_equal(owner, ro)
    assert n == rn


def test_search_batch_device_and_edges(oracle):
    import ctypes as C
    m = ORBmatcher(0.8)
    # empty inputs
    kps, desc, u_right, scale = _frame(oracle, "texture", 320, 240, 8030, 300)
    n, match, owner = m.SearchByProjection(kps, desc, u_right, (0, 0, 320, 240), scale, np.zeros(0, scenes.TRACK_DTYPE))
    assert n == 0 and len(match) == 0 and (owner == -1).all()
    n, match, owner = m.SearchByProjection(kps[:0], desc[:0], None, (0, 0, 320, 240), scale,
                                           scenes.track_points(kps, desc, u_right, 8, 50, 1))
    assert n == 0 and (match == -1).all()
    with pytest.raises(_lib.OrbfeError):
        m.SearchByProjection(kps, desc, u_right, (0, 0, 0, 240), scale, np.zeros(1, scenes.TRACK_DTYPE))
    # a batch of 5 frames with ragged counts equals the single-frame calls
    nfr, cap, pcap = 5, len(kps), 600
    K = np.zeros((nfr, cap), _lib.KP_DTYPE)
    D = np.zeros((nfr, cap, 32), np.uint8)
    U = np.full((nfr, cap), -1, np.float32)
    P = np.zeros((nfr, pcap), scenes.TRACK_DTYPE)
    nn = np.array([cap, cap - 17, 0, cap // 2, cap], np.int32)
    npp = np.array([600, 333, 100, 0, 599], np.int32)
    singles = []
    for f in range(nfr):
        K[f, :nn[f]], D[f, :nn[f]], U[f, :nn[f]] = kps[:nn[f]], desc[:nn[f]], u_right[:nn[f]]
        if nn[f] and npp[f]:
            P[f, :npp[f]] = scenes.track_points(kps[:nn[f]], desc[:nn[f]], u_right[:nn[f]], 8, int(npp[f]), 90 + f)
        singles.append(m.SearchByProjection(K[f, :nn[f]], D[f, :nn[f]], U[f, :nn[f]], (0, 0, 320, 240), scale,
                                            P[f, :npp[f]], th=2.0))
    match = np.zeros((nfr, pcap), np.int32)
    owner = np.zeros((nfr, cap), np.int32)
    cnt = np.zeros(nfr, np.int32)
    m.search_by_projection_batch_raw(K, D, U, None, nn, cap, (0, 0, 320, 240), scale, P, npp, pcap, nfr, 2.0, match, owner,
                                     cnt)
    for f in range(nfr):
        assert cnt[f] == singles[f][0]
        np.testing.assert_array_equal(match[f, :npp[f]], singles[f][1])
        assert (match[f, npp[f]:] == -1).all()
        np.testing.assert_array_equal(owner[f, :nn[f]], singles[f][2])


def _stereo_frame(oracle, kind, w, h, seed, nf):
    left = scenes.make(kind, w, h, seed)
    right = scenes.stereo_right_rectified(left, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kl, dl = ex(left)
    _, kr, dr = ex(right)
    return scenes.stereo_camera_frame(kl, dl, kr, dr, seed) + (ex.scale,)


@pytest.mark.parametrize("kind,w,h,seed,nf,npts,th,ratio", [
    ("texture", 752, 480, 8100, 1200, 2500, 1.0, 0.8),
    ("texture", 640, 480, 8101, 1000, 2000, 3.0, 0.8),
    ("texture", 640, 400, 8102, 1000, 1500, 10.0, 0.9),
    ("checker", 640, 480, 8103, 1000, 2000, 1.0, 0.6),
    ("noise", 320, 240, 8104, 500, 800, 2.0, 0.8),
])
def test_search_by_projection_stereo_camera_matches_oracle(oracle, kind, w, h, seed, nf, npts, th, ratio):
    """Nleft != -1 (ORBmatcher.cc:44-214): left and right searches, partner assignments, the continue past the right
    search, slots freed by points without observations."""
    keys, desc, nleft, ltr, rtl, scale = _stereo_frame(oracle, kind, w, h, seed, nf)
    pts = scenes.track_points_stereo(keys, desc, nleft, ltr, 8, npts, seed, dup_frac=0.3)
    occupied = (scenes.rand_u32(len(keys), seed + 9) % 8 == 0).astype(np.uint8)
    bounds = (0.0, 0.0, float(w), float(h))
    rl, rr, ro, rn = oracle.search_by_projection_stereo(keys, desc, occupied, nleft, ltr, rtl, bounds, scale, pts, th=th,
                                                        nn_ratio=ratio)
    n, ml, mr, owner = ORBmatcher(ratio).SearchByProjectionStereo(keys, desc, nleft, ltr, rtl, bounds, scale, pts, th=th,
                                                                  occupied=occupied)
    np.testing.assert_array_equal(ml, rl)
    np.testing.assert_array_equal(mr, rr)
    np.testing.assert_array_equal(owner, ro)
    assert n == rn
    if kind == "texture":
        assert (ml >= 0).sum() > npts // 5 and (mr >= 0).sum() > 20


def test_search_stereo_camera_contended(oracle):
    """16 copies of every point, one in five without observations: conflicts in every round, slots freed and retaken."""
    keys, desc, nleft, ltr, rtl, scale = _stereo_frame(oracle, "texture", 320, 240, 8120, 400)
    base = scenes.track_points_stereo(keys, desc, nleft, ltr, 8, 40, 8120, dup_frac=0.0)
    pts = np.repeat(base, 16)
    pts["flags"] = np.where(np.arange(len(pts)) % 5 == 0, 9, 13)
    bounds = (0.0, 0.0, 320.0, 240.0)
    rl, rr, ro, rn = oracle.search_by_projection_stereo(keys, desc, None, nleft, ltr, rtl, bounds, scale, pts, th=4.0,
                                                        nn_ratio=0.9)
    n, ml, mr, owner = ORBmatcher(0.9).SearchByProjectionStereo(keys, desc, nleft, ltr, rtl, bounds, scale, pts, th=4.0)
    np.testing.assert_array_equal(ml, rl)
    np.testing.assert_array_equal(mr, rr)
    np.testing.assert_array_equal(owner, ro)
    assert n == rn


def test_search_stereo_camera_edges(oracle):
    keys, desc, nleft, ltr, rtl, scale = _stereo_frame(oracle, "texture", 320, 240, 8130, 300)
    m = ORBmatcher(0.8)
    n, ml, mr, owner = m.SearchByProjectionStereo(keys, desc, nleft, ltr, rtl, (0, 0, 320, 240), scale,
                                                  np.zeros(0, scenes.TRACK_STEREO_DTYPE))
    assert n == 0 and len(ml) == 0 and (owner == -1).all()
    # no right features at all (Nleft == N): the right search finds nothing, the left one still works
    pts = scenes.track_points_stereo(keys[:nleft], desc[:nleft], nleft, np.full(nleft, -1, np.int32), 8, 300, 5)
    none_r = np.zeros(0, np.int32)
    rl, rr, ro, rn = oracle.search_by_projection_stereo(keys[:nleft], desc[:nleft], None, nleft, np.full(nleft, -1, np.int32),
                                                        none_r, (0, 0, 320, 240), scale, pts, th=2.0, nn_ratio=0.8)
    n, ml, mr, owner = m.SearchByProjectionStereo(keys[:nleft], desc[:nleft], nleft, np.full(nleft, -1, np.int32), none_r,
                                                  (0, 0, 320, 240), scale, pts, th=2.0)
    np.testing.assert_array_equal(ml, rl)
    assert (mr == -1).all() and (rr == -1).all() and n == rn
    np.testing.assert_array_equal(owner, ro)
