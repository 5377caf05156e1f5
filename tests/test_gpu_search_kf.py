"""-m gpu: MapPoints projected into a key frame on the GPU (k_search_kf.cu) against the CPU oracle:
ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) (ORBmatcher.cc:428-647),
ORBmatcher::Fuse (:1155-1345, :1347-1462) and ORBmatcher::SearchBySim3 (:1464-1681).  Index work: bit-exact."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBmatcher, _lib, scenes

pytestmark = pytest.mark.gpu


def _key_frame(oracle, kind, w, h, seed, nf):
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kps, desc = ex(scenes.make(kind, w, h, seed))
    return kps, desc, ex.scale.astype(np.float32)


def _points(kps, desc, npts, seed, w, h, dup_frac=0.3):
    pts = scenes.frame_points(kps, desc, len(kps), None, None, 8, npts, seed, dup_frac=dup_frac)
    if npts >= 8:    # projections exactly on the bounds: KeyFrame::IsInImage is >= min and < max
        pts["u"][3], pts["v"][5], pts["u"][6], pts["v"][7] = w, h, 0.0, 0.0
        pts["flags"][[3, 5, 6, 7]] |= 1
    return pts


@pytest.mark.parametrize("kind,w,h,seed,nf,npts,th,ratio", [
    ("texture", 752, 480, 9100, 1200, 1500, 10.0, 1.0),     # LoopClosing: th = 10 / 8 / 5 / 3 with ratioHamming 1.0 .. 2.0
    ("texture", 752, 480, 9101, 1200, 2500, 8.0, 1.5),
    ("texture", 640, 480, 9102, 1000, 1000, 3.0, 2.0),
    ("checker", 640, 480, 9103, 1000, 1500, 5.0, 1.0),      # repeated structure: distance ties
    ("noise", 320, 240, 9104, 500, 700, 10.0, 0.6),
])
def test_sim3_projection_matches_oracle(oracle, kind, w, h, seed, nf, npts, th, ratio):
    kps, desc, scale = _key_frame(oracle, kind, w, h, seed, nf)
    n = len(kps)
    pts = _points(kps, desc, npts, seed, w, h)
    occupied = (scenes.rand_u32(n, seed + 9) % 6 == 0).astype(np.uint8)
    bounds = (0.0, 0.0, float(w), float(h))
    rm, ro, rn = oracle.search_by_projection_sim3(kps, desc, occupied, bounds, scale, pts, th, ratio)
    nm, m, owner = ORBmatcher(0.75, True).SearchByProjectionSim3(kps, desc, bounds, scale, pts, th, ratio, occupied=occupied)
    np.testing.assert_array_equal(m, rm)
    np.testing.assert_array_equal(owner, ro)
    assert nm == rn
    assert m[3] == -1 and m[5] == -1
    if kind == "texture":
        assert nm > npts // 8
    got = m[m >= 0]
    assert len(np.unique(got)) == len(got) and not occupied[got].any()


def test_sim3_projection_contended_is_sequential(oracle):
    """many points aimed at the same few features: every commit round has conflicts; no occupied array."""
    kps, desc, scale = _key_frame(oracle, "texture", 320, 240, 9200, 400)
    base = scenes.frame_points(kps[:40], desc[:40], 40, None, None, 8, 60, 9200, dup_frac=0.0)
    pts = np.repeat(base, 16)
    bounds = (0.0, 0.0, 320.0, 240.0)
    rm, ro, rn = oracle.search_by_projection_sim3(kps, desc, None, bounds, scale, pts, 10.0, 2.0)
    nm, m, owner = ORBmatcher(0.75, True).SearchByProjectionSim3(kps, desc, bounds, scale, pts, 10.0, 2.0)
    np.testing.assert_array_equal(m, rm)
    np.testing.assert_array_equal(owner, ro)
    assert nm == rn and nm > 40


@pytest.mark.parametrize("kind,w,h,seed,nf,npts,th,check", [
    ("texture", 752, 480, 9300, 1200, 2000, 3.0, True),     # LocalMapping::SearchInNeighbors: th = 3
    ("texture", 752, 480, 9301, 1200, 3000, 6.0, True),
    ("texture", 640, 480, 9302, 1000, 1500, 4.0, False),    # LoopClosing::SearchAndFuse: th = 4
    ("checker", 640, 480, 9303, 1000, 1500, 3.0, True),
])
def test_fuse_matches_oracle(oracle, kind, w, h, seed, nf, npts, th, check):
    kps, desc, scale = _key_frame(oracle, kind, w, h, seed, nf)
    n = len(kps)
    pts = _points(kps, desc, npts, seed, w, h)
    u_right = np.where(scenes.rand_u32(n, seed) % 3 == 0, -1.0, kps["x"] - 8.0).astype(np.float32)
    inv_sigma2 = (1.0 / (scale * scale)).astype(np.float32)
    bounds = (0.0, 0.0, float(w), float(h))
    ri, rd, rn = oracle.fuse(kps, desc, u_right, bounds, scale, inv_sigma2, 47.9, pts, th, check)
    cnt, bi, bd = ORBmatcher(0.6, True).Fuse(kps, desc, bounds, scale, inv_sigma2, pts, th, u_right=u_right, bf=47.9,
                                             check_reprojection=check)
    np.testing.assert_array_equal(bi, ri)
    np.testing.assert_array_equal(bd, rd)
    assert cnt == rn
    if kind == "texture":
        assert cnt > npts // 10
        assert len(np.unique(bi[bi >= 0])) < (bi >= 0).sum()    # duplicates share a feature: nothing blocks in Fuse


def test_fuse_without_uright_and_gate_effect(oracle):
    kps, desc, scale = _key_frame(oracle, "texture", 640, 480, 9400, 1000)
    pts = _points(kps, desc, 1500, 9400, 640, 480)
    inv_sigma2 = (1.0 / (scale * scale)).astype(np.float32)
    bounds = (0.0, 0.0, 640.0, 480.0)
    m = ORBmatcher(0.6, True)
    res = {}
    for check in (True, False):
        ri, rd, rn = oracle.fuse(kps, desc, None, bounds, scale, inv_sigma2, 0.0, pts, 6.0, check)
        cnt, bi, bd = m.Fuse(kps, desc, bounds, scale, inv_sigma2, pts, 6.0, check_reprojection=check)
        np.testing.assert_array_equal(bi, ri)
        np.testing.assert_array_equal(bd, rd)
        res[check] = cnt
    assert res[True] < res[False]


def _sim3_pair(oracle, seed, w, h, nf):
    cur = scenes.make("texture", w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k1, d1 = ex(cur)
    _, k2, d2 = ex(scenes.previous_frame(cur, seed))
    p1 = scenes.frame_points(k2, d2, len(k2), None, None, 8, len(k1), seed + 1, dup_frac=0.05)
    p2 = scenes.frame_points(k1, d1, len(k1), None, None, 8, len(k2), seed + 2, dup_frac=0.05)
    r = scenes.rand_u32(len(k1), seed + 3)
    for i1 in range(0, len(k1), 2):    # mutual pairs: point i1 looks at feature j of KF2 and point j looks back
        j = int(r[i1] % len(k2))
        p1[i1]["u"], p1[i1]["v"], p1[i1]["level"], p1[i1]["desc"], p1[i1]["flags"] = k2["x"][j], k2["y"][j], k2["octave"][j], d2[j], 1
        p2[j]["u"], p2[j]["v"], p2[j]["level"], p2[j]["desc"], p2[j]["flags"] = k1["x"][i1], k1["y"][i1], k1["octave"][i1], d1[i1], 1
    return k1, d1, k2, d2, ex.scale.astype(np.float32), p1, p2


@pytest.mark.parametrize("seed,w,h,nf,th", [(9500, 752, 480, 1200, 7.5), (9501, 640, 480, 1000, 3.0), (9502, 320, 240, 400, 7.5)])
def test_search_by_sim3_matches_oracle(oracle, seed, w, h, nf, th):
    k1, d1, k2, d2, scale, p1, p2 = _sim3_pair(oracle, seed, w, h, nf)
    b = (0.0, 0.0, float(w), float(h))
    r12, r1, r2, rn = oracle.search_by_sim3(k1, d1, k2, d2, b, b, scale, p1, p2, th)
    nf_, m12, v1, v2 = ORBmatcher(0.75, True).SearchBySim3(k1, d1, k2, d2, b, b, scale, p1, p2, th)
    np.testing.assert_array_equal(v1, r1)
    np.testing.assert_array_equal(v2, r2)
    np.testing.assert_array_equal(m12, r12)
    assert nf_ == rn and nf_ > len(k1) // 8 and (v1 >= 0).sum() > nf_


def test_kf_searches_ragged_device_batch(oracle):
    """batches of key frames with different feature / point counts (one empty), device-resident buffers."""
    import torch
    frames = [_key_frame(oracle, "texture", 640, 480, 9600 + k, nf) for k, nf in enumerate((1000, 300, 700))]
    cap, pcap, F = max(len(f[0]) for f in frames), 1100, 4
    K = np.zeros((F, cap), _lib.KP_DTYPE)
    D = np.zeros((F, cap, 32), np.uint8)
    U = np.full((F, cap), -1, np.float32)
    O = np.zeros((F, cap), np.uint8)
    P = np.zeros((F, pcap), scenes.FRAME_POINT_DTYPE)
    nn, npp = np.zeros(F, np.int32), np.array([1100, 0, 411, 50], np.int32)
    for f, (k, d, _) in enumerate(frames):
        nn[f] = len(k)
        K[f, :len(k)], D[f, :len(k)] = k, d
        U[f, :len(k)] = np.where(scenes.rand_u32(len(k), 9610 + f) % 2 == 0, -1.0, k["x"] - 6.0)
        O[f, :len(k)] = scenes.rand_u32(len(k), 9620 + f) % 5 == 0
        P[f, :npp[f]] = _points(k, d, int(npp[f]), 9630 + f, 640, 480)
    P[3, :50] = P[0, :50]    # frame 3 has points but no features
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1)).to(dev)    # noqa: E731
    dK, dD, dP, dO = t(K), t(D), t(P), t(O)
    dU, dn, dnp = (torch.from_numpy(a).to(dev) for a in (U, nn, npp))
    z = lambda *s: torch.zeros(s, dtype=torch.int32, device=dev)    # noqa: E731
    bounds, scale = (0.0, 0.0, 640.0, 480.0), frames[0][2]
    inv_sigma2 = (1.0 / (scale * scale)).astype(np.float32)
    m = ORBmatcher(0.75, True)
    flags = _lib.IN_DEVICE | _lib.OUT_DEVICE
    dm, down, dcnt = z(F, pcap), z(F, cap), z(F)
    dbi, dbd, dfc = z(F, pcap), z(F, pcap), z(F)
    torch.cuda.synchronize()    # the fills above run on torch's stream, the library on its own
    m.search_sim3_projection_batch_raw(dK, dD, dO, dn, cap, bounds, scale, dP, dnp, pcap, F, 8.0, 1.5, dm, down, dcnt, flags)
    m.fuse_batch_raw(dK, dD, dU, dn, cap, bounds, scale, inv_sigma2, 47.9, dP, dnp, pcap, F, 3.0, True, dbi, dbd, dfc, flags)
    mm, owner, cnt = dm.cpu().numpy(), down.cpu().numpy(), dcnt.cpu().numpy()
    bi, bd, fc = dbi.cpu().numpy(), dbd.cpu().numpy(), dfc.cpu().numpy()
    for f in range(F):
        n, npt = int(nn[f]), int(npp[f])
        rm, ro, rn = oracle.search_by_projection_sim3(K[f, :n], D[f, :n], O[f, :n], bounds, scale, P[f, :npt], 8.0, 1.5)
        np.testing.assert_array_equal(mm[f, :npt], rm)
        np.testing.assert_array_equal(owner[f, :n], ro)
        assert cnt[f] == rn
        ri, rd, rc = oracle.fuse(K[f, :n], D[f, :n], U[f, :n], bounds, scale, inv_sigma2, 47.9, P[f, :npt], 3.0, True)
        np.testing.assert_array_equal(bi[f, :npt], ri)
        np.testing.assert_array_equal(bd[f, :npt], rd)
        assert fc[f] == rc
    assert cnt[1] == 0 and cnt[3] == 0 and cnt[0] > 100
    assert (mm[np.arange(pcap)[None, :] >= npp[:, None]] == -1).all()    # slots beyond npts stay -1


def test_kf_search_argument_errors():
    m = ORBmatcher(0.75, True)
    k = np.zeros(4, _lib.KP_DTYPE)
    d = np.zeros((4, 32), np.uint8)
    p = np.zeros(2, scenes.FRAME_POINT_DTYPE)
    sf = np.ones(8, np.float32)
    with pytest.raises(Exception):
        m.SearchByProjectionSim3(k, d, (0, 0, 0, 10), sf, p, 10.0)          # empty bounds
    with pytest.raises(Exception):
        m.SearchByProjectionSim3(k, d, (0, 0, 10, 10), sf, p, 10.0, 6.0)    # TH_LOW * ratioHamming >= 256
    nm, mm, owner = m.SearchByProjectionSim3(k, d, (0, 0, 10, 10), sf, p[:0], 10.0)
    assert nm == 0 and len(mm) == 0
