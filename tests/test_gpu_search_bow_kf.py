"""-m gpu: the two key-frame <-> key-frame matchers that walk the DBoW2 feature vectors, on the GPU against the CPU oracle:
ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (ORBmatcher.cc:766-906, orbfe_search_by_bow_kf_batch) and
ORBmatcher::SearchForTriangulation (:908-1153, orbfe_search_for_triangulation_batch).  Index work: bit-exact (the epipolar
gate is fp32 in the reference's operation order, compared in double like the reference)."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBmatcher, _lib, scenes
from test_search_rule import kf_pair_case

pytestmark = pytest.mark.gpu


def _angles(c):
    return c["k1"]["angle"].astype(np.float32), c["k2"]["angle"].astype(np.float32)


@pytest.mark.parametrize("seed,w,h,nf,ratio,ori", [
    (9700, 752, 480, 1200, 0.75, True),     # LoopClosing / relocalisation: ORBmatcher(0.75, true)
    (9701, 752, 480, 1200, 0.9, True),
    (9702, 640, 480, 1000, 0.6, False),
    (9703, 320, 240, 400, 0.75, True),
])
def test_search_by_bow_kf_matches_oracle(oracle, seed, w, h, nf, ratio, ori):
    c = kf_pair_case(oracle, seed, w, h, nf)
    a1, a2 = _angles(c)
    rm, rn = oracle.search_by_bow_kf(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, c["has2"], c["fv2"], ratio, ori)
    nm, m = ORBmatcher(ratio, ori).SearchByBoWKeyFrames(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, c["has2"], c["fv2"])
    np.testing.assert_array_equal(m, rm)
    assert nm == rn and nm > 30
    got = m[m >= 0]
    assert len(np.unique(got)) == len(got)


def test_search_by_bow_kf_contended_node(oracle):
    """every feature in ONE vocabulary node: vbMatched2 decides, in KF1 feature order."""
    c = kf_pair_case(oracle, 9710, 320, 240, 300)
    a1, a2 = _angles(c)
    one = lambda n: (np.array([5], np.int32), np.array([0, n], np.int32), np.arange(n, dtype=np.int32))    # noqa: E731
    f1, f2 = one(len(c["d1"])), one(len(c["d2"]))
    rm, rn = oracle.search_by_bow_kf(c["d1"], a1, c["has1"], f1, c["d2"], a2, c["has2"], f2, 0.9, True)
    nm, m = ORBmatcher(0.9, True).SearchByBoWKeyFrames(c["d1"], a1, c["has1"], f1, c["d2"], a2, c["has2"], f2)
    np.testing.assert_array_equal(m, rm)
    assert nm == rn and nm > 50


def test_search_by_bow_kf_disjoint_and_empty(oracle):
    c = kf_pair_case(oracle, 9720, 320, 240, 300)
    a1, a2 = _angles(c)
    far = (c["fv2"][0] + 1000, c["fv2"][1], c["fv2"][2])
    mt = ORBmatcher(0.75, True)
    nm, m = mt.SearchByBoWKeyFrames(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, c["has2"], far)
    assert nm == 0 and (m == -1).all()
    nm, m = mt.SearchByBoWKeyFrames(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, np.zeros_like(c["has2"]), c["fv2"])
    assert nm == 0 and (m == -1).all()


def _tri_args(c, stereo, coarse, ori, all_free=False):
    n = max(len(c["k1"]), len(c["k2"]))
    if all_free:
        free1 = free2 = np.ones(n, np.uint8)
    else:
        free1, free2 = 1 - c["has1"], 1 - c["has2"]
        free1[::2], free2[::3] = 1, 1
    s1, s2 = (c["st1"], c["st2"]) if stereo else (None, None)
    return (c["k1"], c["d1"], free1[:len(c["k1"])], s1, c["fv1"], c["k2"], c["d2"], free2[:len(c["k2"])], s2, c["fv2"], c["F"],
            c["ep"], c["scale"], c["sigma2"], coarse, ori)


@pytest.mark.parametrize("seed,w,h,nf,coarse,ori,stereo", [
    (9800, 752, 480, 1200, False, True, False),     # LocalMapping::CreateNewMapPoints: ORBmatcher(0.6, false) + bCoarse by state
    (9801, 752, 480, 1200, True, True, False),
    (9802, 640, 480, 1000, False, False, True),
    (9803, 640, 480, 1000, False, True, True),
    (9804, 320, 240, 400, True, False, True),
])
def test_search_for_triangulation_matches_oracle(oracle, seed, w, h, nf, coarse, ori, stereo):
    c = kf_pair_case(oracle, seed, w, h, nf)
    k1, d1, free1, s1, fv1, k2, d2, free2, s2, fv2, F, ep, scale, sigma2, _, _ = a = _tri_args(c, stereo, coarse, ori)
    rm, rn = oracle.search_for_triangulation(*a)
    nm, pairs, m = ORBmatcher(0.6, ori).SearchForTriangulation(k1, d1, free1, fv1, k2, d2, free2, fv2, F, ep, scale, sigma2,
                                                               stereo1=s1, stereo2=s2, bCoarse=coarse)
    np.testing.assert_array_equal(m, rm)
    assert nm == rn and nm > 25 and len(pairs) == nm
    assert (pairs[:, 1] == m[pairs[:, 0]]).all() and (np.diff(pairs[:, 0]) > 0).all()
    assert free1[pairs[:, 0]].all() and free2[pairs[:, 1]].all()


def test_search_for_triangulation_gates(oracle):
    """the epipolar constraint and the epipole gate both remove pairs; a degenerate F12 (den == 0) removes all of them."""
    c = kf_pair_case(oracle, 9810, 640, 480, 1000)
    mt = ORBmatcher(0.6, False)
    counts = {}
    for name, coarse, F, ep in (("strict", False, c["F"], c["ep"]), ("coarse", True, c["F"], c["ep"]),
                                ("far_epipole", True, c["F"], np.array([-5000.0, -5000.0], np.float32)),
                                ("zero_f", False, np.zeros((3, 3), np.float32), c["ep"])):
        k1, d1, free1, s1, fv1, k2, d2, free2, s2, fv2, _, _, scale, sigma2, _, _ = _tri_args(c, False, coarse, False, all_free=True)
        a = (k1, d1, free1, s1, fv1, k2, d2, free2, s2, fv2, F, ep, scale, sigma2, coarse, False)
        rm, rn = oracle.search_for_triangulation(*a)
        nm, pairs, m = mt.SearchForTriangulation(k1, d1, free1, fv1, k2, d2, free2, fv2, F, ep, scale, sigma2, bCoarse=coarse)
        np.testing.assert_array_equal(m, rm)
        assert nm == rn
        counts[name] = nm
    assert counts["zero_f"] == 0 < counts["strict"] < counts["coarse"] <= counts["far_epipole"]


def test_bow_kf_matchers_ragged_device_batch(oracle):
    import torch
    cases = [kf_pair_case(oracle, 9820 + k, 320, 240, nf) for k, nf in enumerate((500, 300, 400))]
    cap = max(max(len(c["k1"]), len(c["k2"])) for c in cases)
    ncap = max(max(len(c["fv1"][0]), len(c["fv2"][0])) for c in cases)
    F = len(cases)
    spec = dict(k=((F, cap), _lib.KP_DTYPE), d=((F, cap, 32), np.uint8), a=((F, cap), np.float32), v=((F, cap), np.uint8),
                fr=((F, cap), np.uint8), s=((F, cap), np.uint8), id=((F, ncap), np.int32), st=((F, ncap + 1), np.int32),
                ix=((F, cap), np.int32), nn=((F,), np.int32))
    A = {side: {k: np.zeros(s, d) for k, (s, d) in spec.items()} for side in ("1", "2")}
    Fm, Ep = np.zeros((F, 9), np.float32), np.zeros((F, 2), np.float32)
    for f, c in enumerate(cases):
        for side in ("1", "2"):
            k, d, fv = c["k" + side], c["d" + side], c["fv" + side]
            n = len(k)
            X = A[side]
            X["k"][f, :n], X["d"][f, :n], X["a"][f, :n], X["v"][f, :n] = k, d, k["angle"], c["has" + side]
            X["fr"][f, :n], X["s"][f, :n] = 1 - c["has" + side], c["st" + side]
            X["id"][f, :len(fv[0])], X["st"][f, :len(fv[1])], X["ix"][f, :len(fv[2])], X["nn"][f] = fv[0], fv[1], fv[2], len(fv[0])
        Fm[f], Ep[f] = c["F"].reshape(9), c["ep"]
    dev = torch.device("cuda", 0)
    tt = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1) if a.dtype == _lib.KP_DTYPE else a).to(dev)    # noqa: E731
    T = {side: {k: tt(v) for k, v in A[side].items()} for side in ("1", "2")}
    dF, dE = torch.from_numpy(Fm).to(dev), torch.from_numpy(Ep).to(dev)
    flags = _lib.IN_DEVICE | _lib.OUT_DEVICE
    dm = torch.zeros((F, cap), dtype=torch.int32, device=dev)
    dc = torch.zeros(F, dtype=torch.int32, device=dev)
    a, b = T["1"], T["2"]
    torch.cuda.synchronize()    # the fills above run on torch's stream, the library on its own
    ORBmatcher(0.75, True).search_by_bow_kf_batch_raw(a["d"], a["a"], a["v"], a["id"], a["st"], a["ix"], a["nn"], b["d"], b["a"],
                                                      b["v"], b["id"], b["st"], b["ix"], b["nn"], cap, ncap, F, dm, dc, flags)
    m, cnt = dm.cpu().numpy(), dc.cpu().numpy()
    dm2 = torch.zeros((F, cap), dtype=torch.int32, device=dev)
    dc2 = torch.zeros(F, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    scale, sigma2 = cases[0]["scale"], cases[0]["sigma2"]
    ORBmatcher(0.6, True).search_for_triangulation_batch_raw(a["k"], a["d"], a["fr"], a["s"], a["id"], a["st"], a["ix"], a["nn"],
                                                             b["k"], b["d"], b["fr"], b["s"], b["id"], b["st"], b["ix"], b["nn"],
                                                             cap, ncap, F, dF, dE, scale, sigma2, False, dm2, dc2, flags)
    m2, cnt2 = dm2.cpu().numpy(), dc2.cpu().numpy()
    for f, c in enumerate(cases):
        a1, a2 = _angles(c)
        n1 = len(c["k1"])
        rm, rn = oracle.search_by_bow_kf(c["d1"], a1, c["has1"], c["fv1"], c["d2"], a2, c["has2"], c["fv2"], 0.75, True)
        np.testing.assert_array_equal(m[f, :n1], rm)
        assert cnt[f] == rn and (m[f, n1:] == -1).all()
        rm2, rn2 = oracle.search_for_triangulation(c["k1"], c["d1"], 1 - c["has1"], c["st1"], c["fv1"], c["k2"], c["d2"],
                                                   1 - c["has2"], c["st2"], c["fv2"], c["F"], c["ep"], scale, sigma2, False, True)
        np.testing.assert_array_equal(m2[f, :n1], rm2)
        assert cnt2[f] == rn2 and (m2[f, n1:] == -1).all()


# ---- two-camera (KannalaBrandt8) key frames: orbfe_search_for_triangulation_kb8_batch -------------------------------------------
# Tolerance: the triangulation is floating point (tanf / atan2f differ by an ulp or two between CUDA and glibc):
# a feature's match may differ from the oracle's only where the oracle reports that one of its gates was decided within 1e-5
# (`slack`); everywhere else the indices are identical.
from test_search_rule import kb8_tri_case    # noqa: E402


@pytest.mark.parametrize("seed,coarse,ori", [(9900, False, True), (9901, False, False), (9902, True, True), (9903, False, True)])
def test_search_for_triangulation_kb8_matches_oracle(oracle, seed, coarse, ori):
    rig, c, fv1, fv2, free1, free2 = kb8_tri_case(seed, npts=1200)
    rm, rn, slack = oracle.search_for_triangulation_kb8(c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"], c["desc2"],
                                                        free2, c["nleft2"], fv2, rig, c["t12"], c["sigma2"], coarse, False)
    nm, pairs, m = ORBmatcher(0.6, False).SearchForTriangulationKB8(c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"],
                                                                    c["desc2"], free2, c["nleft2"], fv2, rig, c["t12"],
                                                                    c["sigma2"], bCoarse=coarse)
    differ = m != rm
    assert not (differ & (slack >= 1e-5)).any(), "match differs away from a gate: features %s" % np.nonzero(differ & (slack >= 1e-5))[0][:10]
    assert differ.sum() <= 3 and abs(nm - rn) <= 3 and nm > 400
    if coarse:
        assert not differ.any()
    if ori and not differ.any():    # the rotation histogram on top (exact when the matches are)
        rm2, rn2, _ = oracle.search_for_triangulation_kb8(c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"], c["desc2"],
                                                          free2, c["nleft2"], fv2, rig, c["t12"], c["sigma2"], coarse, True)
        nm2, _, m2 = ORBmatcher(0.6, True).SearchForTriangulationKB8(c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"],
                                                                     c["desc2"], free2, c["nleft2"], fv2, rig, c["t12"],
                                                                     c["sigma2"], bCoarse=coarse)
        np.testing.assert_array_equal(m2, rm2)
        assert nm2 == rn2 and nm2 <= nm
    got = m >= 0
    r1, r2 = np.nonzero(got)[0] >= c["nleft1"], m[got] >= c["nleft2"]
    assert all(((r1 == a) & (r2 == b)).sum() > 20 for a in (False, True) for b in (False, True))


def test_search_for_triangulation_kb8_errors():
    from orbslam3lib_b200 import _lib as L
    rig, c, fv1, fv2, free1, free2 = kb8_tri_case(9910, npts=200)
    bad = rig.copy()
    bad["cam1"][0] = 0
    with pytest.raises(L.OrbfeError):
        ORBmatcher(0.6, False).SearchForTriangulationKB8(c["keys1"], c["desc1"], free1, c["nleft1"], fv1, c["keys2"], c["desc2"],
                                                         free2, c["nleft2"], fv2, bad, c["t12"], c["sigma2"])
