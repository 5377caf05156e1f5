"""-m gpu: ORBmatcher::SearchForInitialization (ORBmatcher.cc:649-764) on the GPU (orbfe_search_for_initialization_batch)
against the CPU oracle: vnMatches12, the updated vbPrevMatched and the match count, including the order-dependent state
(vMatchedDistance / vnMatches21 stealing) and the rotation-histogram filter.  Index work: bit-exact."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBmatcher, _lib, scenes

pytestmark = pytest.mark.gpu


def _pair(oracle, kind, w, h, seed, nf):
    cur = scenes.make(kind, w, h, seed)
    prv = scenes.previous_frame(cur, seed)
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, k1, d1 = ex(prv)
    _, k2, d2 = ex(cur)
    return k1, d1, k2, d2


@pytest.mark.parametrize("kind,w,h,seed,nf,window,ratio,ori", [
    ("texture", 752, 480, 8100, 5000, 100, 0.9, True),      # Tracking::MonocularInitialization: 5 * nFeatures, window 100
    ("texture", 752, 480, 8101, 2000, 100, 0.9, False),
    ("texture", 640, 480, 8102, 2000, 10, 0.9, True),       # the default windowSize
    ("checker", 640, 480, 8103, 2000, 100, 0.97, True),     # repeated structure: ties and steals
    ("noise", 320, 240, 8104, 1000, 50, 0.9, True),
    ("flat_blobs", 752, 480, 8105, 1000, 100, 0.9, True),
])
def test_search_for_initialization_matches_oracle(oracle, kind, w, h, seed, nf, window, ratio, ori):
    k1, d1, k2, d2 = _pair(oracle, kind, w, h, seed, nf)
    pm = np.stack([k1["x"], k1["y"]], 1)
    bounds = (0.0, 0.0, float(w), float(h))
    rm, rp, rn = oracle.search_for_initialization(k1, d1, k2, d2, bounds, pm, window, ratio, ori)
    n, m12, pm2 = ORBmatcher(ratio, ori).SearchForInitialization(k1, d1, k2, d2, bounds, pm, window)
    np.testing.assert_array_equal(m12, rm)
    np.testing.assert_array_equal(pm2, rp)
    assert n == rn
    if kind == "texture":
        assert n > 100


def test_search_for_initialization_contended(oracle):
    """every keypoint of F1 aimed at the same spot with a window covering most of F2: each commit touches the windows
    of the rest of its chunk, so the ordered prefix commit runs one point at a time; plus a second call fed with the
    first call's vbPrevMatched, as Tracking does while the initial match count stays low."""
    k1, d1, k2, d2 = _pair(oracle, "texture", 320, 240, 8110, 1500)
    pm = np.tile(np.array([[160.0, 120.0]], np.float32), (len(k1), 1))
    bounds = (0.0, 0.0, 320.0, 240.0)
    for ratio in (0.9, 1.0):
        rm, rp, rn = oracle.search_for_initialization(k1, d1, k2, d2, bounds, pm, 150, ratio, True)
        n, m12, pm2 = ORBmatcher(ratio, True).SearchForInitialization(k1, d1, k2, d2, bounds, pm, 150)
        np.testing.assert_array_equal(m12, rm)
        np.testing.assert_array_equal(pm2, rp)
        assert n == rn and n > 20
    rm, rp2, rn = oracle.search_for_initialization(k1, d1, k2, d2, bounds, rp, 20, 0.9, True)
    n, m12, pm3 = ORBmatcher(0.9, True).SearchForInitialization(k1, d1, k2, d2, bounds, pm2, 20)
    np.testing.assert_array_equal(m12, rm)
    np.testing.assert_array_equal(pm3, rp2)
    # the same frame against itself: every level-0 keypoint finds itself at distance 0
    n, m12, _ = ORBmatcher(0.9, True).SearchForInitialization(k1, d1, k1, d1, bounds, np.stack([k1["x"], k1["y"]], 1), 5)
    lv0 = k1["octave"] == 0
    rm, _, rn = oracle.search_for_initialization(k1, d1, k1, d1, bounds, np.stack([k1["x"], k1["y"]], 1), 5, 0.9, True)
    np.testing.assert_array_equal(m12, rm)
    assert n == rn and n > 0.9 * lv0.sum() and (m12[~lv0] == -1).all()


def test_search_for_initialization_batch_and_edges(oracle):
    m = ORBmatcher(0.9, True)
    k1, d1, k2, d2 = _pair(oracle, "texture", 320, 240, 8120, 600)
    bounds = (0.0, 0.0, 320.0, 240.0)
    n, m12, pm = m.SearchForInitialization(k1[:0], d1[:0], k2, d2, bounds, np.zeros((0, 2), np.float32), 100)
    assert n == 0 and len(m12) == 0
    n, m12, pm = m.SearchForInitialization(k1, d1, k2[:0], d2[:0], bounds, np.stack([k1["x"], k1["y"]], 1), 100)
    assert n == 0 and (m12 == -1).all()
    with pytest.raises(_lib.OrbfeError):
        m.SearchForInitialization(k1, d1, k2, d2, (0, 0, 320, 0), np.stack([k1["x"], k1["y"]], 1), 100)
    # a window centre far outside the image finds nothing and is left alone
    far = np.full((len(k1), 2), 1e6, np.float32)
    n, m12, pm = m.SearchForInitialization(k1, d1, k2, d2, bounds, far, 100)
    assert n == 0 and (pm == far).all()
    # ragged batch of 4 pairs == the single calls
    nfr, cap = 4, max(len(k1), len(k2))
    K1, K2 = np.zeros((nfr, cap), _lib.KP_DTYPE), np.zeros((nfr, cap), _lib.KP_DTYPE)
    D1, D2 = np.zeros((nfr, cap, 32), np.uint8), np.zeros((nfr, cap, 32), np.uint8)
    PM = np.zeros((nfr, cap, 2), np.float32)
    n1 = np.array([len(k1), len(k1) - 31, 0, len(k1) // 2], np.int32)
    n2 = np.array([len(k2), len(k2) // 3, len(k2), len(k2) - 5], np.int32)
    singles = []
    for f in range(nfr):
        K1[f, :n1[f]], D1[f, :n1[f]] = k1[:n1[f]], d1[:n1[f]]
        K2[f, :n2[f]], D2[f, :n2[f]] = k2[:n2[f]], d2[:n2[f]]
        PM[f, :n1[f], 0], PM[f, :n1[f], 1] = k1["x"][:n1[f]] + f, k1["y"][:n1[f]] - f
        singles.append(oracle.search_for_initialization(k1[:n1[f]], d1[:n1[f]], k2[:n2[f]], d2[:n2[f]], bounds,
                                                        PM[f, :n1[f]], 40 + 20 * f, 0.9, True))
    M = np.zeros((nfr, cap), np.int32)
    cnt = np.zeros(nfr, np.int32)
    for f in range(nfr):    # windowSize is per call: one call per distinct window
        pmf = PM[f:f + 1].copy()
        m.search_for_initialization_batch_raw(K1[f:f + 1], D1[f:f + 1], n1[f:f + 1], K2[f:f + 1], D2[f:f + 1], n2[f:f + 1],
                                              cap, 1, bounds, 40 + 20 * f, pmf, M[f:f + 1], cnt[f:f + 1])
        np.testing.assert_array_equal(M[f, :n1[f]], singles[f][0])
        assert (M[f, n1[f]:] == -1).all()
        np.testing.assert_array_equal(pmf[0, :n1[f]], singles[f][1])
        assert cnt[f] == singles[f][2]
    # all four pairs in one launch with a common window
    ref = [oracle.search_for_initialization(k1[:n1[f]], d1[:n1[f]], k2[:n2[f]], d2[:n2[f]], bounds, PM[f, :n1[f]], 60, 0.9,
                                            True) for f in range(nfr)]
    pmb = PM.copy()
    m.search_for_initialization_batch_raw(K1, D1, n1, K2, D2, n2, cap, nfr, bounds, 60, pmb, M, cnt)
    for f in range(nfr):
        np.testing.assert_array_equal(M[f, :n1[f]], ref[f][0])
        np.testing.assert_array_equal(pmb[f, :n1[f]], ref[f][1])
        assert cnt[f] == ref[f][2]
