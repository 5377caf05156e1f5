"""CPU (not gpu): the restated C++ oracle (oracle/cpp) against the REFERENCE'S OWN extractor glue, compiled from
/root/reference/app/src/main/cpp/src/ORBextractor_old.cc into oracle/_ref/libext_old.so by
oracle/ref_ext/make_ref_ext.py (`make -C oracle ref_ext`).

This is what pins the ORB-SLAM glue — cell loop, DistributeOctTree / DivideNode / compareNodes incl. the std::sort tie
order, IC_Angle, computeOrbDescriptor, the mono-front / stereo-back packing of operator() — to the reference's code
rather than to a second reading of it.  The OpenCV primitives under the glue are pinned against genuine OpenCV by
tests/test_oracle_vs_cv2.py and tests/golden/primitives.npz.

The library is built where /root/reference exists and travels with the snapshot; where it is missing the tests skip
(the committed fixtures tests/golden/extract_*.npz, generated from it, still apply: test_oracle_golden.py)."""
import numpy as np
import pytest

import ref_ext_lib
from orbslam3lib_b200 import scenes

pytestmark = pytest.mark.skipif(not ref_ext_lib.available(), reason="oracle/_ref/libext_old.so not built (no /root/reference)")

# the scenes of tests/test_gpu_extractor.py::CASES, plus lapping-area, small-image and threshold variants
CASES = [
    # kind, w, h, seed, nfeatures, nlevels, iniTh, minTh, lapping
    ("texture", 640, 480, 1000, 1000, 8, 20, 7, (0, 0)),
    ("texture", 752, 480, 1001, 1200, 8, 20, 7, (0, 0)),
    ("checker", 640, 480, 1002, 1000, 8, 20, 7, (0, 0)),
    ("noise", 320, 240, 1003, 500, 8, 20, 7, (0, 0)),
    ("flat_blobs", 752, 480, 1004, 1000, 8, 20, 7, (0, 0)),
    ("texture", 211, 173, 1005, 300, 5, 20, 7, (0, 0)),
    ("noise", 752, 480, 1006, 1200, 8, 20, 7, (0, 0)),
    ("texture", 1280, 720, 1007, 2000, 10, 20, 7, (0, 0)),
    ("texture", 640, 480, 1010, 1000, 8, 20, 7, (200, 440)),
    ("checker", 752, 480, 1011, 1200, 8, 20, 7, (0, 511)),
    ("noise", 640, 400, 1012, 1500, 8, 20, 7, (100, 1000)),
    ("flat_blobs", 320, 240, 1013, 500, 8, 40, 12, (0, 0)),
    ("texture", 160, 120, 1014, 200, 4, 20, 7, (0, 0)),
    ("texture", 640, 480, 1015, 500, 8, 20, 7, (0, 0)),        # C3 shape: 500 features
]


def _same_keypoints(a, b, what):
    assert len(a) == len(b), "%s: %d vs %d keypoints" % (what, len(a), len(b))
    for f in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        np.testing.assert_array_equal(a[f].view(np.uint32), b[f].view(np.uint32), "%s field %s" % (what, f))


@pytest.mark.parametrize("kind,w,h,seed,nf,nl,ini,mn,lap", CASES)
def test_oracle_equals_reference_extractor(oracle, kind, w, h, seed, nf, nl, ini, mn, lap):
    img = scenes.make(kind, w, h, seed)
    ours = oracle.OracleExtractor(nf, 1.2, nl, ini, mn)
    ref = ref_ext_lib.RefExtractor(nf, 1.2, nl, ini, mn)
    # a-0: constructor tables
    for f in ("scale", "inv_scale", "sigma2", "inv_sigma2"):
        np.testing.assert_array_equal(getattr(ours, f).view(np.uint32), getattr(ref, f).view(np.uint32), f)
    np.testing.assert_array_equal(ours.features_per_level, ref.features_per_level)
    np.testing.assert_array_equal(ours.umax, ref.umax)
    mono_o, kps_o, desc_o = ours(img, lap)
    mono_r, kps_r, desc_r = ref(img, lap)
    # a-7: packing (mono front / stereo back) — order included; a-4: the angle bit for bit (same fastAtan2 under both)
    assert mono_o == mono_r
    _same_keypoints(kps_o, kps_r, "operator()")
    np.testing.assert_array_equal(desc_o, desc_r)                       # a-6
    for l in range(nl):
        assert ours.level_size(l) == ref.level_size(l)
        np.testing.assert_array_equal(ours.level_padded(l), ref.level_padded(l), "ComputePyramid level %d" % l)   # a-1
        _same_keypoints(ours.level_keypoints(l), ref.level_keypoints(l), "level %d after the quadtree" % l)      # a-2/3


def test_sort_tie_path_is_exercised(oracle):
    """compareNodes is not a strict weak order on (count, UL.x) alone: equal pairs reach std::sort and the result then
    depends on the sort algorithm.  The scenes above must actually contain such ties, or the equality of the quadtree
    output says nothing about the tie order."""
    total = 0
    for kind, w, h, seed, nf, nl, ini, mn, lap in CASES[:4]:
        ours = oracle.OracleExtractor(nf, 1.2, nl, ini, mn)
        ours(scenes.make(kind, w, h, seed), lap)
        total += ours.sort_ties()
    assert total > 100


@pytest.mark.parametrize("seed,n,N", [(7001, 400, 60), (7002, 1500, 217), (7003, 3000, 261), (7004, 90, 200),
                                      (7005, 2500, 31), (7006, 1, 10), (7007, 5000, 939), (7008, 2, 1), (7009, 800, 800)])
def test_distribute_octtree_stress(oracle, seed, n, N):
    """DistributeOctTree alone on lattice-clustered integer positions with few distinct responses: the stress case for
    DivideNode's boundaries, the compareNodes tie order inside std::sort and the first-maximum rule of the final pick.
    Restated oracle == reference build, kept keys AND their order."""
    r = scenes.rand_u32(3 * n, seed).astype(np.int64)
    W, H = 720, 448
    keys = np.zeros(n, ref_ext_lib.KP_DTYPE)
    keys["x"] = (r[:n] % (W // 4)) * 4            # a lattice: many nodes with equal counts and equal UL.x
    keys["y"] = (r[n:2 * n] % (H // 2)) * 2
    keys["response"] = (r[2 * n:] % 8 + 20).astype(np.float32)
    keys["size"], keys["angle"], keys["class_id"] = 7, -1, -1
    ref = ref_ext_lib.RefExtractor(1000, 1.2, 8, 20, 7)
    ours = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    a = ref.distribute(keys, 16, 16 + W, 16, 16 + H, N)
    b, ties = ours.distribute(keys, 16, 16 + W, 16, 16 + H, N)
    _same_keypoints(b, a, "DistributeOctTree (%d compareNodes ties)" % ties)
    assert n < 300 or N < 100 or ties > 0


def test_committed_fixtures_are_reference_output():
    """tests/golden/extract_*.npz carry "source": oracle/_ref and are exactly what the reference build produces now."""
    import golden_util as gu
    paths = gu.extract_fixtures()
    assert len(paths) >= 5
    for path in paths:
        g = gu.load(path)
        assert str(g["source"]).startswith("oracle/_ref"), path
        w, h, seed, nf, nl, ini, mn, lap0, lap1 = [int(v) for v in g["meta"]]
        img = scenes.make(str(g["scene"]), w, h, seed)
        ref = ref_ext_lib.RefExtractor(nf, 1.2, nl, ini, mn)
        mono, kps, desc = ref(img, (lap0, lap1))
        assert mono == int(g["mono"])
        np.testing.assert_array_equal(np.stack([kps[f] for f in ("x", "y", "size", "angle", "response")], 1), g["kps"])
        np.testing.assert_array_equal(kps["octave"], g["octave"])
        np.testing.assert_array_equal(desc, g["desc"])
        for l in range(nl):
            assert gu.sha(ref.level_padded(l)) == str(g["padded_sha256_%d" % l])
