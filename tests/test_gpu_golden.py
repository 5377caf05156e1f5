"""-m gpu: the CUDA path (through the C ABI) against the committed golden fixtures (cv2-generated, tests/golden/):
independent of the oracle library, so a defect shared by oracle and kernels cannot hide."""
import os

import numpy as np
import pytest

import golden_util as gu
from orbslam3lib_b200 import BFMatcher, ORBextractor, scenes

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", gu.extract_fixtures(), ids=lambda p: os.path.basename(p)[8:-4])
def test_extractor_fixture(path):
    g = gu.load(path)
    w, h, seed, nf, nl, ini, mn, lap0, lap1 = [int(v) for v in g["meta"]]
    img = scenes.make(str(g["scene"]), w, h, seed)
    ex = ORBextractor(nf, 1.2, nl, ini, mn)
    mono, kps, desc = ex(img, (lap0, lap1))
    np.testing.assert_array_equal(ex.features_per_level, g["features_per_level"])
    np.testing.assert_array_equal(ex.umax, g["umax"])
    np.testing.assert_array_equal(ex.GetScaleFactors(), g["scale"])
    gu.check_extractor_against_fixture(ex, g, img)
    gu.check_outputs_against_fixture(mono, kps, desc, g)


def test_knn_fixture():
    g = gu.load(os.path.join(gu.GOLDEN, "primitives.npz"))
    nq, nt, seed = [int(v) for v in g["knn_meta"]]
    q, t = scenes.map_descriptors(nq, nt, seed)
    idx, dist = BFMatcher().knnMatch(q, t)
    np.testing.assert_array_equal(idx, g["knn_idx"])
    np.testing.assert_array_equal(dist, g["knn_dist"])


@pytest.mark.parametrize("path", gu.stereo_fixtures(), ids=lambda p: os.path.basename(p)[7:-4])
def test_stereo_fixture(path):
    """Frame::ComputeStereoMatches on the GPU against the fixture made with cv2.norm + numpy float32."""
    g = gu.load(path)
    w, h, seed, nf, nl = [int(v) for v in g["meta"]]
    left, right = gu.stereo_images(g)
    mb, mbf = [float(v) for v in g["mb_mbf"]]
    ex = ORBextractor(nf, 1.2, nl, 20, 7)
    counts, kps, desc, u, z = ex.extract_stereo_batch(np.stack([left, right])[None], mb, mbf)
    nL, nR = int(counts[0, 0]), int(counts[0, 1])
    gu.check_stereo_against_fixture(nL, nR, desc[0, 0, :nL], u[0, :nL], z[0, :nL], g)


def test_undistort_fixture():
    """Frame::UndistortKeyPoints on the GPU against the committed cv2.undistortPoints fixture, bit for bit."""
    from orbslam3lib_b200 import ORBmatcher, _lib
    g = gu.load(os.path.join(gu.GOLDEN, "undistort.npz"))
    K, pts = g["K"], g["pts"]
    kps = np.zeros(len(pts), _lib.KP_DTYPE)
    kps["x"], kps["y"] = pts[:, 0], pts[:, 1]
    kps["octave"], kps["angle"], kps["class_id"] = np.arange(len(pts)) % 8, 33.25, -1
    m = ORBmatcher()
    for i in range(3):
        out = m.UndistortKeyPoints(kps, K, g["dist_%d" % i])
        ref = g["un_%d" % i]
        np.testing.assert_array_equal(out["x"].view(np.uint32), np.ascontiguousarray(ref[:, 0]).view(np.uint32))
        np.testing.assert_array_equal(out["y"].view(np.uint32), np.ascontiguousarray(ref[:, 1]).view(np.uint32))
        for f in ("size", "angle", "response", "octave", "class_id"):
            np.testing.assert_array_equal(out[f], kps[f])
    same = m.UndistortKeyPoints(kps, K, [0.0, 0.1, 0.0, 0.0])    # mDistCoef[0] == 0 -> copy
    np.testing.assert_array_equal(same, kps)
    assert len(m.UndistortKeyPoints(kps[:0], K, g["dist_0"])) == 0
