"""Helpers shared by the golden-fixture tests (tests/golden/*.npz, written by oracle/gen_golden.py from cv2)."""
import glob
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
ANGLE_TOL_DEG = 1e-3    # BASELINE.json north_star: angles within 1e-3 degrees


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def extract_fixtures():
    return sorted(glob.glob(os.path.join(GOLDEN, "extract_*.npz")))


def load(path):
    return np.load(path, allow_pickle=False)


def angle_diff(a, b):
    d = np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))
    return np.minimum(d, 360.0 - d)


def check_extractor_against_fixture(ex, g, image):
    """ex: an object with the stage accessors shared by oracle_lib.OracleExtractor and orbslam3lib_b200.ORBextractor
    (called already on `image`); g: a loaded extract_*.npz."""
    w, h, seed, nf, nl, ini, mn, lap0, lap1 = [int(v) for v in g["meta"]]
    assert sha(image) == str(g["image_sha256"]), "scene generator drifted from the fixture"
    for l in range(nl):
        assert tuple(ex.level_size(l)) == tuple(int(v) for v in g["level_size_%d" % l])
        assert sha(ex.level_padded(l)) == str(g["padded_sha256_%d" % l]), "padded pyramid level %d" % l
        np.testing.assert_array_equal(ex.level_candidates(l), g["cand_%d" % l], "FAST candidates level %d" % l)
        lk, gk = ex.level_keypoints(l), g["lkps_%d" % l]
        assert len(lk) == len(gk), "level %d keypoint count" % l
        if len(gk):
            np.testing.assert_array_equal(lk["x"], gk[:, 0])
            np.testing.assert_array_equal(lk["y"], gk[:, 1])
            np.testing.assert_array_equal(lk["response"], gk[:, 3])
            assert angle_diff(lk["angle"], gk[:, 2]).max() <= ANGLE_TOL_DEG
            assert sha(ex.level_blurred(l)) == str(g["blur_sha256_%d" % l]), "blurred level %d" % l
    np.testing.assert_array_equal(ex.level_padded(nl - 1), g["padded_last"])


def check_outputs_against_fixture(mono, kps, desc, g):
    gk = g["kps"]
    assert mono == int(g["mono"])
    assert len(kps) == len(gk)
    np.testing.assert_array_equal(kps["x"], gk[:, 0])
    np.testing.assert_array_equal(kps["y"], gk[:, 1])
    np.testing.assert_array_equal(kps["size"], gk[:, 2])
    np.testing.assert_array_equal(kps["response"], gk[:, 4])
    np.testing.assert_array_equal(kps["octave"], g["octave"])
    assert len(gk) == 0 or angle_diff(kps["angle"], gk[:, 3]).max() <= ANGLE_TOL_DEG
    np.testing.assert_array_equal(desc, g["desc"])


def stereo_fixtures():
    return sorted(glob.glob(os.path.join(GOLDEN, "stereo_*.npz")))


def stereo_images(g):
    """the rectified pair of a stereo_*.npz fixture, checked against the stored hashes."""
    from orbslam3lib_b200 import scenes
    w, h, seed, nf, nl = [int(v) for v in g["meta"]]
    left = scenes.make(str(g["scene"]), w, h, seed)
    right = scenes.stereo_right_rectified(left, seed)
    assert sha(left) == str(g["left_sha256"]) and sha(right) == str(g["right_sha256"]), "scene generator drifted"
    return left, right


def check_stereo_against_fixture(n_left, n_right, desc_left, u_right, depth, g):
    """mvuRight / mvDepth must equal the cv2-generated fixture bit for bit."""
    assert (n_left, n_right) == (int(g["n_left"]), int(g["n_right"]))
    assert sha(desc_left) == str(g["desc_left_sha256"])
    np.testing.assert_array_equal(np.asarray(u_right, np.float32).view(np.uint32), g["u_right"].view(np.uint32))
    np.testing.assert_array_equal(np.asarray(depth, np.float32).view(np.uint32), g["depth"].view(np.uint32))
