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
