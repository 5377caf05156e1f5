"""CPU (not gpu): the C++ oracle against the committed golden fixtures that oracle/gen_golden.py produced from genuine
OpenCV (cv2) primitives + the restated ORB-SLAM glue.  This is what pins the oracle (the reference
itself ships no vectors for this path); the GPU parity tests then compare the CUDA path with the pinned oracle."""
import hashlib
import os
import re

import numpy as np
import pytest

import golden_util as gu
from orbslam3lib_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def prim():
    return gu.load(os.path.join(gu.GOLDEN, "primitives.npz"))


def test_resize_linear(oracle, prim):
    w, h, seed = [int(v) for v in prim["resize_src_meta"]]
    src = scenes.make("noise", w, h, seed)
    for i, (dw, dh) in enumerate(prim["resize_sizes"]):
        np.testing.assert_array_equal(oracle.resize_linear(src, int(dw), int(dh)), prim["resize_%d" % i])


def test_border_reflect101(oracle, prim):
    w, h, seed = [int(v) for v in prim["resize_src_meta"]]
    src = scenes.make("noise", w, h, seed)
    np.testing.assert_array_equal(oracle.border_reflect101(src[:40, :50].copy(), 19), prim["border"])


def test_gaussian7x7(oracle, prim):
    w, h, seed = [int(v) for v in prim["blur_meta"]]
    np.testing.assert_array_equal(oracle.gaussian7x7(scenes.make("texture", w, h, seed)), prim["blur"])
    w, h, seed = [int(v) for v in prim["resize_src_meta"]]
    np.testing.assert_array_equal(oracle.gaussian7x7(scenes.make("noise", w, h, seed)), prim["blur_noise"])


@pytest.mark.parametrize("kind,seed", [("noise", 5003), ("texture", 5004), ("checker", 5005)])
@pytest.mark.parametrize("t", [7, 20])
def test_fast9_nms(oracle, prim, kind, seed, t):
    w, h = [int(v) for v in prim["fast_meta"]]
    got = oracle.fast9_nms(scenes.make(kind, w, h, seed), t)
    np.testing.assert_array_equal(got, prim["fast_%s_%d" % (kind, t)])    # order included


def test_fast_atan2(oracle, prim):
    got = np.array([oracle.fast_atan2(y, x) for y, x in zip(prim["atan_y"], prim["atan_x"])], np.float32)
    np.testing.assert_array_equal(got, prim["atan"])    # the polynomial is restated op by op: bit exact


def test_knn2(oracle, prim):
    nq, nt, seed = [int(v) for v in prim["knn_meta"]]
    q, t = scenes.map_descriptors(nq, nt, seed)
    for threads in (1, 3):
        idx, dist = oracle.knn2_hamming(q, t, threads=threads)
        np.testing.assert_array_equal(idx, prim["knn_idx"])
        np.testing.assert_array_equal(dist, prim["knn_dist"])


@pytest.mark.parametrize("path", gu.extract_fixtures(), ids=lambda p: os.path.basename(p)[8:-4])
def test_extractor_fixture(oracle, path):
    g = gu.load(path)
    w, h, seed, nf, nl, ini, mn, lap0, lap1 = [int(v) for v in g["meta"]]
    img = scenes.make(str(g["scene"]), w, h, seed)
    ex = oracle.OracleExtractor(nf, 1.2, nl, ini, mn)
    mono, kps, desc = ex(img, (lap0, lap1))
    np.testing.assert_array_equal(ex.features_per_level, g["features_per_level"])
    np.testing.assert_array_equal(ex.umax, g["umax"])
    np.testing.assert_array_equal(ex.scale, g["scale"])
    gu.check_extractor_against_fixture(ex, g, img)
    gu.check_outputs_against_fixture(mono, kps, desc, g)


def test_pattern_table():
    """The rBRIEF table both sides sample with: its digest is pinned, and (in the build container, where the
    reference is mounted) it equals the table of ORBextractor_old.cc:151-409."""
    src = open(os.path.join(ROOT, "orbslam3lib_b200", "csrc", "orb_pattern.h")).read()
    vals = [int(v) for v in re.findall(r"-?\d+", src[src.index("{") + 1:src.rindex("}")])]
    assert len(vals) == 1024
    digest = hashlib.sha256(np.array(vals, np.int8).tobytes()).hexdigest()
    assert digest == open(os.path.join(gu.GOLDEN, "pattern.sha256")).read().strip()
    ref = "/root/reference/app/src/main/cpp/src/ORBextractor_old.cc"
    if os.path.exists(ref):
        m = re.search(r"bit_pattern_31_\[256\*4\]\s*=\s*\{(.*?)\};", open(ref).read(), re.S)
        body = re.sub(r"/\*.*?\*/", "", m.group(1), flags=re.S)
        body = re.sub(r"//.*", "", body)
        assert [int(v) for v in re.findall(r"-?\d+", body)] == vals


@pytest.mark.parametrize("path", gu.stereo_fixtures(), ids=lambda p: os.path.basename(p)[7:-4])
def test_stereo_matches_fixture(oracle, path):
    """orc_stereo_matches (Frame::ComputeStereoMatches restated in C++) against the cv2.norm / numpy-float32 fixture."""
    g = gu.load(path)
    w, h, seed, nf, nl = [int(v) for v in g["meta"]]
    left, right = gu.stereo_images(g)
    mb, mbf = [float(v) for v in g["mb_mbf"]]
    eL, eR = oracle.OracleExtractor(nf, 1.2, nl, 20, 7), oracle.OracleExtractor(nf, 1.2, nl, 20, 7)
    _, kl, dl = eL(left)
    _, kr, dr = eR(right)
    u, z, sad, faults = oracle.stereo_matches(eL, eR, kl, dl, kr, dr, mb, mbf)
    assert faults == 0
    gu.check_stereo_against_fixture(len(kl), len(kr), dl, u, z, g)
    assert int((u >= 0).sum()) > 50
    assert ((sad >= 0) >= (u >= 0)).all()    # the median cut only removes accepted matches


def test_undistort_fixture(oracle):
    """orc_undistort_keypoints against the committed cv2.undistortPoints fixture."""
    g = gu.load(os.path.join(gu.GOLDEN, "undistort.npz"))
    K, pts = g["K"], g["pts"]
    kps = np.zeros(len(pts), oracle.KP_DTYPE)
    kps["x"], kps["y"] = pts[:, 0], pts[:, 1]
    for i in range(3):
        out = oracle.undistort_keypoints(kps, K[0, 0], K[1, 1], K[0, 2], K[1, 2], g["dist_%d" % i])
        ref = g["un_%d" % i]
        np.testing.assert_array_equal(out["x"].view(np.uint32), np.ascontiguousarray(ref[:, 0]).view(np.uint32))
        np.testing.assert_array_equal(out["y"].view(np.uint32), np.ascontiguousarray(ref[:, 1]).view(np.uint32))
