"""CPU (not gpu): the dependency-free C++ oracle against GENUINE OpenCV (cv2) on randomised inputs — every primitive
the reference path calls (SURVEY.md §8c self-checks) and the whole extractor against the cv2-backed Python glue.
Skipped where cv2 is not installed (the committed fixtures of tests/golden/ cover that case)."""
import os
import sys

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from orbslam3lib_b200 import scenes  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def _pyramid_sizes(w, h, nlevels):
    s, out = np.float32(1.0), []
    for _ in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(float(s) * float(np.float32(1.2)))
    return out


@pytest.mark.parametrize("w,h,nl", [(752, 480, 8), (640, 480, 8), (211, 173, 5), (960, 540, 10)])
def test_resize_every_level(oracle, w, h, nl):
    img = scenes.make("noise", w, h, 7000 + w)
    sizes = _pyramid_sizes(w, h, nl)
    for (dw, dh) in sizes[1:]:
        ref = cv2.resize(img, (dw, dh), 0, 0, cv2.INTER_LINEAR)
        np.testing.assert_array_equal(oracle.resize_linear(img, dw, dh), ref)
        img = ref


def test_resize_random_pairs(oracle):
    rng = np.random.default_rng(1)
    for k in range(40):
        sw, sh = int(rng.integers(8, 300)), int(rng.integers(8, 300))
        dw, dh = max(2, int(sw / rng.uniform(1.05, 1.9))), max(2, int(sh / rng.uniform(1.05, 1.9)))
        # (an exact 2x pair switches OpenCV to INTER_AREA: the oracle follows, test_resize_exact_half_is_the_box_average)
        img = scenes.make("noise" if k % 2 else "texture", sw, sh, 7100 + k)
        np.testing.assert_array_equal(oracle.resize_linear(img, dw, dh), cv2.resize(img, (dw, dh), 0, 0, cv2.INTER_LINEAR))


def test_border_and_blur(oracle):
    rng = np.random.default_rng(2)
    for k in range(30):
        w, h = int(rng.integers(20, 260)), int(rng.integers(20, 200))
        img = scenes.make(("noise", "texture", "checker")[k % 3], w, h, 7200 + k)
        np.testing.assert_array_equal(oracle.border_reflect101(img, 19),
                                      cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101))
        np.testing.assert_array_equal(oracle.gaussian7x7(img),
                                      cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101))


@pytest.mark.parametrize("t", [1, 7, 20, 60])
def test_fast_cells(oracle, t):
    det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    rng = np.random.default_rng(3)
    for k in range(60):
        w, h = int(rng.integers(7, 60)), int(rng.integers(7, 60))
        img = scenes.make(("noise", "texture", "checker", "flat_blobs")[k % 4], w, h, 7300 + k)
        kp = det.detect(img)
        ref = np.array([[int(p.pt[0]), int(p.pt[1]), int(p.response)] for p in kp], np.int32).reshape(-1, 3)
        np.testing.assert_array_equal(oracle.fast9_nms(img, t), ref)


def test_fast_atan2_moments(oracle):
    r = scenes.rand_u32(20000, 7400).astype(np.int64)
    ys = (r[:10000] % 2000001 - 1000000).astype(np.float32)
    xs = (r[10000:] % 2000001 - 1000000).astype(np.float32)
    got = np.array([oracle.fast_atan2(y, x) for y, x in zip(ys, xs)], np.float32)
    ref = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(ys, xs)], np.float32)
    np.testing.assert_array_equal(got, ref)


@pytest.mark.parametrize("nq,nt", [(1, 1), (5, 2), (200, 333), (1000, 1000)])
def test_knn2_vs_bfmatcher(oracle, nq, nt):
    import cv2_oracle
    q, t = scenes.map_descriptors(nq, nt, 7500 + nq)
    ri, rd = cv2_oracle.knn2(q, t)
    idx, dist = oracle.knn2_hamming(q, t, threads=2)
    np.testing.assert_array_equal(idx, ri)
    np.testing.assert_array_equal(dist, rd)


@pytest.mark.parametrize("kind,w,h,seed,nf,nl,lap", [
    ("texture", 400, 300, 7601, 500, 7, (0, 0)),
    ("checker", 320, 240, 7602, 400, 6, (100, 220)),
    ("noise", 200, 150, 7603, 300, 5, (0, 0)),
    ("flat_blobs", 752, 480, 7604, 1200, 8, (0, 0)),
])
def test_extractor_vs_cv2_glue(oracle, kind, w, h, seed, nf, nl, lap):
    import cv2_oracle
    import gen_golden
    img = scenes.make(kind, w, h, seed)
    ref = cv2_oracle.Cv2Extractor(nf, 1.2, nl, 20, 7)
    ref.set_pattern(gen_golden.load_pattern())
    rmono, rk, rd = ref(img, lap)
    ex = oracle.OracleExtractor(nf, 1.2, nl, 20, 7)
    mono, kps, desc = ex(img, lap)
    assert mono == rmono and len(kps) == len(rk)
    for l in range(nl):
        np.testing.assert_array_equal(ex.level_padded(l), ref.pyr_padded[l])
        np.testing.assert_array_equal(ex.level_candidates(l), np.array(ref.candidates[l], np.int32).reshape(-1, 3))
        if ref.blurred[l] is not None:
            np.testing.assert_array_equal(ex.level_blurred(l), ref.blurred[l])
    rk = np.array([[k[0], k[1], k[2], k[3], k[4], k[5]] for k in rk], np.float32).reshape(-1, 6)
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        np.testing.assert_array_equal(kps[f], rk[:, i], f)
    np.testing.assert_array_equal(kps["octave"], rk[:, 5].astype(np.int32))
    np.testing.assert_array_equal(desc, rd)


@pytest.mark.parametrize("kind,w,h,seed,nf,rectified", [
    ("texture", 752, 480, 7301, 1200, True),
    ("texture", 640, 480, 7302, 1000, False),
    ("checker", 376, 240, 7303, 500, True),
    ("noise", 320, 240, 7304, 500, True),
    ("flat_blobs", 752, 480, 7305, 1000, True),
])
def test_stereo_matches_vs_cv2(oracle, kind, w, h, seed, nf, rectified):
    """Frame::ComputeStereoMatches: the C++ restatement against the cv2.norm(NORM_L1) + numpy-float32 restatement on
    the oracle's own keypoints and pyramids (both eyes), bit for bit."""
    import cv2_oracle
    left = scenes.make(kind, w, h, seed)
    right = scenes.stereo_right_rectified(left, seed) if rectified else scenes.stereo_right(left, seed)
    eL, eR = oracle.OracleExtractor(nf, 1.2, 8, 20, 7), oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kl, dl = eL(left)
    _, kr, dr = eR(right)
    for mb, mbf in ((0.11, 0.11 * 435.2), (0.05, 0.05 * 20.0)):
        u, z, sad, faults = oracle.stereo_matches(eL, eR, kl, dl, kr, dr, mb, mbf)
        assert faults == 0
        pl = [eL.level_pixels(l) for l in range(8)]
        pr = [eR.level_pixels(l) for l in range(8)]
        u2, z2 = cv2_oracle.stereo_matches_cv2(pl, pr, eL.scale, eL.inv_scale, kl, dl, kr, dr, mb, mbf)
        np.testing.assert_array_equal(u.view(np.uint32), u2.view(np.uint32))
        np.testing.assert_array_equal(z.view(np.uint32), z2.view(np.uint32))


@pytest.mark.parametrize("dist", [
    [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05],           # EuRoC cam0
    [-0.28, 0.07, 0.0002, 1.7e-05, 0.01],                            # with k3
    [0.9, -2.5, 0.01, -0.02],                                        # strong: icdist < 0 near the corners
    [0.0, 0.07, 0.0002, 1.7e-05],                                    # mDistCoef[0] == 0 -> copy (Frame.cc:765-769)
])
def test_undistort_keypoints_vs_cv2(oracle, dist):
    """Frame::UndistortKeyPoints: the restated cv::undistortPoints (5 iterations, double) against cv2, bit for bit."""
    fx, fy, cx, cy = 458.654, 457.296, 367.215, 248.375
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    D = np.array(dist, np.float32)
    r = scenes.rand_u32(8000, 7400).astype(np.float64) / 4294967296.0
    kps = np.zeros(4000, oracle.KP_DTYPE)
    kps["x"], kps["y"] = (r[:4000] * 752).astype(np.float32), (r[4000:] * 480).astype(np.float32)
    kps["octave"], kps["angle"] = np.arange(4000) % 8, 17.5
    out = oracle.undistort_keypoints(kps, K[0, 0], K[1, 1], K[0, 2], K[1, 2], D)
    if D[0] == 0:
        ref = np.stack([kps["x"], kps["y"]], 1)
    else:
        ref = cv2.undistortPoints(np.stack([kps["x"], kps["y"]], 1).reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    np.testing.assert_array_equal(out["x"].view(np.uint32), np.ascontiguousarray(ref[:, 0]).view(np.uint32))
    np.testing.assert_array_equal(out["y"].view(np.uint32), np.ascontiguousarray(ref[:, 1]).view(np.uint32))
    for f in ("size", "angle", "response", "octave", "class_id"):
        np.testing.assert_array_equal(out[f], kps[f])


def test_resize_exact_half_is_the_box_average(oracle):
    """cv::resize(INTER_LINEAR) turns into INTER_AREA when both axes shrink by exactly 2 (scaleFactor 2.0 on even sizes);
    one axis exact and the other not stays bilinear."""
    rng = np.random.RandomState(77)
    for (w, h, dw, dh) in ((640, 480, 320, 240), (322, 250, 161, 125), (64, 62, 32, 31), (640, 481, 320, 240), (641, 480, 320, 240)):
        img = rng.randint(0, 256, (h, w)).astype(np.uint8)
        np.testing.assert_array_equal(oracle.resize_linear(img, dw, dh), cv2.resize(img, (dw, dh), 0, 0, cv2.INTER_LINEAR))
