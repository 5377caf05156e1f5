#!/usr/bin/env python
"""Generate tests/golden/*.npz.  Test infrastructure.

The reference ships no golden vectors for this path, so the fixtures are outputs of the reference
itself run here, and of the genuine OpenCV it depends on:
  * extract_*.npz   the reference's own ORBextractor_old.cc, compiled from /root/reference into
                    oracle/_ref/libext_old.so (oracle/ref_ext/make_ref_ext.py), cross-checked at generation time
                    against cv2 primitives + the Python statement of the glue (oracle/py/cv2_oracle.py);
  * primitives / stereo / undistort   genuine cv2.resize, copyMakeBorder, FastFeatureDetector, GaussianBlur,
                    fastAtan2, BFMatcher, norm, undistortPoints;
  * fisheye_*.npz   the reference's vendored Eigen 3.4.0 (oracle/_ref/libeigen_svd.so).
cv2 and /root/reference exist only in the build container; the fixtures travel, this script does not need to.

    python oracle/gen_golden.py          # rewrites tests/golden/
"""
import hashlib
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))

GOLDEN = os.path.join(ROOT, "tests", "golden")

# (name, scene, w, h, seed, nfeatures, nlevels, iniTh, minTh, lapping)
EXTRACT_CASES = [
    ("texture_320x240", "texture", 320, 240, 4001, 400, 6, 20, 7, (0, 0)),
    ("checker_256x200", "checker", 256, 200, 4002, 300, 5, 20, 7, (60, 180)),
    ("blobs_376x240", "flat_blobs", 376, 240, 4003, 500, 8, 20, 7, (0, 0)),
    ("noise_160x120", "noise", 160, 120, 4004, 200, 4, 20, 7, (0, 1000)),
    ("texture_640x480", "texture", 640, 480, 1000, 1000, 8, 20, 7, (0, 0)),    # BASELINE configs[0] shape
]


def load_pattern():
    src = open(os.path.join(ROOT, "orbslam3lib_b200", "csrc", "orb_pattern.h")).read()
    body = src[src.index("{") + 1:src.rindex("}")]
    vals = [int(v) for v in re.findall(r"-?\d+", body)]
    assert len(vals) == 1024
    return np.array(vals, np.int8)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gen_extract(case):
    """extract_*.npz: outputs of the REFERENCE'S OWN extractor (oracle/_ref/libext_old.so = ORBextractor_old.cc compiled
    from /root/reference, `make -C oracle ref_ext`) — tables, padded pyramid, per-level keypoints after the quadtree,
    final keypoints / descriptors / monoIndex.  The same image goes through the cv2-backed Python statement
    (genuine cv2 primitives + restated glue); generation FAILS unless the two agree bit for bit, so every fixture is
    pinned from both sides: reference glue over cv2-pinned primitives == restated glue over real cv2 primitives.
    The two intermediates the reference does not keep (FAST candidates before the quadtree, blurred planes) come from
    the cv2 side."""
    import cv2
    import cv2_oracle
    import ref_ext_lib
    from orbslam3lib_b200 import scenes
    name, kind, w, h, seed, nf, nl, ini, mn, lap = case
    img = scenes.make(kind, w, h, seed)
    if not ref_ext_lib.available():
        raise SystemExit("extract fixtures need oracle/_ref/libext_old.so (make -C oracle ref_ext; needs /root/reference)")
    ref = ref_ext_lib.RefExtractor(nf, 1.2, nl, ini, mn)
    mono, rk, desc = ref(img, lap)
    ex = cv2_oracle.Cv2Extractor(nf, 1.2, nl, ini, mn)
    ex.set_pattern(load_pattern())
    mono_c, kps_c, desc_c = ex(img, lap)
    kc = np.array([[k[0], k[1], k[2], k[3], k[4]] for k in kps_c], np.float32).reshape(-1, 5)
    rk5 = np.stack([rk["x"], rk["y"], rk["size"], rk["angle"], rk["response"]], 1)
    assert mono == mono_c and np.array_equal(rk5, kc) and np.array_equal(desc, desc_c), "reference build != cv2 statement"
    assert np.array_equal(rk["octave"], np.array([k[5] for k in kps_c], np.int32))
    assert np.array_equal(ref.features_per_level, ex.features_per_level) and np.array_equal(ref.umax, ex.umax)
    out = {
        "meta": np.array([w, h, seed, nf, nl, ini, mn, lap[0], lap[1]], np.int64),
        "scene": np.array(kind),
        "source": np.array("oracle/_ref (ORBextractor_old.cc compiled from /root/reference, g++ libstdc++); "
                           "cross-checked bit for bit against cv2 " + cv2.__version__ + " + oracle/py/cv2_oracle.py"),
        "opencv_version": np.array(cv2.__version__),
        "image_sha256": np.array(sha(img)),
        "mono": np.array(mono, np.int32),
        "kps": rk5,
        "octave": rk["octave"].astype(np.int32),
        "desc": desc,
        "features_per_level": ref.features_per_level.astype(np.int32),
        "umax": ref.umax.astype(np.int32),
        "scale": ref.scale.astype(np.float32),
    }
    for l in range(nl):
        padded = ref.level_padded(l)
        assert np.array_equal(padded, ex.pyr_padded[l]), "ComputePyramid level %d" % l
        out["padded_sha256_%d" % l] = np.array(sha(padded))
        out["level_size_%d" % l] = np.array(ref.level_size(l), np.int32)
        out["cand_%d" % l] = np.array(ex.candidates[l], np.int32).reshape(-1, 3)
        lk = ref.level_keypoints(l)
        lc = np.array([[k[0], k[1], k[3], k[4]] for k in ex.level_kps[l]], np.float32).reshape(-1, 4)
        lr = np.stack([lk["x"], lk["y"], lk["angle"], lk["response"]], 1).astype(np.float32).reshape(-1, 4)
        assert np.array_equal(lr, lc), "level %d keypoints after the quadtree" % l
        out["lkps_%d" % l] = lr
        if ex.blurred[l] is not None:
            out["blur_sha256_%d" % l] = np.array(sha(ex.blurred[l]))
    # one small level stored verbatim so that a hash mismatch can be localised
    out["padded_last"] = ref.level_padded(nl - 1)
    np.savez_compressed(os.path.join(GOLDEN, "extract_%s.npz" % name), **out)
    print("extract_%s: %d keypoints, mono %d (reference build == cv2 statement)" % (name, len(rk), mono))


def gen_primitives():
    import cv2
    from orbslam3lib_b200 import scenes
    out = {"opencv_version": np.array(cv2.__version__)}
    # resize: every (src -> dst) step of a 1.2 pyramid on an awkward size + a generic pair
    src = scenes.make("noise", 211, 173, 5001)
    sizes = [(176, 144), (147, 120), (97, 61), (210, 172), (106, 87)]
    out["resize_src_meta"] = np.array([211, 173, 5001], np.int64)
    out["resize_sizes"] = np.array(sizes, np.int32)
    for i, (dw, dh) in enumerate(sizes):
        out["resize_%d" % i] = cv2.resize(src, (dw, dh), 0, 0, cv2.INTER_LINEAR)
    out["border"] = cv2.copyMakeBorder(src[:40, :50].copy(), 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    small = scenes.make("texture", 123, 77, 5002)
    out["blur_meta"] = np.array([123, 77, 5002], np.int64)
    out["blur"] = cv2.GaussianBlur(small, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    out["blur_noise"] = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    # FAST on cell-sized crops, both thresholds, order included
    for kind, seed in (("noise", 5003), ("texture", 5004), ("checker", 5005)):
        cell = scenes.make(kind, 41, 43, seed)
        for t in (7, 20):
            det = cv2.FastFeatureDetector_create(t, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kp = det.detect(cell)
            out["fast_%s_%d" % (kind, t)] = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kp],
                                                      np.int32).reshape(-1, 3)
    out["fast_meta"] = np.array([41, 43], np.int64)
    # fastAtan2 on moment-like inputs
    r = scenes.rand_u32(4000, 5006).astype(np.int64)
    ys = (r[:2000] % 200001 - 100000).astype(np.float32)
    xs = (r[2000:] % 200001 - 100000).astype(np.float32)
    ys[:8] = [0, 0, 1, -1, 5, -5, 0, 3]
    xs[:8] = [0, 1, 0, 0, 5, -5, -7, 3]
    out["atan_y"], out["atan_x"] = ys, xs
    out["atan"] = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(ys, xs)], np.float32)
    # BFMatcher kNN (ties included)
    q, t = scenes.map_descriptors(300, 700, 5007)
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
    out["knn_meta"] = np.array([300, 700, 5007], np.int64)
    out["knn_idx"] = np.array([[d.trainIdx for d in row] for row in m], np.int32)
    out["knn_dist"] = np.array([[int(d.distance) for d in row] for row in m], np.int32)
    np.savez_compressed(os.path.join(GOLDEN, "primitives.npz"), **out)
    print("primitives: ok")


# (name, scene, w, h, seed, nfeatures, nlevels, mb, fx)
STEREO_CASES = [
    ("texture_376x240", "texture", 376, 240, 4101, 500, 8, 0.11, 435.2),
    ("blobs_320x240", "flat_blobs", 320, 240, 4102, 400, 6, 0.11, 435.2),
    ("texture_320x200_short_baseline", "texture", 320, 200, 4103, 400, 6, 0.05, 14.0),
]


def gen_stereo(case):
    """Frame::ComputeStereoMatches fixture: cv2-backed extraction of both eyes + stereo_matches_cv2 (cv2.norm)."""
    import cv2
    import cv2_oracle
    from orbslam3lib_b200 import scenes
    name, kind, w, h, seed, nf, nl, mb, fx = case
    left = scenes.make(kind, w, h, seed)
    right = scenes.stereo_right_rectified(left, seed)
    kp_dtype = np.dtype([("x", "<f4"), ("y", "<f4"), ("octave", "<i4")])
    eyes = []
    for img in (left, right):
        ex = cv2_oracle.Cv2Extractor(nf, 1.2, nl, 20, 7)
        ex.set_pattern(load_pattern())
        _, kps, desc = ex(img, (0, 0))
        ka = np.zeros(len(kps), kp_dtype)
        ka["x"] = [k[0] for k in kps]
        ka["y"] = [k[1] for k in kps]
        ka["octave"] = [k[5] for k in kps]
        eyes.append((ex, ka, desc))
    (exl, kl, dl), (exr, kr, dr) = eyes
    mbf = float(np.float32(mb) * np.float32(fx))
    u, z = cv2_oracle.stereo_matches_cv2(exl.pyr, exr.pyr, exl.scale, exl.inv_scale, kl, dl, kr, dr, mb, mbf)
    np.savez_compressed(os.path.join(GOLDEN, "stereo_%s.npz" % name),
                        meta=np.array([w, h, seed, nf, nl], np.int64), scene=np.array(kind),
                        mb_mbf=np.array([mb, mbf], np.float32), opencv_version=np.array(cv2.__version__),
                        left_sha256=np.array(sha(left)), right_sha256=np.array(sha(right)),
                        n_left=np.array(len(kl), np.int32), n_right=np.array(len(kr), np.int32),
                        desc_left_sha256=np.array(sha(dl)), u_right=u, depth=z)
    print("stereo_%s: %d / %d keypoints, %d matches" % (name, len(kl), len(kr), int((u >= 0).sum())))


def gen_undistort():
    """Frame::UndistortKeyPoints fixture: genuine cv2.undistortPoints on seeded positions, three coefficient sets."""
    import cv2
    from orbslam3lib_b200 import scenes
    fx, fy, cx, cy = 458.654, 457.296, 367.215, 248.375
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    r = scenes.rand_u32(6000, 4201).astype(np.float64) / 4294967296.0
    pts = np.stack([(r[:3000] * 752).astype(np.float32), (r[3000:] * 480).astype(np.float32)], 1)
    out = {"K": K, "pts": pts, "opencv_version": np.array(cv2.__version__), "seed": np.array(4201)}
    for i, d in enumerate(([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], [-0.28, 0.07, 0.0002, 1.7e-05, 0.01],
                           [0.9, -2.5, 0.01, -0.02])):
        D = np.array(d, np.float32)
        out["dist_%d" % i] = D
        out["un_%d" % i] = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    np.savez_compressed(os.path.join(GOLDEN, "undistort.npz"), **out)
    print("undistort: ok")


def gen_fisheye():
    """Frame::ComputeStereoFishEyeMatches / KannalaBrandt8::TriangulateMatches fixtures: the oracle's float restatement
    with the 4x4 null vector solved by the REAL Eigen::JacobiSVD<Matrix4f> of the reference's vendored Eigen 3.4.0
    (oracle/ref_eigen -> oracle/_ref/libeigen_svd.so; `make -C oracle ref`, needs /root/reference)."""
    import subprocess
    from orbslam3lib_b200 import scenes
    sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
    import oracle_lib
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"])
    if not oracle_lib.use_eigen_svd(True):
        raise SystemExit("fisheye: oracle/_ref/libeigen_svd.so missing (no /root/reference here)")
    try:
        for name, w, h, seed, nf, ml, mr, base in (("a", 640, 400, 4301, 1500, 120, 80, 0.11),
                                                   ("b", 752, 480, 4302, 1000, 0, 0, 0.064)):
            img = scenes.make("texture", w, h, seed)
            _, kl, _ = oracle_lib.OracleExtractor(nf, 1.2, 8, 20, 7)(img)
            rig = scenes.fisheye_rig(w, h, base, seed)
            kr, idx, dist = scenes.fisheye_stereo_case(kl, rig, seed, ml, mr)
            ltr, rtl, depth, p3d, status, n, slack = oracle_lib.stereo_fisheye_matches(kl, ml, kr, mr, idx, dist, rig)
            np.savez_compressed(os.path.join(GOLDEN, "fisheye_%s.npz" % name), keys_left=kl, keys_right=kr, knn_idx=idx,
                                knn_dist=dist, rig=rig, mono=np.array([ml, mr]), ltr=ltr, rtl=rtl, depth=depth, p3d=p3d,
                                status=status, nmatches=np.array(n), slack=slack, eigen_version=np.array("3.4.0"))
            print("fisheye_%s: %d matches of %d rows" % (name, n, len(idx)))
    finally:
        oracle_lib.use_eigen_svd(False)


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    if "--extract-only" in sys.argv:
        for c in EXTRACT_CASES:
            gen_extract(c)
        return
    if "--fisheye-only" in sys.argv:
        return gen_fisheye()
    if "--undistort-only" in sys.argv:
        return gen_undistort()
    if "--stereo-only" in sys.argv:
        for c in STEREO_CASES:
            gen_stereo(c)
        return
    gen_primitives()
    gen_undistort()
    gen_fisheye()
    for c in STEREO_CASES:
        gen_stereo(c)
    for c in EXTRACT_CASES:
        gen_extract(c)
    pat = load_pattern()
    open(os.path.join(GOLDEN, "pattern.sha256"), "w").write(hashlib.sha256(pat.tobytes()).hexdigest() + "\n")


if __name__ == "__main__":
    main()
