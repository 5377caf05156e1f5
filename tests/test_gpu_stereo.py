"""-m gpu: Frame::ComputeStereoMatches (Frame.cc:827-997) on the GPU (orbfe_extract_stereo_batch) against the CPU oracle.

Bars: the matched set, mvuRight and mvDepth are bit-exact (the float arithmetic is restated op by op, IEEE fp32 without
FMA on both sides); keypoints / descriptors of both eyes as in the extractor tests."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBextractor, scenes

pytestmark = pytest.mark.gpu

MB, FX = 0.11, 435.2    # EuRoC-like baseline (m) and focal length (px): mbf = mb * fx, maxD = fx


def _oracle_pair(oracle, left, right, nf, nl, mb, mbf):
    eL, eR = oracle.OracleExtractor(nf, 1.2, nl, 20, 7), oracle.OracleExtractor(nf, 1.2, nl, 20, 7)
    _, kl, dl = eL(left)
    _, kr, dr = eR(right)
    u, z, sad, faults = oracle.stereo_matches(eL, eR, kl, dl, kr, dr, mb, mbf)
    assert faults == 0
    return kl, dl, kr, dr, u, z


def _check_pair(oracle, ex_out, p, left, right, nf, nl, mb, mbf):
    counts, kps, desc, u_right, depth = ex_out
    kl, dl, kr, dr, u, z = _oracle_pair(oracle, left, right, nf, nl, mb, mbf)
    nL, nR = int(counts[p, 0]), int(counts[p, 1])
    assert (nL, nR) == (len(kl), len(kr))
    for f in ("x", "y", "octave"):
        np.testing.assert_array_equal(kps[p, 0, :nL][f], kl[f])
        np.testing.assert_array_equal(kps[p, 1, :nR][f], kr[f])
    np.testing.assert_array_equal(desc[p, 0, :nL], dl)
    np.testing.assert_array_equal(desc[p, 1, :nR], dr)
    np.testing.assert_array_equal(u_right[p, :nL] >= 0, u >= 0, "matched set of pair %d" % p)
    np.testing.assert_array_equal(u_right[p, :nL].view(np.uint32), u.view(np.uint32), "mvuRight bits, pair %d" % p)
    np.testing.assert_array_equal(depth[p, :nL].view(np.uint32), z.view(np.uint32), "mvDepth bits, pair %d" % p)
    return int((u >= 0).sum())


@pytest.mark.parametrize("kind,w,h,seed,nf,nl", [
    ("texture", 752, 480, 2000, 1200, 8),
    ("checker", 640, 480, 2001, 1000, 8),
    ("noise", 320, 240, 2002, 500, 8),
    ("flat_blobs", 752, 480, 2003, 1000, 8),
    ("texture", 1280, 720, 2004, 2000, 10),
])
def test_stereo_matches_oracle(oracle, kind, w, h, seed, nf, nl):
    left = scenes.make(kind, w, h, seed)
    right = scenes.stereo_right_rectified(left, seed)
    ex = ORBextractor(nf, 1.2, nl, 20, 7)
    out = ex.extract_stereo_batch(np.stack([left, right])[None], MB, MB * FX)
    matched = _check_pair(oracle, out, 0, left, right, nf, nl, MB, MB * FX)
    if kind == "texture":
        assert matched > 100    # the scene must exercise the SAD / parabola / median path


def test_stereo_row_dependent_disparity(oracle):
    """the C2 bench scene (disparity changes every row: most SAD windows are poor, many parabola rejections)."""
    left = scenes.make("texture", 752, 480, 2010)
    right = scenes.stereo_right(left, 2010)
    ex = ORBextractor(1200, 1.2, 8, 20, 7)
    out = ex.extract_stereo_batch(np.stack([left, right])[None], MB, MB * FX)
    _check_pair(oracle, out, 0, left, right, 1200, 8, MB, MB * FX)


def test_stereo_small_max_disparity(oracle):
    """a short baseline: maxD = 12 px cuts the disparity window (Frame.cc:857-859, 876-877, 969)."""
    left = scenes.make("texture", 640, 480, 2020)
    right = scenes.stereo_right_rectified(left, 2020)
    ex = ORBextractor(1000, 1.2, 8, 20, 7)
    mb, mbf = 0.05, 0.05 * 12.0
    out = ex.extract_stereo_batch(np.stack([left, right])[None], mb, mbf)
    _check_pair(oracle, out, 0, left, right, 1000, 8, mb, mbf)


def test_stereo_batch_chunked_and_pipelined(oracle):
    """300 pairs through the pipelined host path (128-image stages) and 10 pairs through 4-image chunks: every pair
    equals the single-pair result, and sampled pairs equal the oracle."""
    npairs, w, h, nf = 300, 320, 240, 400
    imgs = np.empty((npairs, 2, h, w), np.uint8)
    for p in range(npairs):
        imgs[p, 0] = scenes.make("texture", w, h, 3000 + (p % 7))
        imgs[p, 1] = scenes.stereo_right_rectified(imgs[p, 0], 3000 + (p % 7))
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    out = ex.extract_stereo_batch(imgs, MB, MB * FX)
    for p in (0, 1, 6, 127, 128, 299):
        _check_pair(oracle, out, p, imgs[p, 0], imgs[p, 1], nf, 8, MB, MB * FX)
    for p in range(7, npairs):    # seeds repeat with period 7
        np.testing.assert_array_equal(out[3][p], out[3][p % 7])
        np.testing.assert_array_equal(out[4][p], out[4][p % 7])
    ex2 = ORBextractor(nf, 1.2, 8, 20, 7)
    per_image = 3 * 1024 * 1024
    ex2.set_workspace_limit(4 * per_image)
    out2 = ex2.extract_stereo_batch(imgs[:10], MB, MB * FX)
    np.testing.assert_array_equal(out2[0], out[0][:10])
    for p in range(10):
        for e in range(2):
            n = int(out[0][p, e])
            np.testing.assert_array_equal(out2[1][p, e, :n], out[1][p, e, :n])    # slots past the count are undefined
            np.testing.assert_array_equal(out2[2][p, e, :n], out[2][p, e, :n])
    np.testing.assert_array_equal(out2[3], out[3][:10])
    np.testing.assert_array_equal(out2[4], out[4][:10])


def test_stereo_bad_arguments():
    from orbslam3lib_b200._lib import OrbfeError
    ex = ORBextractor(500, 1.2, 8, 20, 7)
    img = np.zeros((1, 2, 120, 160), np.uint8)
    with pytest.raises(OrbfeError):
        ex.extract_stereo_batch(img, 0.0, 10.0)    # mb must be positive (maxD = mbf / mb)
    counts, kps, desc, u, z = ex.extract_stereo_batch(img, MB, MB * FX)    # flat images: no keypoints, no matches
    assert counts.sum() == 0 and (u == -1).all() and (z == -1).all()
