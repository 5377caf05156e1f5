"""-m gpu: Frame::ComputeStereoFishEyeMatches after BFMatchORB + KannalaBrandt8::TriangulateMatches on the GPU
(orbfe_stereo_fisheye_matches_batch) against the CPU oracle and the Eigen-made fixtures.
Tolerances (floating point): vs the oracle (same algorithm, fp64 null vector on both sides; only tanf / atan2f / cosf /
sinf differ by an ulp or two between CUDA and glibc, i.e. the rays by ~1e-7) mvDepth within 1e-5 relative OR 1e-6 rad of
parallax (|1/z - 1/z'| * |tlr|), the bearing of mvStereo3Dpoints within 1e-6, decisions identical except within 1e-5 of a gate; vs the Eigen fixtures the bounds of test_oracle_fisheye.py."""
import os

import numpy as np
import pytest

from golden_util import GOLDEN
from orbslam3lib_b200 import BFMatcher, ORBextractor, ORBmatcher, _lib, scenes
from test_oracle_fisheye import FISHEYE_TOL, check_against

pytestmark = pytest.mark.gpu


def _close(got, ref, slack, rig, rtol=1e-5):
    n, ltr, rtl, depth, p3d, status = got
    rltr, rrtl, rdepth, rp3d, rstatus, rn, _ = ref
    edge = slack < 1e-5
    same = status == rstatus
    assert (same | edge).all(), "decision differs away from a gate: rows %s" % np.nonzero(~(same | edge))[0][:10]
    np.testing.assert_array_equal(ltr[same], rltr[same])
    both = (status == 1) & (rstatus == 1)
    z, zr = depth[both].astype(np.float64), rdepth[both].astype(np.float64)
    base = float(np.linalg.norm(rig["tlr"]))
    ok = (np.abs(z - zr) <= rtol * zr) | (np.abs(1 / z - 1 / zr) * base <= 1e-6)
    assert ok.all(), "depth: %s vs %s" % (z[~ok][:5], zr[~ok][:5])
    if both.any():
        b, br = p3d[both].astype(np.float64), rp3d[both].astype(np.float64)
        b, br = b / np.linalg.norm(b, axis=1, keepdims=True), br / np.linalg.norm(br, axis=1, keepdims=True)
        assert np.abs(b - br).max() <= 1e-6
    assert (depth[status != 1] == -1).all() and (ltr[status != 1] == -1).all() and (p3d[status != 1] == 0).all()
    assert n == int((status == 1).sum())
    if same.all():
        np.testing.assert_array_equal(rtl, rrtl)
        assert n == rn


@pytest.mark.parametrize("w,h,seed,nf,ml,mr,base", [
    (640, 400, 8200, 1500, 120, 80, 0.11),
    (752, 480, 8201, 1200, 0, 0, 0.064),
    (640, 400, 8202, 2000, 500, 0, 0.2),
    (320, 240, 8203, 400, 0, 37, 0.05),
])
def test_fisheye_matches_vs_oracle(oracle, w, h, seed, nf, ml, mr, base):
    img = scenes.make("texture", w, h, seed)
    _, kl, _ = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)(img)
    rig = scenes.fisheye_rig(w, h, base, seed)
    kr, idx, dist = scenes.fisheye_stereo_case(kl, rig, seed, ml, mr)
    ref = oracle.stereo_fisheye_matches(kl, ml, kr, mr, idx, dist, rig)
    got = ORBmatcher().ComputeStereoFishEyeMatches(kl, ml, kr, mr, idx, dist, rig)
    _close(got, ref, ref[6], rig)
    assert got[0] > 50


@pytest.mark.parametrize("name", ["a", "b"])
def test_fisheye_matches_vs_eigen_fixture(name):
    fix = np.load(os.path.join(GOLDEN, "fisheye_%s.npz" % name))
    ml, mr = (int(v) for v in fix["mono"])
    n, ltr, rtl, depth, p3d, status = ORBmatcher().ComputeStereoFishEyeMatches(
        fix["keys_left"], ml, fix["keys_right"], mr, fix["knn_idx"], fix["knn_dist"], fix["rig"])
    check_against(fix, ltr, rtl, depth, p3d, status, n, FISHEYE_TOL, 1e-3)


def test_fisheye_from_images_end_to_end(oracle):
    """the whole Lynx stereo frame: two extractions with lapping areas, BFMatchORB on the stereo-area descriptors, then
    the triangulation — every stage on the GPU, every stage checked against the oracle."""
    w, h = 640, 400
    left = scenes.make("texture", w, h, 8210)
    right = scenes.stereo_right(left, 8210)
    ex = ORBextractor(1000, 1.2, 8, 20, 7)
    lap = (200, 640), (0, 440)
    ml, kl, dl = ex(left, lap[0])
    mr, kr, dr = ex(right, lap[1])
    bf = BFMatcher()
    idx, dist = bf.knnMatch(dl[ml:], dr[mr:])
    rig = scenes.fisheye_rig(w, h, 0.11, 8210)
    rig["level_sigma2"][:8] = np.asarray(ex.GetScaleSigmaSquares(), np.float32)
    got = ORBmatcher().ComputeStereoFishEyeMatches(kl, ml, kr, mr, idx, dist, rig)
    oex = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    oml, okl, odl = oex(left, lap[0])
    omr, okr, odr = oex(right, lap[1])
    assert (oml, omr) == (ml, mr) and 0 < ml < len(kl)
    np.testing.assert_array_equal(dl, odl)
    oidx, odist = oracle.knn2_hamming(odl[oml:], odr[omr:])
    np.testing.assert_array_equal(idx, oidx)
    ref = oracle.stereo_fisheye_matches(okl, oml, okr, omr, oidx, odist, rig)
    _close(got, ref, ref[6], rig)


def test_fisheye_batch_device_and_edges(oracle):
    import torch
    m = ORBmatcher()
    img = scenes.make("texture", 320, 240, 8220)
    _, kl, _ = oracle.OracleExtractor(500, 1.2, 8, 20, 7)(img)
    rig = scenes.fisheye_rig(320, 240, 0.1, 8220)
    kr, idx, dist = scenes.fisheye_stereo_case(kl, rig, 8220, 10, 5)
    # no right keypoints / no knn rows / index -1 / monoLeft == Nleft
    n, ltr, rtl, depth, p3d, st = m.ComputeStereoFishEyeMatches(kl, 10, kr[:0], 0, np.full_like(idx, -1), np.full_like(dist, -1), rig)
    assert n == 0 and (ltr == -1).all() and (depth == -1).all() and len(rtl) == 0
    n, ltr, rtl, depth, p3d, st = m.ComputeStereoFishEyeMatches(kl, len(kl), kr, 5, idx[:0], dist[:0], rig)
    assert n == 0 and (st == 0).all()
    n, ltr, rtl, depth, p3d, st = m.ComputeStereoFishEyeMatches(kl[:0], 0, kr, 5, idx[:0], dist[:0], rig)
    assert n == 0 and len(ltr) == 0 and (rtl == -1).all()
    bad = rig.copy()
    bad["cam2"][0] = 0
    with pytest.raises(_lib.OrbfeError):
        m.ComputeStereoFishEyeMatches(kl, 10, kr, 5, idx, dist, bad)
    # ragged batch, device-resident in and out
    nfr, cap = 4, max(len(kl), len(kr))
    KL, KR = np.zeros((nfr, cap), _lib.KP_DTYPE), np.zeros((nfr, cap), _lib.KP_DTYPE)
    I, D = np.full((nfr, cap, 2), -1, np.int32), np.full((nfr, cap, 2), -1, np.int32)
    nl = np.array([len(kl), len(kl) - 40, 0, len(kl) // 2], np.int32)
    mls = np.array([10, 0, 0, 33], np.int32)
    mrs = np.array([5, 17, 0, 0], np.int32)
    nr = np.zeros(nfr, np.int32)
    refs = []
    for f in range(nfr):
        k = kl[:nl[f]]
        r, i2, d2 = scenes.fisheye_stereo_case(k, rig, 8230 + f, int(mls[f]), int(mrs[f]))
        nr[f] = len(r)
        KL[f, :nl[f]], KR[f, :nr[f]] = k, r
        I[f, :len(i2)], D[f, :len(d2)] = i2, d2
        refs.append(oracle.stereo_fisheye_matches(k, int(mls[f]), r, int(mrs[f]), i2, d2, rig))
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1)).to(dev)    # noqa: E731
    dKL, dKR, dI, dD = t(KL), t(KR), torch.from_numpy(I).to(dev), torch.from_numpy(D).to(dev)
    dn = [torch.from_numpy(a).to(dev) for a in (nl, nr, mls, mrs)]
    o_ltr = torch.zeros((nfr, cap), dtype=torch.int32, device=dev)
    o_rtl = torch.zeros((nfr, cap), dtype=torch.int32, device=dev)
    o_st = torch.zeros((nfr, cap), dtype=torch.int32, device=dev)
    o_z = torch.zeros((nfr, cap), dtype=torch.float32, device=dev)
    o_p = torch.zeros((nfr, cap, 3), dtype=torch.float32, device=dev)
    o_n = torch.zeros(nfr, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()    # the fills above run on torch's stream, the library on its own
    m.stereo_fisheye_matches_batch_raw(dKL, dKR, dn[0], dn[1], dn[2], dn[3], cap, dI, dD, cap, nfr, rig, o_ltr, o_rtl, o_z,
                                       o_p, o_st, o_n, flags=_lib.IN_DEVICE | _lib.OUT_DEVICE)
    torch.cuda.synchronize()
    for f in range(nfr):
        a, b = int(nl[f]), int(nr[f])
        got = (int(o_n[f].item()), o_ltr[f, :a].cpu().numpy(), o_rtl[f, :b].cpu().numpy(), o_z[f, :a].cpu().numpy(),
               o_p[f, :a].cpu().numpy(), o_st[f, :a].cpu().numpy())
        _close(got, refs[f], refs[f][6], rig)
        assert (o_ltr[f, a:] == -1).all() and (o_rtl[f, b:] == -1).all() and (o_z[f, a:] == -1).all()
