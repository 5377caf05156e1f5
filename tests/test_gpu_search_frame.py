"""-m gpu: the per-frame tracking matcher ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono)
(ORBmatcher.cc:1683-1921) and the key-frame form (:1923-2059) on the GPU (orbfe_search_by_projection_frame_batch /
_keyframe_batch) against the CPU oracle: slot chosen per point after the rotation check, final mvpMapPoints owner of every
slot and nmatches.  Index work: bit-exact."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBmatcher, _lib, scenes

pytestmark = pytest.mark.gpu


def _mono_frame(oracle, kind, w, h, seed, nf):
    ex = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    _, kps, desc = ex(scenes.make(kind, w, h, seed))
    u_right = np.where(scenes.rand_u32(len(kps), seed) % 3 == 0, -1.0, kps["x"] - 9.25).astype(np.float32)
    return kps, desc, u_right, ex.scale


@pytest.mark.parametrize("kind,w,h,seed,nf,npts,th,direction,ori", [
    ("texture", 752, 480, 8100, 1200, 1200, 7.0, 0, True),      # Tracking::TrackWithMotionModel: th = 7 (stereo) / 15 (mono)
    ("texture", 752, 480, 8101, 1200, 1200, 15.0, 0, True),
    ("texture", 752, 480, 8102, 1200, 1500, 14.0, 1, True),     # bForward
    ("texture", 640, 480, 8103, 1000, 1000, 14.0, 2, True),     # bBackward
    ("checker", 640, 480, 8104, 1000, 1500, 7.0, 0, True),      # repeated structure: distance ties
    ("noise", 320, 240, 8105, 500, 600, 15.0, 0, False),        # mbCheckOrientation off
    ("flat_blobs", 752, 480, 8106, 1000, 300, 7.0, 0, True),
])
def test_frame_search_mono_matches_oracle(oracle, kind, w, h, seed, nf, npts, th, direction, ori):
    kps, desc, u_right, scale = _mono_frame(oracle, kind, w, h, seed, nf)
    n = len(kps)
    pts = scenes.frame_points(kps, desc, n, None, u_right, 8, npts, seed, dup_frac=0.3)
    occupied = (scenes.rand_u32(n, seed + 9) % 10 == 0).astype(np.uint8)
    bounds = (0.0, 0.0, float(w), float(h))
    rl, rr, ro, rn = oracle.search_by_projection_frame(kps, desc, u_right, occupied, n, bounds, scale, 47.9, pts, th=th,
                                                       direction=direction, check_orientation=ori)
    nm, ml, mr, owner = ORBmatcher(0.9, ori).SearchByProjectionFrame(kps, desc, n, bounds, scale, pts, th, mbf=47.9,
                                                                     u_right=u_right, occupied=occupied, direction=direction)
    np.testing.assert_array_equal(ml, rl)
    np.testing.assert_array_equal(owner, ro)
    assert nm == rn and (mr == -1).all()
    if kind == "texture":
        assert nm > npts // 5
        if ori:
            assert nm < int((ml >= 0).sum()) + 1 and (rl >= 0).sum() < (pts["flags"] & 1).sum()


@pytest.mark.parametrize("seed,th,direction", [(8200, 7.0, 0), (8201, 15.0, 1), (8202, 10.0, 2)])
def test_frame_search_stereo_camera_matches_oracle(oracle, seed, th, direction):
    """Nleft != -1 (the Lynx fisheye rig): left search on mGrid, right search on mGridRight, slots of both cameras."""
    imgl = scenes.make("texture", 640, 400, seed)
    ex = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    _, kl, dl = ex(imgl)
    _, kr, dr = ex(scenes.stereo_right(imgl, seed))
    keys, desc, nleft, ltr, rtl = scenes.stereo_camera_frame(kl, dl, kr, dr, seed)
    pts = scenes.frame_points(keys, desc, nleft, ltr, None, 8, 1500, seed, dup_frac=0.3)
    occupied = (scenes.rand_u32(len(keys), seed + 9) % 12 == 0).astype(np.uint8)
    bounds = (0.0, 0.0, 640.0, 400.0)
    rl, rr, ro, rn = oracle.search_by_projection_frame(keys, desc, None, occupied, nleft, bounds, ex.scale, 0.0, pts, th=th,
                                                       direction=direction)
    nm, ml, mr, owner = ORBmatcher(0.9, True).SearchByProjectionFrame(keys, desc, nleft, bounds, ex.scale, pts, th,
                                                                      occupied=occupied, direction=direction)
    np.testing.assert_array_equal(ml, rl)
    np.testing.assert_array_equal(mr, rr)
    np.testing.assert_array_equal(owner, ro)
    assert nm == rn and (mr >= 0).sum() > 100


def test_frame_search_contended_is_sequential(oracle):
    """many points aimed at the same few features, all with observations: every commit round has conflicts."""
    kps, desc, u_right, scale = _mono_frame(oracle, "texture", 320, 240, 8300, 400)
    n = len(kps)
    base = scenes.frame_points(kps[:40], desc[:40], 40, None, u_right[:40], 8, 60, 8300, dup_frac=0.0)
    pts = np.repeat(base, 16)
    pts["flags"] |= 4
    bounds = (0.0, 0.0, 320.0, 240.0)
    rl, rr, ro, rn = oracle.search_by_projection_frame(kps, desc, u_right, None, n, bounds, scale, 47.9, pts, th=15.0)
    nm, ml, mr, owner = ORBmatcher(0.9, True).SearchByProjectionFrame(kps, desc, n, bounds, scale, pts, 15.0, mbf=47.9,
                                                                      u_right=u_right)
    np.testing.assert_array_equal(ml, rl)
    np.testing.assert_array_equal(owner, ro)
    assert nm == rn


@pytest.mark.parametrize("seed,th,dist", [(8400, 10.0, 100), (8401, 3.0, 64), (8402, 15.0, 50)])
def test_keyframe_search_matches_oracle(oracle, seed, th, dist):
    kps, desc, _, scale = _mono_frame(oracle, "texture", 752, 480, seed, 1200)
    n = len(kps)
    pts = scenes.frame_points(kps, desc, n, None, None, 8, 1500, seed, dup_frac=0.3)
    occupied = (scenes.rand_u32(n, seed + 9) % 5 == 0).astype(np.uint8)
    bounds = (-4.5, -3.0, 756.0, 484.0)
    rm, ro, rn = oracle.search_by_projection_keyframe(kps, desc, occupied, bounds, scale, pts, th=th, orb_dist=dist)
    nm, m, owner = ORBmatcher(0.9, True).SearchByProjectionKeyFrame(kps, desc, bounds, scale, pts, th, dist, occupied=occupied)
    np.testing.assert_array_equal(m, rm)
    np.testing.assert_array_equal(owner, ro)
    assert nm == rn and nm > 100
    assert (owner[occupied.astype(bool)] == -1).all()


def test_frame_search_ragged_device_batch(oracle):
    """a batch of frames with different feature / point counts, device-resident buffers."""
    import torch
    frames = [_mono_frame(oracle, "texture", 640, 480, 8500 + k, nf) for k, nf in enumerate((1000, 300, 700))]
    cap, pcap, F = max(len(f[0]) for f in frames), 900, 3
    K = np.zeros((F, cap), _lib.KP_DTYPE)
    D = np.zeros((F, cap, 32), np.uint8)
    U = np.full((F, cap), -1, np.float32)
    P = np.zeros((F, pcap), scenes.FRAME_POINT_DTYPE)
    nn, npp, dirs = np.zeros(F, np.int32), np.array([900, 0, 411], np.int32), np.array([0, 1, 2], np.int32)
    for f, (k, d, u, _) in enumerate(frames):
        nn[f] = len(k)
        K[f, :len(k)], D[f, :len(k)], U[f, :len(k)] = k, d, u
        P[f, :npp[f]] = scenes.frame_points(k, d, len(k), None, u, 8, int(npp[f]), 8510 + f)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(a.view(np.uint8).reshape(a.shape[0], -1)).to(dev)    # noqa: E731
    dK, dD, dP = t(K), t(D), t(P)
    dU, dn, dnp, ddir = (torch.from_numpy(a).to(dev) for a in (U, nn, npp, dirs))
    dml = torch.zeros((F, pcap), dtype=torch.int32, device=dev)
    dmr = torch.zeros((F, pcap), dtype=torch.int32, device=dev)
    down = torch.zeros((F, cap), dtype=torch.int32, device=dev)
    dcnt = torch.zeros(F, dtype=torch.int32, device=dev)
    m = ORBmatcher(0.9, True)
    bounds, scale = (0.0, 0.0, 640.0, 480.0), frames[0][3]
    torch.cuda.synchronize()    # the fills above run on torch's stream, the library on its own
    m.search_frame_batch_raw(dK, dD, dU, None, dn, None, ddir, cap, bounds, scale, 47.9, dP, dnp, pcap, F, 12.0, dml, dmr,
                             down, dcnt, flags=_lib.IN_DEVICE | _lib.OUT_DEVICE)
    ml, owner, cnt = dml.cpu().numpy(), down.cpu().numpy(), dcnt.cpu().numpy()
    for f, (k, d, u, _) in enumerate(frames):
        rl, rr, ro, rn = oracle.search_by_projection_frame(k, d, u, None, len(k), bounds, scale, 47.9, P[f, :npp[f]], th=12.0,
                                                           direction=int(dirs[f]))
        np.testing.assert_array_equal(ml[f, :npp[f]], rl)
        np.testing.assert_array_equal(owner[f, :len(k)], ro)
        assert cnt[f] == rn
