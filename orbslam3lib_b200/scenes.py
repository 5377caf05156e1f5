"""Seeded synthetic scenes for parity tests and the benchmark.

Everything derives from a splitmix64 counter hash, so a scene is a pure function of (kind, w, h, seed).
The reference has no fixtures for this path (dsp/debug_scripts/prepare_input.py:16-17 names two PNGs that are
not in the repo), so these scenes are the shared inputs of the oracle and of the CUDA path.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(idx, seed):
    """Counter-based hash: element `idx` (array of uint64) of stream `seed` -> uint64."""
    with np.errstate(over="ignore"):
        z = np.asarray(idx, dtype=np.uint64) + np.uint64((int(seed) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def rand_u8(n, seed, offset=0):
    return (splitmix64(np.arange(offset, offset + n, dtype=np.uint64), seed) >> np.uint64(56)).astype(np.uint8)


def rand_u32(n, seed, offset=0):
    return (splitmix64(np.arange(offset, offset + n, dtype=np.uint64), seed) >> np.uint64(32)).astype(np.uint32)


def noise(w, h, seed):
    return rand_u8(w * h, seed).reshape(h, w)


def _signed_noise(w, h, seed, amp):
    r = rand_u8(w * h, seed ^ 0x5EED).reshape(h, w).astype(np.int32)
    return r % (2 * amp + 1) - amp


def checker(w, h, seed, period=16, amp=8):
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.where(((xx // period) + (yy // period)) % 2 == 0, 64, 192).astype(np.int32)
    return np.clip(base + _signed_noise(w, h, seed, amp), 0, 255).astype(np.uint8)


def _box3(img):
    p = np.pad(img.astype(np.int32), 1, mode="edge")
    s = sum(p[dy:dy + img.shape[0], dx:dx + img.shape[1]] for dy in range(3) for dx in range(3))
    return ((s + 4) // 9).astype(np.uint8)


def _rects(w, h, seed, n, base, smooth):
    img = np.full((h, w), base, np.uint8)
    r = rand_u32(5 * n, seed)
    for i in range(n):
        x0 = int(r[5 * i] % w)
        y0 = int(r[5 * i + 1] % h)
        rw = 4 + int(r[5 * i + 2] % max(8, w // 6))
        rh = 4 + int(r[5 * i + 3] % max(8, h // 6))
        img[y0:y0 + rh, x0:x0 + rw] = r[5 * i + 4] & 0xFF
    return _box3(img) if smooth else img


def texture(w, h, seed):
    """Random gray rectangles painted in seed order, then a 3x3 box filter: multi-scale corners, some empty cells."""
    n = max(8, int(round(400.0 * (w * h) / 360960.0)))
    return _rects(w, h, seed, n, 128, True)


def flat_blobs(w, h, seed):
    """Constant 128 with 20 rectangles: mostly empty FAST cells (minThFAST fallback, levels with < N candidates)."""
    return _rects(w, h, seed, 20, 128, False)


def stereo_right(left, seed):
    """Right view: row y shifted left by the integer disparity d(y) = 4 + (y mod 37), edge replicated, noise +-2."""
    h, w = left.shape
    yy, xx = np.mgrid[0:h, 0:w]
    src = np.clip(xx + (4 + (yy % 37)), 0, w - 1)
    out = left[yy, src].astype(np.int32) + _signed_noise(w, h, seed ^ 0xA11CE, 2)
    return np.clip(out, 0, 255).astype(np.uint8)


def stereo_right_rectified(left, seed, band=48):
    """Right view of a rectified pair for Frame::ComputeStereoMatches (Frame.cc:827-997): the disparity is constant
    inside horizontal bands of `band` rows, d = 4 + (7 * (y // band)) mod 37, so 11x11 windows away from the band edges
    match exactly up to the +-2 noise; edge replicated."""
    h, w = left.shape
    yy, xx = np.mgrid[0:h, 0:w]
    src = np.clip(xx + (4 + (7 * (yy // band)) % 37), 0, w - 1)
    out = left[yy, src].astype(np.int32) + _signed_noise(w, h, seed ^ 0x5E7E0, 2)
    return np.clip(out, 0, 255).astype(np.uint8)


def previous_frame(cur, seed):
    """Previous frame of C1: the current frame shifted by (+3, +2) px, edge replicated, noise +-2."""
    h, w = cur.shape
    yy, xx = np.mgrid[0:h, 0:w]
    out = cur[np.clip(yy - 2, 0, h - 1), np.clip(xx - 3, 0, w - 1)].astype(np.int32)
    return np.clip(out + _signed_noise(w, h, seed ^ 0xB0B, 2), 0, 255).astype(np.uint8)


TRACK_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_xr", "<f4"), ("view_cos", "<f4"),
                        ("track_depth", "<f4"), ("level", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])


def track_points(kps, desc, u_right, nlevels, npts, seed, jitter=3.0, max_flips=30, dup_frac=0.15):
    """Synthetic local-map points for ORBmatcher::SearchByProjection (ORBmatcher.cc:44-143): point i is a frame feature
    seen again -- projected at the feature's position + uniform jitter, predicted level = octave or octave + 1,
    descriptor = the feature's with 0..max_flips random bit flips, mTrackProjXR = the feature's uRight + jitter.  A
    fraction of the points are duplicates of an earlier point's feature (two points competing for one feature: the
    sequential already-matched rule decides), some are not in view, bad or far.  Deterministic in `seed`."""
    n = len(kps)
    pts = np.zeros(npts, TRACK_DTYPE)
    if n == 0 or npts == 0:
        return pts
    r = rand_u32(npts * 8, seed).reshape(8, npts)
    feat = (r[0] % np.uint32(n)).astype(np.int64)
    ndup = int(npts * dup_frac)
    if ndup > 0 and npts > 1:
        dst = (rand_u32(ndup, seed + 1) % np.uint32(npts - 1)).astype(np.int64) + 1
        feat[dst] = feat[dst - 1]
    u = lambda a: (a.astype(np.float64) / 4294967296.0)    # noqa: E731
    pts["proj_x"] = (kps["x"][feat] + (u(r[1]) * 2 - 1) * jitter).astype(np.float32)
    pts["proj_y"] = (kps["y"][feat] + (u(r[2]) * 2 - 1) * jitter).astype(np.float32)
    ur = u_right[feat] if u_right is not None else np.full(npts, -1.0, np.float32)
    pts["proj_xr"] = (ur + (u(r[3]) * 2 - 1) * jitter).astype(np.float32)
    pts["view_cos"] = (0.9975 + u(r[4]) * 0.0025).astype(np.float32)
    pts["track_depth"] = (1.0 + u(r[5]) * 39.0).astype(np.float32)
    pts["level"] = np.minimum(kps["octave"][feat] + (r[6] & 1).astype(np.int32), nlevels - 1)
    sel = r[7] % np.uint32(100)
    pts["flags"] = np.where(sel < 4, 0, 1) | np.where((sel >= 4) & (sel < 7), 2, 0) | np.where(sel < 93, 4, 0)
    d = desc[feat].copy()
    nflip = (rand_u32(npts, seed + 2) % np.uint32(max_flips + 1)).astype(np.int64)
    pos = (rand_u32(npts * max_flips, seed + 3) % np.uint32(256)).reshape(npts, max_flips).astype(np.int64)
    for k in range(max_flips):
        rows = np.nonzero(nflip > k)[0]
        p = pos[rows, k]
        d[rows, p >> 3] ^= (1 << (p & 7)).astype(np.uint8)
    pts["desc"] = d
    return pts


TRACK_STEREO_DTYPE = np.dtype([("proj_x", "<f4"), ("proj_y", "<f4"), ("proj_xr", "<f4"), ("proj_yr", "<f4"),
                               ("view_cos", "<f4"), ("view_cos_r", "<f4"), ("track_depth", "<f4"), ("level", "<i4"),
                               ("level_r", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])


def stereo_camera_frame(kps_l, desc_l, kps_r, desc_r, seed, max_dist=60):
    """A stereo-camera frame (Nleft != -1) for SearchByProjection: features = left then right, and a consistent pair of
    mvLeftToRightMatch / mvRightToLeftMatch tables (mutual nearest descriptors below max_dist; -1 elsewhere).
    -> (keys (N,), desc (N, 32), nleft, left_to_right, right_to_left)"""
    nl, nr = len(kps_l), len(kps_r)
    keys = np.concatenate([kps_l, kps_r])
    desc = np.concatenate([desc_l, desc_r])
    ltr = np.full(nl, -1, np.int32)
    rtl = np.full(nr, -1, np.int32)
    if nl and nr:
        bl = np.unpackbits(desc_l, axis=1).astype(np.int16)
        br = np.unpackbits(desc_r, axis=1).astype(np.int16)
        dist = bl.sum(1)[:, None] + br.sum(1)[None, :] - 2 * (bl @ br.T)
        best_r = dist.argmin(1)
        best_l = dist.argmin(0)
        for i in range(nl):
            j = int(best_r[i])
            if best_l[j] == i and dist[i, j] < max_dist:
                ltr[i], rtl[j] = j, i
    return keys, desc, nl, ltr, rtl


def track_points_stereo(keys, desc, nleft, ltr, nlevels, npts, seed, jitter=3.0, max_flips=30, dup_frac=0.2):
    """Local-map points for the stereo-camera branch: point i re-observes left feature f (projection = its position +
    jitter, level = octave or + 1) and, in the right camera, f's stereo partner when it has one (else the left position
    shifted by 12 px); flags: bit 0 in view left, bit 3 in view right, bit 1 bad, bit 2 has observations."""
    pts = np.zeros(npts, TRACK_STEREO_DTYPE)
    if nleft == 0 or npts == 0:
        return pts
    r = rand_u32(npts * 10, seed).reshape(10, npts)
    feat = (r[0] % np.uint32(nleft)).astype(np.int64)
    ndup = int(npts * dup_frac)
    if ndup > 0 and npts > 1:
        dst = (rand_u32(ndup, seed + 1) % np.uint32(npts - 1)).astype(np.int64) + 1
        feat[dst] = feat[dst - 1]
    u = lambda a: (a.astype(np.float64) / 4294967296.0)    # noqa: E731
    part = ltr[feat]
    has = part >= 0
    pr = np.where(has, part + nleft, feat)
    pts["proj_x"] = (keys["x"][feat] + (u(r[1]) * 2 - 1) * jitter).astype(np.float32)
    pts["proj_y"] = (keys["y"][feat] + (u(r[2]) * 2 - 1) * jitter).astype(np.float32)
    pts["proj_xr"] = (keys["x"][pr] - np.where(has, 0.0, 12.0) + (u(r[3]) * 2 - 1) * jitter).astype(np.float32)
    pts["proj_yr"] = (keys["y"][pr] + (u(r[8]) * 2 - 1) * jitter).astype(np.float32)
    pts["view_cos"] = (0.9975 + u(r[4]) * 0.0025).astype(np.float32)
    pts["view_cos_r"] = (0.9975 + u(r[9]) * 0.0025).astype(np.float32)
    pts["track_depth"] = (1.0 + u(r[5]) * 39.0).astype(np.float32)
    pts["level"] = np.minimum(keys["octave"][feat] + (r[6] & 1).astype(np.int32), nlevels - 1)
    lr = np.minimum(keys["octave"][pr] + ((r[6] >> 1) & 1).astype(np.int32), nlevels - 1)
    sel = r[7] % np.uint32(100)
    pts["level_r"] = np.where(sel % 17 == 0, -1, lr)
    in_l = np.where(sel % 11 == 0, 0, 1)
    in_r = np.where(sel % 7 == 0, 0, 8)
    pts["flags"] = in_l | in_r | np.where((sel >= 4) & (sel < 7), 2, 0) | np.where(sel < 93, 4, 0)
    d = desc[feat].copy()
    nflip = (rand_u32(npts, seed + 2) % np.uint32(max_flips + 1)).astype(np.int64)
    pos = (rand_u32(npts * max_flips, seed + 3) % np.uint32(256)).reshape(npts, max_flips).astype(np.int64)
    for k in range(max_flips):
        rows = np.nonzero(nflip > k)[0]
        p = pos[rows, k]
        d[rows, p >> 3] ^= (1 << (p & 7)).astype(np.uint8)
    pts["desc"] = d
    return pts


# == orbfe_frame_point / orc_frame_point
FRAME_POINT_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("ur", "<f4"), ("vr", "<f4"), ("invz", "<f4"), ("angle", "<f4"),
                              ("level", "<i4"), ("flags", "<i4"), ("desc", "u1", (32,))])
assert FRAME_POINT_DTYPE.itemsize == 64


def frame_points(keys, desc, nleft, ltr, u_right, nlevels, npts, seed, jitter=4.0, max_flips=40, dup_frac=0.2, mbf=47.9,
                 rot_deg=7.0, outlier_rot_frac=0.12):
    """MapPoints of the LAST frame as ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) sees them
    (ORBmatcher.cc:1722-1917), synthesised from the CURRENT frame's features: point i re-observes left feature f --
    projection = its position + jitter, right-camera projection = f's stereo partner when there is one, level = f's
    octave -1 / 0 / +1, angle = f's angle + a common in-plane rotation (a fraction gets a random angle: the rotation
    histogram has losers), descriptor = f's with 0..max_flips bit flips, 1/z consistent with f's mvuRight when it has
    one (so the mvuRight gate passes for most and fails for some).  Duplicates compete for one feature; some points
    have no observations (they do not block), some do not reach the search, some project outside the image."""
    pts = np.zeros(npts, FRAME_POINT_DTYPE)
    if nleft == 0 or npts == 0:
        return pts
    r = rand_u32(npts * 12, seed).reshape(12, npts)
    feat = (r[0] % np.uint32(nleft)).astype(np.int64)
    ndup = int(npts * dup_frac)
    if ndup > 0 and npts > 1:
        dst = (rand_u32(ndup, seed + 1) % np.uint32(npts - 1)).astype(np.int64) + 1
        feat[dst] = feat[dst - 1]
    u = lambda a: (a.astype(np.float64) / 4294967296.0)    # noqa: E731
    pts["u"] = (keys["x"][feat] + (u(r[1]) * 2 - 1) * jitter).astype(np.float32)
    pts["v"] = (keys["y"][feat] + (u(r[2]) * 2 - 1) * jitter).astype(np.float32)
    if ltr is not None:
        part = ltr[feat]
        has = part >= 0
        pr = np.where(has, part + nleft, feat)
        pts["ur"] = (keys["x"][pr] - np.where(has, 0.0, 12.0) + (u(r[3]) * 2 - 1) * jitter).astype(np.float32)
        pts["vr"] = (keys["y"][pr] + (u(r[4]) * 2 - 1) * jitter).astype(np.float32)
    disp = np.full(npts, 8.0)
    if u_right is not None:
        ur = u_right[feat]
        disp = np.where(ur > 0, keys["x"][feat] - ur, 8.0)
    # u - mbf * invz ~ mvuRight +- a few pixels; every 9th point is far off (gated out)
    off = (u(r[5]) * 2 - 1) * 2.0 + np.where(r[6] % np.uint32(9) == 0, 40.0, 0.0)
    pts["invz"] = np.maximum((disp + off) / mbf, 1e-3).astype(np.float32)
    ang = keys["angle"][feat].astype(np.float64) + rot_deg + (u(r[7]) * 2 - 1) * 3.0
    ang = np.where(u(r[8]) < outlier_rot_frac, u(r[9]) * 360.0, ang)
    pts["angle"] = np.mod(ang, 360.0).astype(np.float32)
    pts["level"] = np.clip(keys["octave"][feat] + (r[10] % np.uint32(3)).astype(np.int32) - 1, 0, nlevels - 1)
    sel = r[11] % np.uint32(100)
    pts["flags"] = np.where(sel < 5, 0, 1) | np.where(sel < 85, 4, 0)
    edge = sel == 50    # a few projections outside the image bounds
    pts["u"] = np.where(edge, -3.0, pts["u"]).astype(np.float32)
    d = desc[feat].copy()
    nflip = (rand_u32(npts, seed + 2) % np.uint32(max_flips + 1)).astype(np.int64)
    pos = (rand_u32(npts * max_flips, seed + 3) % np.uint32(256)).reshape(npts, max_flips).astype(np.int64)
    for k in range(max_flips):
        rows = np.nonzero(nflip > k)[0]
        q = pos[rows, k]
        d[rows, q >> 3] ^= (1 << (q & 7)).astype(np.uint8)
    pts["desc"] = d
    return pts


def feature_vector(desc, nnodes=96, order_seed=0):
    """A stand-in for DBoW2's FeatureVector (vocabulary node at levelsup 4 -> indices of the features under it): the node
    of a descriptor is a function of a few of its bits, so re-observed descriptors with a few flipped bits mostly stay in
    their node.  Returned as CSR: (node_ids ascending int32, start int32 (len + 1), idx int32) with the features of a node
    in ascending index order (DBoW2 appends them in feature order)."""
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    node = ((d[:, 0].astype(np.int64) >> 2) * 7 + (d[:, 13].astype(np.int64) >> 5) + order_seed) % nnodes
    ids = np.unique(node)
    order = np.argsort(node, kind="stable")
    counts = np.array([(node == i).sum() for i in ids], np.int64)
    start = np.zeros(len(ids) + 1, np.int32)
    start[1:] = np.cumsum(counts)
    return ids.astype(np.int32), start, order.astype(np.int32)


# == orbfe_fisheye_rig / orc_fisheye_rig
FISHEYE_RIG_DTYPE = np.dtype([("cam1", "<f4", 8), ("cam2", "<f4", 8), ("Rlr", "<f4", 9), ("tlr", "<f4", 3),
                              ("precision", "<f4"), ("nlevels", "<i4"), ("level_sigma2", "<f4", 16)])
assert FISHEYE_RIG_DTYPE.itemsize == 184


def fisheye_rig(w=640, h=400, baseline=0.11, seed=0, nlevels=8, scale=1.2):
    """a Kannala-Brandt stereo rig like the Lynx / TUM-VI ones: ~190 px focal length, small relative rotation."""
    r = rand_u32(16, seed ^ 0xF15E).astype(np.float64) / 2.0**32 - 0.5
    rig = np.zeros((), FISHEYE_RIG_DTYPE)
    for cam, o in (("cam1", 0), ("cam2", 4)):
        rig[cam] = (190.9 + 4 * r[o], 190.2 + 4 * r[o + 1], w / 2 + 6 * r[o + 2], h / 2 + 6 * r[o + 3],
                    0.0034, 0.0007, -0.0020, 0.0002)
    ax = 0.02 * r[8:11]                                   # Rodrigues vector of Rlr
    th = float(np.linalg.norm(ax))
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]]) / max(th, 1e-12)
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    rig["Rlr"] = R.reshape(-1)
    rig["tlr"] = (baseline, 0.004 * r[11], 0.004 * r[12])
    rig["precision"] = 1e-6
    rig["nlevels"] = nlevels
    rig["level_sigma2"][:nlevels] = (np.float32(scale) ** np.arange(nlevels, dtype=np.float32)) ** 2
    return rig


def _kb_unproject64(prm, x, y):
    px, py = (x - prm[2]) / prm[0], (y - prm[3]) / prm[1]
    td = np.minimum(np.sqrt(px * px + py * py), np.pi / 2)
    th = td.copy()
    for _ in range(12):
        t2 = th * th
        f = th * (1 + prm[4] * t2 + prm[5] * t2**2 + prm[6] * t2**3 + prm[7] * t2**4) - td
        df = 1 + 3 * prm[4] * t2 + 5 * prm[5] * t2**2 + 7 * prm[6] * t2**3 + 9 * prm[7] * t2**4
        th = th - f / df
    sc = np.where(td > 1e-8, np.tan(th) / np.maximum(td, 1e-12), 1.0)
    return np.stack([px * sc, py * sc, np.ones_like(px)], 1)


def _kb_project64(prm, X):
    th = np.arctan2(np.sqrt(X[:, 0]**2 + X[:, 1]**2), X[:, 2])
    psi = np.arctan2(X[:, 1], X[:, 0])
    r = th + prm[4] * th**3 + prm[5] * th**5 + prm[6] * th**7 + prm[7] * th**9
    return np.stack([prm[0] * r * np.cos(psi) + prm[2], prm[1] * r * np.sin(psi) + prm[3]], 1)


def fisheye_stereo_case(kps_left, rig, seed, mono_left=0, mono_right=0, n_right=None):
    """Right keypoints and a BFMatchORB result for a left keypoint set seen by `rig`: every stereo-area left keypoint
    gets a depth (0.3 m .. 60 m, a few at infinity = no parallax), is re-projected into camera 2 with 0 .. 4 px of noise
    (some fail the chi-square gate), a few partners are behind the camera or shared by two left keypoints, and the
    descriptor distances cover 0 ("bugged point"), < 70 and >= 70.  -> (kps_right, knn_idx (ns, 2), knn_dist (ns, 2))"""
    from ._lib import KP_DTYPE
    nl = len(kps_left)
    ns = nl - mono_left
    n_right = ns + mono_right if n_right is None else n_right
    nsr = n_right - mono_right
    u = rand_u32(8 * max(ns, 1), seed ^ 0xF1E).reshape(8, -1).astype(np.float64) / 2.0**32
    kl = kps_left[mono_left:]
    c1, c2 = rig["cam1"].astype(np.float64), rig["cam2"].astype(np.float64)
    R12, t12 = rig["Rlr"].astype(np.float64).reshape(3, 3), rig["tlr"].astype(np.float64)
    ray = _kb_unproject64(c1, kl["x"].astype(np.float64), kl["y"].astype(np.float64))
    z = 0.3 * (200.0 ** u[0])                               # log-uniform 0.3 .. 60 m
    z = np.where(u[1] < 0.05, 1e5, z)                       # no parallax
    z = np.where((u[1] >= 0.05) & (u[1] < 0.08), -z, z)     # behind camera 1: lands anywhere
    X1 = ray * z[:, None]
    X2 = (X1 - t12) @ R12                                   # R21 * (X1 - t12), R21 = R12^T
    uv = _kb_project64(c2, X2)
    amp = np.where(u[2] < 0.7, 0.4, 4.0)                    # most matches are sub-pixel accurate
    uv = uv + amp[:, None] * (np.stack([u[3], u[4]], 1) - 0.5) * 2
    kr = np.zeros(n_right, KP_DTYPE)
    kr["x"][:mono_right] = 5.0
    kr["y"][:mono_right] = 5.0
    idx = np.full((ns, 2), -1, np.int32)
    dist = np.full((ns, 2), -1, np.int32)
    if nsr > 0 and ns > 0:
        perm = (rand_u32(ns, seed ^ 0xBEEF) % np.uint32(nsr)).astype(np.int64)     # collisions: shared right partners
        first = np.full(nsr, -1, np.int64)
        for i in range(ns - 1, -1, -1):
            first[perm[i]] = i                              # the right keypoint sits where its FIRST left partner projects
        has = first >= 0
        kr["x"][mono_right:][has] = uv[first[has], 0].astype(np.float32)
        kr["y"][mono_right:][has] = uv[first[has], 1].astype(np.float32)
        kr["octave"][mono_right:] = (rand_u32(nsr, seed ^ 0x0C7) % 8).astype(np.int32)
        idx[:, 0] = perm
        idx[:, 1] = (perm + 1) % nsr
        d = (u[5] * 90).astype(np.int32) + 1
        d = np.where(u[6] < 0.02, 0, d)                     # bugged points
        dist[:, 0] = d
        dist[:, 1] = d + (u[7] * 40).astype(np.int32)
    return kr, idx, dist


def kb8_keyframe_pair(rig, w, h, seed, npts=900, nlevels=8):
    """Two key frames of a two-camera KannalaBrandt8 rig that see the same 3-D points: the data of
    ORBmatcher::SearchForTriangulation on stereo-camera key frames.  Key frame 2 is key frame 1 moved by ~0.2 m and a small
    rotation.  Every point gets a base descriptor; each of its observations (left / right camera of either key frame, each
    present with probability ~0.7) is the projection + up to 0.5 px of noise with a few flipped descriptor bits; a tenth of
    the observations of key frame 2 are moved by 6 - 40 px (same descriptor, wrong geometry: the constraint must reject them).
    -> dict(keys1, desc1, nleft1, keys2, desc2, nleft2, t12 (4, 12): R12 | t12 of Tll, Tlr, Trl, Trr)"""
    from ._lib import KP_DTYPE
    u = rand_u32(16 * npts, seed ^ 0x7B8).reshape(16, npts).astype(np.float64) / 2.0**32
    c1, c2 = rig["cam1"].astype(np.float64), rig["cam2"].astype(np.float64)
    Rlr, tlr = rig["Rlr"].astype(np.float64).reshape(3, 3), rig["tlr"].astype(np.float64)
    Rrl, trl = Rlr.T, -Rlr.T @ tlr
    # points in the left camera of key frame 1
    px, py = 30 + u[0] * (w - 60), 30 + u[1] * (h - 60)
    z = 0.5 * (16.0 ** u[2])
    X1 = _kb_unproject64(c1, px, py) * z[:, None]
    ax = np.array([0.01, -0.02, 0.015])    # key frame 2's left camera: x_l1 = M_R x_l2 + M_t
    th = float(np.linalg.norm(ax))
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]]) / th
    MR = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    Mt = np.array([0.18, 0.03, 0.06])
    comp = lambda Ra, ta, Rb, tb: (Ra @ Rb, Ra @ tb + ta)    # noqa: E731  (a o b)(x) = Ra (Rb x + tb) + ta
    Tll = (MR, Mt)
    Tlr = comp(MR, Mt, Rlr, tlr)
    Trl = comp(Rrl, trl, MR, Mt)
    Trr = comp(Rrl, trl, *Tlr)
    t12 = np.stack([np.concatenate([R.reshape(-1), t]) for R, t in (Tll, Tlr, Trl, Trr)]).astype(np.float32)
    Xl2 = (X1 - Mt) @ MR            # MR^T (X1 - Mt)
    cams = {"l1": (c1, X1), "r1": (c2, X1 @ Rrl.T + trl), "l2": (c1, Xl2), "r2": (c2, Xl2 @ Rrl.T + trl)}
    base = rand_u32(npts * 8, seed + 1).reshape(npts, 8).view(np.uint8).reshape(npts, 32)
    octave = (rand_u32(npts, seed + 2) % np.uint32(4)).astype(np.int32)
    out = {"t12": t12}
    for kf, sides, col in (("1", ("l1", "r1"), 3), ("2", ("l2", "r2"), 8)):
        ks, ds = [], []
        for si, side in enumerate(sides):
            cam, X = cams[side]
            uv = _kb_project64(cam, X)
            noise = (np.stack([u[col + 2 * si], u[col + 2 * si + 1]], 1) - 0.5)
            if kf == "2":
                bad = u[12 + si] < 0.1
                noise = np.where(bad[:, None], np.sign(noise) * (6 + 68 * np.abs(noise)), noise)
            uv = uv + noise
            vis = (u[14 + si] < 0.7) & (X[:, 2] > 0.05) & (uv[:, 0] > 5) & (uv[:, 0] < w - 5) & (uv[:, 1] > 5) & (uv[:, 1] < h - 5)
            ids = np.nonzero(vis)[0]
            k = np.zeros(len(ids), KP_DTYPE)
            k["x"], k["y"] = uv[ids, 0].astype(np.float32), uv[ids, 1].astype(np.float32)
            k["octave"] = octave[ids]
            k["size"] = 31.0
            k["angle"] = (rand_u32(npts, seed + 3)[ids] % np.uint32(360)).astype(np.float32) + (7.0 if kf == "2" else 0.0)
            k["class_id"] = -1
            d = base[ids].copy()
            fl = rand_u32(len(ids) * 6, seed + 10 * int(kf) + si).reshape(len(ids), 6)
            nfl = fl[:, 0] % np.uint32(6)
            for j in range(5):
                rows = np.nonzero(nfl > j)[0]
                q = (fl[rows, j + 1] % np.uint32(256)).astype(np.int64)
                d[rows, q >> 3] ^= (1 << (q & 7)).astype(np.uint8)
            ks.append(k)
            ds.append(d)
        out["keys" + kf], out["desc" + kf], out["nleft" + kf] = np.concatenate(ks), np.concatenate(ds), len(ks[0])
    out["sigma2"] = ((np.float32(1.2) ** np.arange(nlevels, dtype=np.float32)) ** 2).astype(np.float32)
    return out


SCENES = {"noise": noise, "checker": checker, "texture": texture, "flat_blobs": flat_blobs}


def make(kind, w, h, seed):
    return np.ascontiguousarray(SCENES[kind](w, h, seed))


def descriptors(n, seed):
    """n iid 256-bit descriptors, uint8 (n, 32)."""
    return rand_u8(n * 32, seed).reshape(n, 32)


def map_descriptors(nq, nt, seed, max_flips=40, dup_frac=0.01):
    """C5 workload: train iid (1 % of the rows duplicated verbatim -> exact ties), query i = a train row with
    0..max_flips random bit flips.  Returns (query, train)."""
    train = descriptors(nt, seed)
    ndup = int(nt * dup_frac)
    if ndup > 0 and nt > 1:
        src = (rand_u32(ndup, seed + 1) % np.uint32(nt)).astype(np.int64)
        dst = (rand_u32(ndup, seed + 2) % np.uint32(nt)).astype(np.int64)
        train[dst] = train[src]
    pick = (rand_u32(nq, seed + 3) % np.uint32(nt)).astype(np.int64)
    query = train[pick].copy()
    nflip = (rand_u32(nq, seed + 4) % np.uint32(max_flips + 1)).astype(np.int64)
    pos = (rand_u32(nq * max_flips, seed + 5) % np.uint32(256)).reshape(nq, max_flips).astype(np.int64)
    for k in range(max_flips):
        m = nflip > k
        rows = np.nonzero(m)[0]
        p = pos[rows, k]
        query[rows, p >> 3] ^= (1 << (p & 7)).astype(np.uint8)
    return query, train
