"""-m gpu: the CUDA extractor against the CPU oracle, stage by stage and end to end, through the C ABI.

Bars (BASELINE.json north_star): bit-exact pyramid pixels, FAST responses, keypoint (x, y, octave) sets and order
after distribution, descriptors; angles within 1e-3 degrees."""
import numpy as np
import pytest

from orbslam3lib_b200 import KP_DTYPE as KP
from orbslam3lib_b200 import ORBextractor, scenes

pytestmark = pytest.mark.gpu

ANGLE_TOL_DEG = 1e-3

CASES = [
    # kind, w, h, seed, nfeatures, nlevels
    ("texture", 640, 480, 1000, 1000, 8),
    ("texture", 752, 480, 1001, 1200, 8),
    ("checker", 640, 480, 1002, 1000, 8),
    ("noise", 320, 240, 1003, 500, 8),
    ("flat_blobs", 752, 480, 1004, 1000, 8),
    ("texture", 211, 173, 1005, 300, 5),
    ("noise", 752, 480, 1006, 1200, 8),
    ("texture", 1280, 720, 1007, 2000, 10),
]


def _compare(oracle, kind, w, h, seed, nf, nl, lap=(0, 0), stages=True):
    img = scenes.make(kind, w, h, seed)
    ref = oracle.OracleExtractor(nf, 1.2, nl, 20, 7)
    mono_r, kps_r, desc_r = ref(img, lap)
    ex = ORBextractor(nf, 1.2, nl, 20, 7)
    mono_g, kps_g, desc_g = ex(img, lap)
    if stages:
        for l in range(nl):
            assert ex.level_size(l) == ref.level_size(l), "level %d size" % l
            np.testing.assert_array_equal(ex.level_padded(l), ref.level_padded(l), "padded pyramid level %d" % l)
            cr, cg = ref.level_candidates(l), ex.level_candidates(l)
            assert cg.shape == cr.shape, "level %d: %d FAST candidates, oracle %d" % (l, len(cg), len(cr))
            np.testing.assert_array_equal(cg, cr, "FAST candidates level %d" % l)
            kr, kg = ref.level_keypoints(l), ex.level_keypoints(l)
            assert len(kg) == len(kr), "level %d: %d keypoints after the quadtree, oracle %d" % (l, len(kg), len(kr))
            for f in ("x", "y", "response", "octave", "size"):
                np.testing.assert_array_equal(kg[f], kr[f], "level %d keypoint field %s" % (l, f))
            if len(kr):
                np.testing.assert_array_equal(ex.level_blurred(l), ref.level_blurred(l), "blurred level %d" % l)
                d = np.abs(kg["angle"] - kr["angle"])
                d = np.minimum(d, 360.0 - d)
                assert d.max() <= ANGLE_TOL_DEG, "level %d angle diff %g" % (l, d.max())
    assert mono_g == mono_r
    assert len(kps_g) == len(kps_r)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        np.testing.assert_array_equal(kps_g[f], kps_r[f], "keypoint field %s" % f)
    d = np.abs(kps_g["angle"] - kps_r["angle"])
    d = np.minimum(d, 360.0 - d)
    assert len(d) == 0 or d.max() <= ANGLE_TOL_DEG
    np.testing.assert_array_equal(desc_g, desc_r, "descriptors")
    return len(kps_g), ref.sort_ties()


@pytest.mark.parametrize("kind,w,h,seed,nf,nl", CASES)
def test_extractor_matches_oracle(oracle, kind, w, h, seed, nf, nl):
    n, ties = _compare(oracle, kind, w, h, seed, nf, nl)
    assert n > 0


def test_lapping_area_packing(oracle):
    # ORBextractor_old.cc:1171-1180: keypoints inside vLappingArea go to the back
    _compare(oracle, "texture", 640, 480, 1010, 1000, 8, lap=(200, 400), stages=False)
    _compare(oracle, "texture", 640, 480, 1010, 1000, 8, lap=(0, 1000), stages=False)


def test_batch_equals_single(oracle):
    imgs = np.stack([scenes.make("texture", 640, 480, 2000 + i) for i in range(6)])
    ex = ORBextractor(1000, 1.2, 8, 20, 7)
    counts, monos, kps, desc = ex.extract_batch(imgs)
    rc, rm, rk, rd = oracle.extract_batch(imgs, 1000, 1.2, 8, 20, 7, threads=4)
    np.testing.assert_array_equal(counts, rc)
    np.testing.assert_array_equal(monos, rm)
    for i in range(len(imgs)):
        n = counts[i]
        for f in ("x", "y", "size", "response", "octave"):
            np.testing.assert_array_equal(kps[i, :n][f], rk[i, :n][f])
        np.testing.assert_array_equal(desc[i, :n], rd[i, :n])


def test_chunked_batch(oracle):
    # a tiny workspace limit forces several chunks; results must not depend on the chunking
    imgs = np.stack([scenes.make("checker", 320, 240, 3000 + i) for i in range(5)])
    ex = ORBextractor(500, 1.2, 8, 20, 7)
    a = ex.extract_batch(imgs)
    ex2 = ORBextractor(500, 1.2, 8, 20, 7)
    ex2.set_workspace_limit(3 << 20)
    b = ex2.extract_batch(imgs)
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])
    for i in range(len(imgs)):    # rows beyond counts[i] are unspecified
        n = a[0][i]
        np.testing.assert_array_equal(a[2][i, :n], b[2][i, :n])
        np.testing.assert_array_equal(a[3][i, :n], b[3][i, :n])


def test_pipelined_host_batch_and_device_outputs(oracle):
    # >= 256 host images take the pipelined path (copy-in / kernels / copy-out overlapped over 128-image stages);
    # results must equal image-by-image extraction, and the device-resident outputs must feed the matcher
    import ctypes as C
    from orbslam3lib_b200 import BFMatcher, _lib
    base = [scenes.make(k, 160, 120, 5000 + i) for i, k in enumerate(("texture", "checker", "noise", "flat_blobs", "texture"))]
    n = 300
    imgs = np.stack([base[i % len(base)] for i in range(n)])
    ex = ORBextractor(200, 1.2, 4, 20, 7)
    counts, monos, kps, desc = ex.extract_batch(imgs)
    single = ORBextractor(200, 1.2, 4, 20, 7)
    refs = [single(b) for b in base]
    for i in range(n):
        mono, rk, rd = refs[i % len(base)]
        c = int(counts[i])
        assert c == len(rk) and int(monos[i]) == mono
        np.testing.assert_array_equal(kps[i, :c], rk)
        np.testing.assert_array_equal(desc[i, :c], rd)
    # oracle on the distinct images
    for b, (mono, rk, rd) in zip(base, refs):
        omono, ok, od = oracle.OracleExtractor(200, 1.2, 4, 20, 7)(b)
        assert omono == mono
        np.testing.assert_array_equal(rd, od)
    _, dd, dc, dn, dcap = ex.device_outputs()
    assert dn == n and dcap == ex.max_keypoints
    P = n // 2
    m = BFMatcher()
    idx = np.full((P, dcap, 2), -1, np.int32)
    dist = np.full((P, dcap, 2), -1, np.int32)
    m.knn2_batch_raw(dd, dc, dcap, dd + P * dcap * 32, dc + 4 * P, dcap, P, idx, dist, flags=_lib.IN_DEVICE)
    for p in (0, 1, 77, P - 1):
        ri, rdist = oracle.knn2_hamming(desc[p, :counts[p]], desc[P + p, :counts[P + p]])
        np.testing.assert_array_equal(idx[p, :counts[p]], ri)
        np.testing.assert_array_equal(dist[p, :counts[p]], rdist)


# BASELINE.json configs as parity cases (C1 640x480 nf=1000; C3 640x480 stereo nf=500; C4 3840x2160
# nf=5000 12 levels; C2 is covered by CASES[1] and by the benchmark scene)
def test_config_c1_previous_frame_matching(oracle):
    from orbslam3lib_b200 import BFMatcher
    cur = scenes.make("texture", 640, 480, 1000)
    prev = scenes.previous_frame(cur, 1000)
    ex = ORBextractor(1000, 1.2, 8, 20, 7)
    ref = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    outs = []
    for img in (cur, prev):
        mono, kps, desc = ex(img)
        rm, rk, rd = ref(img)
        assert mono == rm
        np.testing.assert_array_equal(desc, rd)
        np.testing.assert_array_equal(kps["x"], rk["x"])
        outs.append(desc)
    idx, dist = BFMatcher().knnMatch(outs[0], outs[1])
    ri, rdist = oracle.knn2_hamming(outs[0], outs[1])
    np.testing.assert_array_equal(idx, ri)
    np.testing.assert_array_equal(dist, rdist)
    good = BFMatcher.ratio_test(idx, dist, 0.7)
    assert (good >= 0).sum() > 100    # the shifted frame really matches


def test_config_c3_stereo_500(oracle):
    left = scenes.make("texture", 640, 480, 1003)
    right = scenes.stereo_right(left, 1003)
    ex = ORBextractor(500, 1.2, 8, 20, 7)
    counts, monos, kps, desc = ex.extract_batch(np.stack([left, right]))
    rc, rm, rk, rd = oracle.extract_batch(np.stack([left, right]), 500, 1.2, 8, 20, 7, threads=2)
    np.testing.assert_array_equal(counts, rc)
    for i in range(2):
        n = counts[i]
        np.testing.assert_array_equal(desc[i, :n], rd[i, :n])
        for f in ("x", "y", "octave", "response"):
            np.testing.assert_array_equal(kps[i, :n][f], rk[i, :n][f])


def test_config_c4_4k_12_levels(oracle):
    n, ties = _compare(oracle, "texture", 3840, 2160, 1004, 5000, 12)
    assert n >= 5000


def test_config_c2_full_batch_properties(oracle):
    """BASELINE configs[1] at full size: 512 stereo pairs of 752x480, 1200 features, through the host (pipelined) path
    and through the rectified-stereo entry point.  Size-independent properties: the batch is 8 distinct pairs repeated,
    so every output must repeat with period 8 (batch position, pipeline stage and workspace slot must not matter); the
    8 distinct pairs equal the oracle; a checksum over all descriptors equals 64 x the checksum of one period."""
    P, w, h, nf = 512, 752, 480, 1200
    distinct = 8
    imgs = np.empty((P, 2, h, w), np.uint8)
    for k in range(distinct):
        imgs[k, 0] = scenes.make("texture", w, h, 9100 + k)
        imgs[k, 1] = scenes.stereo_right_rectified(imgs[k, 0], 9100 + k)
    for p in range(distinct, P):
        imgs[p] = imgs[p % distinct]
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    mb, mbf = 0.11, 0.11 * 435.2
    counts, kps, desc, u_right, depth = ex.extract_stereo_batch(imgs, mb, mbf)
    assert counts.shape == (P, 2) and counts.min() > 1000
    for p in range(distinct, P):
        q = p % distinct
        np.testing.assert_array_equal(counts[p], counts[q])
        for e in range(2):
            n = int(counts[q, e])
            np.testing.assert_array_equal(kps[p, e, :n], kps[q, e, :n])
            np.testing.assert_array_equal(desc[p, e, :n], desc[q, e, :n])
        np.testing.assert_array_equal(u_right[p].view(np.uint32), u_right[q].view(np.uint32))
        np.testing.assert_array_equal(depth[p].view(np.uint32), depth[q].view(np.uint32))
    for q in range(distinct):
        eL, eR = oracle.OracleExtractor(nf, 1.2, 8, 20, 7), oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
        _, kl, dl = eL(imgs[q, 0])
        _, kr, dr = eR(imgs[q, 1])
        assert (int(counts[q, 0]), int(counts[q, 1])) == (len(kl), len(kr))
        np.testing.assert_array_equal(desc[q, 0, :len(kl)], dl)
        np.testing.assert_array_equal(desc[q, 1, :len(kr)], dr)
        np.testing.assert_array_equal(kps[q, 0, :len(kl)]["x"], kl["x"])
        u, z, _, faults = oracle.stereo_matches(eL, eR, kl, dl, kr, dr, mb, mbf)
        assert faults == 0
        np.testing.assert_array_equal(u_right[q, :len(kl)].view(np.uint32), u.view(np.uint32))
        np.testing.assert_array_equal(depth[q, :len(kl)].view(np.uint32), z.view(np.uint32))

    def checksum(lo, hi):
        s = 0
        for p in range(lo, hi):
            for e in range(2):
                s += int(desc[p, e, :counts[p, e]].astype(np.uint64).sum())
        return s
    assert checksum(0, P) == (P // distinct) * checksum(0, distinct)


@pytest.mark.parametrize("npairs,lap", [(3, (0, 0)), (70, (100, 500)), (300, (0, 0))])
def test_extract_match_batch(oracle, npairs, lap):
    """orbfe_extract_match_batch: extraction of both eyes + left -> right kNN per pair in one call, through the chunked
    (3 pairs), and the pipelined two-stream host path (70 / 300 pairs); every pair against the oracle's extraction and kNN."""
    from orbslam3lib_b200 import scenes
    w, h, nf = (752, 480, 1200) if npairs <= 70 else (320, 240, 400)
    base = [scenes.make("texture", w, h, 9600 + k) for k in range(min(npairs, 6))]
    pairs = np.stack([np.stack([base[k % len(base)], scenes.stereo_right(base[k % len(base)], 9600 + k % len(base))])
                      for k in range(npairs)])
    pairs[-1, 1] = 0    # a right image without any keypoint: every match of that pair is -1
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    counts, monos, kps, desc, idx, dist = ex.extract_match_batch(pairs, lapping=lap)
    cap = ex.max_keypoints
    refs = {}
    for p in list(range(min(npairs, len(base)))) + [npairs - 1]:
        out = []
        for e in range(2):
            key = (p % len(base), e) if p != npairs - 1 or e == 0 else ("blank", 1)
            if key not in refs:
                refs[key] = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)(pairs[p, e], lap)
            out.append(refs[key])
        for e, (mono, rk, rd) in enumerate(out):
            n = int(counts[p, e])
            assert n == len(rk) and int(monos[p, e]) == mono
            np.testing.assert_array_equal(desc[p, e, :n], rd)
            np.testing.assert_array_equal(kps[p, e, :n]["x"], rk["x"])
        nl, nr = len(out[0][1]), len(out[1][1])
        ri, rdist = oracle.knn2_hamming(out[0][2], out[1][2]) if nl else (np.zeros((0, 2), np.int32),) * 2
        np.testing.assert_array_equal(idx[p, :nl], ri)
        np.testing.assert_array_equal(dist[p, :nl], rdist)
        assert (idx[p, nl:] == -1).all() and (dist[p, nl:] == -1).all()
    assert (idx[npairs - 1] == -1).all()
    # every other pair repeats one of the checked ones
    for p in range(len(base), npairs - 1):
        q = p % len(base)
        assert np.array_equal(counts[p], counts[q]) and np.array_equal(idx[p], idx[q]) and np.array_equal(dist[p], dist[q])


def test_small_batch_cuda_graph_path(oracle):
    """BASELINE configs[2] shape: one 640x480 stereo pair per call, 500 features, PINNED host buffers.  The second call with
    the same buffers captures the call into a CUDA graph, later calls replay it; every call must return the oracle's
    keypoints, descriptors and matches, also when the images in the (same) buffer change and when a ring of buffers is
    used (one graph per slot)."""
    import torch
    from orbslam3lib_b200 import scenes
    w, h, nf = 640, 480, 500
    ex = ORBextractor(nf, 1.2, 8, 20, 7)
    cap = ex.max_keypoints
    ring = [torch.zeros((2, h, w), dtype=torch.uint8).pin_memory() for _ in range(3)]
    kps = torch.zeros((2, cap, 28), dtype=torch.uint8).pin_memory()
    desc = torch.zeros((2, cap, 32), dtype=torch.uint8).pin_memory()
    counts = torch.zeros(2, dtype=torch.int32).pin_memory()
    monos = torch.zeros(2, dtype=torch.int32).pin_memory()
    idx = torch.zeros((1, cap, 2), dtype=torch.int32).pin_memory()
    dist = torch.zeros((1, cap, 2), dtype=torch.int32).pin_memory()
    ref = oracle.OracleExtractor(nf, 1.2, 8, 20, 7)
    launches = []
    for it in range(9):
        left = scenes.make("texture" if it % 4 else "checker", w, h, 9700 + it)
        pair = np.stack([left, scenes.stereo_right(left, 9700 + it)])
        buf = ring[it % 3]
        buf.copy_(torch.from_numpy(pair))
        ex.extract_match_batch_raw(buf, 1, w, h, kps, desc, cap, counts, monos, idx, dist, lapping=(0, 0))
        launches.append(ex.last_launches())
        exp = [ref(pair[e]) for e in range(2)]
        k = kps.numpy().view(KP).reshape(2, cap)
        for e in range(2):
            mono, rk, rd = exp[e]
            n = int(counts[e])
            assert n == len(rk) and int(monos[e]) == mono, (it, e)
            np.testing.assert_array_equal(desc.numpy()[e, :n], rd)
            np.testing.assert_array_equal(k[e, :n]["x"], rk["x"])
            np.testing.assert_array_equal(k[e, :n]["octave"], rk["octave"])
        ri, rdist = oracle.knn2_hamming(exp[0][2], exp[1][2])
        np.testing.assert_array_equal(idx.numpy()[0, :len(ri)], ri)
        np.testing.assert_array_equal(dist.numpy()[0, :len(ri)], rdist)
    assert min(launches) > 10    # replayed graphs still report the kernels they launch


@pytest.mark.parametrize("scale,nl,w,h,nf", [(1.1, 8, 333, 251, 400), (1.5, 5, 752, 480, 800), (1.33, 4, 160, 150, 100),
                                             (1.2, 3, 1000, 90, 300), (1.25, 7, 517, 389, 500)])
def test_other_scale_factors_and_sizes_single_and_large_batch(oracle, scale, nl, w, h, nf):
    """scaleFactor / nlevels other than the 1.2 / 8 of every shipped configuration, odd image sizes (levels narrower than one
    resize tile, tiles whose source window is a single 16-byte vector), through the single-image kernels (4 rows per thread)
    and the large-batch ones (>= 32 images: 48-row resize tiles): whole output and every padded pyramid plane against the
    oracle."""
    _compare(oracle, "texture", w, h, 8100, nf, nl)    # the same size at scale 1.2 (what _compare builds)
    kinds = ("texture", "checker", "noise")
    base = [scenes.make(kinds[i % 3], w, h, 8100 + i) for i in range(3)]
    n = 35
    imgs = np.stack([base[i % 3] for i in range(n)])
    ex = ORBextractor(nf, scale, nl, 20, 7)
    counts, monos, kps, desc = ex.extract_batch(imgs)
    refs = []
    for b in base:
        r = oracle.OracleExtractor(nf, scale, nl, 20, 7)
        refs.append((r, r(b)))
    for i in (0, 1, 2, n - 2, n - 1):
        r, (mono, rk, rd) = refs[i % 3]
        c = int(counts[i])
        assert c == len(rk) and int(monos[i]) == mono
        np.testing.assert_array_equal(desc[i, :c], rd)
        for f in ("x", "y", "size", "response", "octave"):
            np.testing.assert_array_equal(kps[i, :c][f], rk[f])
        for l in range(nl):
            np.testing.assert_array_equal(ex.level_padded(l, index=i), r.level_padded(l), "image %d level %d" % (i, l))
    single = ORBextractor(nf, scale, nl, 20, 7)
    mono, sk, sd = single(base[1])
    np.testing.assert_array_equal(sd, refs[1][1][2])
    for l in range(nl):
        np.testing.assert_array_equal(single.level_padded(l), refs[1][0].level_padded(l), "single image level %d" % l)


def test_random_geometries(oracle):
    """seeded random image sizes, feature counts, level counts and scale factors (every level at least 61 px on both sides, the
    smallest the reference's cell arithmetic accepts): the table-driven resize tiles, the FAST work items, the blur strips and
    the TMA boxes of the descriptor stage at image borders, single image and a 33-image batch, against the oracle."""
    rng = np.random.RandomState(20260)
    done = 0
    while done < 10:
        w, h = int(rng.randint(120, 1000)), int(rng.randint(100, 720))
        scale = float(rng.choice([1.1, 1.2, 1.2, 1.25, 1.4]))
        nl = int(rng.randint(2, 9))
        if min(w, h) / scale ** (nl - 1) < 61:
            continue
        nf = int(rng.randint(50, 1500))
        kind = ("texture", "checker", "noise", "flat_blobs")[done % 4]
        img = scenes.make(kind, w, h, 8200 + done)
        ref = oracle.OracleExtractor(nf, scale, nl, 20, 7)
        mono_r, kps_r, desc_r = ref(img, (0, 0))
        ex = ORBextractor(nf, scale, nl, 20, 7)
        mono_g, kps_g, desc_g = ex(img, (0, 0))
        tag = "%s %dx%d nf=%d nl=%d scale=%g" % (kind, w, h, nf, nl, scale)
        assert mono_g == mono_r and len(kps_g) == len(kps_r), tag
        np.testing.assert_array_equal(desc_g, desc_r, tag)
        for f in ("x", "y", "size", "response", "octave"):
            np.testing.assert_array_equal(kps_g[f], kps_r[f], tag)
        if done % 3 == 0:    # the large-batch kernels on the same geometry
            counts, monos, kps, desc = ex.extract_batch(np.stack([img] * 33))
            for i in (0, 16, 32):
                c = int(counts[i])
                assert c == len(kps_r) and int(monos[i]) == mono_r, tag
                np.testing.assert_array_equal(desc[i, :c], desc_r, tag)
        done += 1


@pytest.mark.parametrize("env", [{"ORBFE_DESCRIBE_TMA": "0"}, {"ORBFE_DESCRIBE_V1": "1"}, {"ORBFE_FAST_TMAP": "0"},
                                 {"ORBFE_RESIZE_TILE": "1"}, {"ORBFE_KNN_KEYED": "0"}])
def test_alternative_extractor_kernels_still_agree(env):
    """the kernels the defaults replaced stay selectable (A/B measurements, and the fallbacks taken when a tensor map cannot
    be encoded): describe_fused_kernel / describe_kernel instead of describe_tma_kernel, FAST strips by per-row bulk copies,
    resize_tile_kernel, the unkeyed per-pair kNN.  Same outputs as the defaults on a 40-pair batch (large-batch kernels) and a
    single pair; the switches are read once per process, hence the subprocesses."""
    import os
    import subprocess
    import sys
    code = (
        "import numpy as np, sys, zlib\n"
        "from orbslam3lib_b200 import ORBextractor, scenes\n"
        "base = [scenes.make(k, 333, 251, 8300 + i) for i, k in enumerate(('texture', 'checker', 'noise'))]\n"
        "pairs = np.stack([np.stack([base[i % 3], scenes.stereo_right(base[i % 3], 8300 + i % 3)]) for i in range(40)])\n"
        "ex = ORBextractor(400, 1.2, 6, 20, 7)\n"
        "h = 0\n"
        "for batch in (pairs, pairs[:1]):\n"
        "    counts, monos, kps, desc, idx, dist = ex.extract_match_batch(batch)\n"
        "    for p in range(len(batch)):\n"
        "        for e in range(2):\n"
        "            c = int(counts[p, e])\n"
        "            h = zlib.crc32(desc[p, e, :c].tobytes(), h); h = zlib.crc32(kps[p, e, :c].tobytes(), h)\n"
        "        c = int(counts[p, 0])\n"
        "        h = zlib.crc32(idx[p, :c].tobytes(), h); h = zlib.crc32(dist[p, :c].tobytes(), h)\n"
        "    h = zlib.crc32(counts.tobytes(), h); h = zlib.crc32(monos.tobytes(), h)\n"
        "sys.stdout.write('%d %d' % (h, int(counts.sum())))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for e in ({}, env):
        full = dict(os.environ, **e)
        for k in ("ORBFE_DESCRIBE_TMA", "ORBFE_DESCRIBE_V1", "ORBFE_FAST_TMAP", "ORBFE_RESIZE_TILE", "ORBFE_KNN_KEYED"):
            if k not in e:
                full.pop(k, None)
        outs.append(subprocess.run([sys.executable, "-c", code], cwd=root, env=full, capture_output=True, text=True,
                                   timeout=300))
    assert outs[0].returncode == 0 and outs[1].returncode == 0, outs[1].stderr[-400:]
    assert outs[0].stdout == outs[1].stdout and int(outs[0].stdout.split()[1]) > 100    # keypoints of the single pair


@pytest.mark.parametrize("w,h,nl", [(752, 480, 3), (640, 488, 4), (333, 251, 2)])
def test_scale_factor_two_takes_the_box_average(oracle, w, h, nl):
    """scaleFactor 2.0: where a level is an exact 2x decimation of the one below, cv::resize(INTER_LINEAR) is INTER_AREA (2x2
    box average, halve_kernel); odd sizes stay bilinear (333 -> 167 is not exact).  Whole output and every padded plane, single
    image and a 33-image batch."""
    img = scenes.make("texture", w, h, 8400)
    ref = oracle.OracleExtractor(500, 2.0, nl, 20, 7)
    mono_r, kps_r, desc_r = ref(img, (0, 0))
    ex = ORBextractor(500, 2.0, nl, 20, 7)
    mono_g, kps_g, desc_g = ex(img, (0, 0))
    assert mono_g == mono_r and len(kps_g) == len(kps_r) and len(kps_r) > 50
    np.testing.assert_array_equal(desc_g, desc_r)
    for f in ("x", "y", "size", "response", "octave"):
        np.testing.assert_array_equal(kps_g[f], kps_r[f])
    for l in range(nl):
        assert ex.level_size(l) == ref.level_size(l)
        np.testing.assert_array_equal(ex.level_padded(l), ref.level_padded(l), "padded pyramid level %d" % l)
    counts, monos, kps, desc = ex.extract_batch(np.stack([img] * 33))
    for i in (0, 32):
        c = int(counts[i])
        assert c == len(kps_r)
        np.testing.assert_array_equal(desc[i, :c], desc_r)
        for l in range(nl):
            np.testing.assert_array_equal(ex.level_padded(l, index=i), ref.level_padded(l), "image %d level %d" % (i, l))
