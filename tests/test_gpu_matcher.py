"""-m gpu: Hamming kNN (k = 2) against the CPU oracle through the C ABI: bit-exact indices and distances, ties
broken by the lowest train index (OpenCV's order)."""
import numpy as np
import pytest

from orbslam3lib_b200 import BFMatcher, scenes

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nq,nt", [(1, 1), (1, 2), (7, 5), (500, 500), (1000, 1000), (1200, 1200), (300, 4099), (4096, 20000)])
def test_knn2_matches_oracle(oracle, nq, nt):
    q, t = scenes.map_descriptors(nq, nt, 77 + nq)
    idx, dist = BFMatcher().knnMatch(q, t)
    ri, rd = oracle.knn2_hamming(q, t, threads=4)
    np.testing.assert_array_equal(idx, ri)
    np.testing.assert_array_equal(dist, rd)


def test_knn2_ties_lowest_index(oracle):
    t = scenes.descriptors(2000, 5)
    t[700] = t[5]
    t[1500] = t[5]
    q = t[[5, 5, 1500]].copy()
    idx, dist = BFMatcher().knnMatch(q, t)
    assert idx[0].tolist() == [5, 700] and dist[0].tolist() == [0, 0]
    assert idx[2].tolist() == [5, 700]
    # all-equal train set
    t2 = np.repeat(t[:1], 300, axis=0)
    idx, dist = BFMatcher().knnMatch(t[:3], t2)
    assert (idx == np.array([0, 1])).all()


def test_knn2_empty_and_short():
    m = BFMatcher()
    q = scenes.descriptors(4, 1)
    idx, dist = m.knnMatch(q, np.zeros((0, 32), np.uint8))
    assert (idx == -1).all() and (dist == -1).all()
    idx, dist = m.knnMatch(q, q[:1])
    assert (idx[:, 0] == 0).all() and (idx[:, 1] == -1).all() and (dist[:, 1] == -1).all()
    idx, dist = m.knnMatch(np.zeros((0, 32), np.uint8), q)
    assert idx.shape == (0, 2)


def test_knn2_batch_matches_single(oracle):
    p, cap = 6, 640
    rng = np.random.default_rng(3)
    nq = rng.integers(1, cap, p).astype(np.int32)
    nt = rng.integers(2, cap, p).astype(np.int32)
    nq[0], nt[0] = cap, cap
    q = scenes.descriptors(p * cap, 11).reshape(p, cap, 32)
    t = scenes.descriptors(p * cap, 12).reshape(p, cap, 32)
    idx, dist = BFMatcher().knnMatchBatch(q, nq, t, nt)
    for k in range(p):
        ri, rd = oracle.knn2_hamming(q[k, :nq[k]], t[k, :nt[k]])
        np.testing.assert_array_equal(idx[k, :nq[k]], ri)
        np.testing.assert_array_equal(dist[k, :nq[k]], rd)
        assert (idx[k, nq[k]:] == -1).all()


def test_large_kernel_path(oracle):
    # enough query rows for the thread-per-query kernel (>= 296 CTAs), checked on a sample of rows
    q, t = scenes.map_descriptors(80000, 3000, 9)
    idx, dist = BFMatcher().knnMatch(q, t)
    sel = np.arange(0, 80000, 97)
    ri, rd = oracle.knn2_hamming(q[sel], t, threads=4)
    np.testing.assert_array_equal(idx[sel], ri)
    np.testing.assert_array_equal(dist[sel], rd)


def test_ratio_test():
    idx = np.array([[3, 9], [4, -1], [5, 6]], np.int32)
    dist = np.array([[10, 20], [5, -1], [70, 100]], np.int32)
    out = BFMatcher.ratio_test(idx, dist, 0.7)
    assert out.tolist() == [3, -1, -1]    # 10 < 14; no second neighbour; 70 < 70 is false


def test_tensor_core_path_ties_and_ragged(oracle):
    """the IMMA kernel (taken from 296 CTAs up): duplicate and all-equal train rows (ties -> lowest index, also across
    the lanes of a quad and across shared-memory tiles), train counts that are no multiple of 8 or 256, partial query
    tiles, all-zero and all-one descriptors (|a| = 0 / 256)."""
    p, cap = 320, 300
    rng = np.random.default_rng(17)
    nq = rng.integers(1, cap + 1, p).astype(np.int32)
    nt = rng.integers(1, cap + 1, p).astype(np.int32)
    nq[:4], nt[:4] = (cap, 257, 1, 255), (cap, 263, 2, 1)
    q = scenes.descriptors(p * cap, 21).reshape(p, cap, 32)
    t = scenes.descriptors(p * cap, 22).reshape(p, cap, 32)
    t[0, 100:200] = t[0, 7]                 # a run of identical train rows
    q[0, :50] = t[0, 7]
    t[1, :] = t[1, 0]                       # all-equal train set
    q[5, 0], q[5, 1] = 0, 255
    t[5, 3], t[5, 4] = 0, 255
    idx, dist = BFMatcher().knnMatchBatch(q, nq, t, nt)
    for k in list(range(8)) + list(range(8, p, 13)):
        ri, rd = oracle.knn2_hamming(q[k, :nq[k]], t[k, :nt[k]])
        np.testing.assert_array_equal(idx[k, :nq[k]], ri)
        np.testing.assert_array_equal(dist[k, :nq[k]], rd)
    assert idx[0, 0].tolist() == [7, 100] and dist[0, 0].tolist() == [0, 0]


def test_tensor_core_path_large_single(oracle):
    q, t = scenes.map_descriptors(100001, 5003, 31, dup_frac=0.05)
    idx, dist = BFMatcher().knnMatch(q, t)
    sel = np.concatenate([np.arange(0, 100001, 211), np.arange(99990, 100001)])
    ri, rd = oracle.knn2_hamming(q[sel], t, threads=4)
    np.testing.assert_array_equal(idx[sel], ri)
    np.testing.assert_array_equal(dist[sel], rd)


@pytest.mark.parametrize("env", [{"ORBFE_KNN_TC": "0"}, {"ORBFE_KNN_TC": "0", "ORBFE_KNN_MMA": "0"}])
def test_alternative_kernels_still_agree(env):
    """the mma.sync IMMA kernel and the POPC kernel stay selectable for A/B measurements: same answers as the default
    tcgen05 kernel on a ragged batch with ties (the selection is read once per process, hence the subprocess)."""
    import os
    import subprocess
    import sys
    code = (
        "import numpy as np, sys\n"
        "from orbslam3lib_b200 import BFMatcher, scenes\n"
        "p, cap = 320, 300\n"
        "rng = np.random.default_rng(5)\n"
        "nq = rng.integers(1, cap + 1, p).astype(np.int32); nt = rng.integers(1, cap + 1, p).astype(np.int32)\n"
        "q = scenes.descriptors(p * cap, 41).reshape(p, cap, 32); t = scenes.descriptors(p * cap, 42).reshape(p, cap, 32)\n"
        "t[0, 100:200] = t[0, 7]; q[0, :50] = t[0, 7]\n"
        "idx, dist = BFMatcher().knnMatchBatch(q, nq, t, nt)\n"
        "sys.stdout.write('%d %d' % (int((idx.astype(np.int64) * 31 + dist).sum()), int(idx[0, 0, 1])))\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for e in ({}, env):
        full = dict(os.environ, **e)
        for k in ("ORBFE_KNN_TC", "ORBFE_KNN_MMA"):
            if k not in e:
                full.pop(k, None)
        outs.append(subprocess.run([sys.executable, "-c", code], cwd=root, env=full, capture_output=True, text=True,
                                   timeout=300))
    assert outs[0].returncode == 0 and outs[1].returncode == 0, outs[1].stderr[-400:]
    assert outs[0].stdout == outs[1].stdout and outs[0].stdout.split()[1] == "100"


def test_tensor_core_path_super_blocks(oracle):
    """more than 2^21 train rows: keys are re-based per super block; a tie that straddles the boundary keeps the lower
    train index first."""
    nt = (1 << 21) + 1500
    q = scenes.descriptors(80000, 51)
    t = scenes.descriptors(nt, 52)
    t[(1 << 21) - 1] = q[3]          # last row of super block 0
    t[(1 << 21) + 7] = q[3]          # and a copy in super block 1
    t[(1 << 21) + 9] = q[11]
    t[5] = q[11]
    idx, dist = BFMatcher().knnMatch(q, t)
    assert idx[3].tolist() == [(1 << 21) - 1, (1 << 21) + 7] and dist[3].tolist() == [0, 0]
    assert idx[11].tolist() == [5, (1 << 21) + 9] and dist[11].tolist() == [0, 0]
    sel = np.arange(0, 80000, 1250)
    ri, rd = oracle.knn2_hamming(q[sel], t, threads=8)
    np.testing.assert_array_equal(idx[sel], ri)
    np.testing.assert_array_equal(dist[sel], rd)


def test_tensor_core_path_train_segments(oracle):
    """one large problem on few query tiles: the launcher cuts the train set into segments (grid.z) to fill the last
    wave of CTAs and the last CTA of a tile merges the partial top 2 in segment order.  Forced here with
    ORBFE_KNN_SEGS (read per call); ties that straddle a segment boundary keep the lower train index first."""
    import os
    nq, nt = 80000, 40001          # >= 296 x 256 query rows: the tcgen05 path
    q = scenes.descriptors(nq, 61)
    t = scenes.descriptors(nt, 62)
    seg3 = ((nt + 2) // 3 + 127) // 128 * 128          # segment length for 3 segments
    t[seg3 - 1] = q[5]
    t[seg3] = q[5]                                     # exact copies on both sides of the first boundary
    t[2 * seg3 + 5] = q[9]
    t[17] = q[9]
    t[nt - 1] = q[77]                                  # best match in the last (short) segment
    old = os.environ.get("ORBFE_KNN_SEGS")
    try:
        os.environ["ORBFE_KNN_SEGS"] = "1"
        base_i, base_d = BFMatcher().knnMatch(q, t)
        for segs in (3, 8, 64):
            os.environ["ORBFE_KNN_SEGS"] = str(segs)
            idx, dist = BFMatcher().knnMatch(q, t)
            np.testing.assert_array_equal(idx, base_i)
            np.testing.assert_array_equal(dist, base_d)
    finally:
        if old is None:
            os.environ.pop("ORBFE_KNN_SEGS", None)
        else:
            os.environ["ORBFE_KNN_SEGS"] = old
    assert base_i[5].tolist() == [seg3 - 1, seg3] and base_d[5].tolist() == [0, 0]
    assert base_i[9].tolist() == [17, 2 * seg3 + 5] and base_d[9].tolist() == [0, 0]
    assert base_i[77, 0] == nt - 1 and base_d[77, 0] == 0
    sel = np.arange(0, nq, 937)
    ri, rd = oracle.knn2_hamming(q[sel], t, threads=8)
    np.testing.assert_array_equal(base_i[sel], ri)
    np.testing.assert_array_equal(base_d[sel], rd)


def test_knn2_c5_shape_sample(oracle):
    """BASELINE configs[4] / SURVEY §8(d): the map-scale problem at its full size, 1M query x 1M train 256-bit
    descriptors, k = 2.  All 1M result rows come from the tcgen05 kernel; a 4096-row sample (every 244th row, plus the
    first and last 128 so that both ends of the shard are covered) is compared with the CPU oracle bit for bit, and a
    64-row subsample with cv2.BFMatcher where cv2 imports (it refuses 2^18 or more train rows, so on a 262143-row train
    prefix).  Size-independent properties on ALL rows: best <= second, indices in range and distinct, distances <= 256."""
    nq = nt = 1000000
    q, t = scenes.map_descriptors(nq, nt, 9000)
    idx, dist = BFMatcher().knnMatch(q, t)
    assert idx.shape == (nq, 2)
    assert (dist[:, 0] <= dist[:, 1]).all() and (dist >= 0).all() and (dist <= 256).all()
    assert (idx >= 0).all() and (idx < nt).all() and (idx[:, 0] != idx[:, 1]).all()
    sel = np.unique(np.concatenate([np.arange(130, nq, 260)[:3840], np.arange(128), np.arange(nq - 128, nq)]))
    assert len(sel) == 4096
    ri, rd = oracle.knn2_hamming(q[sel], t, threads=oracle.hardware_threads())
    np.testing.assert_array_equal(idx[sel], ri)
    np.testing.assert_array_equal(dist[sel], rd)
    try:
        import cv2
    except ImportError:
        return
    ntc = 262143
    gi, gd = BFMatcher().knnMatch(q[sel[:64]], t[:ntc])
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q[sel[:64]], t[:ntc], k=2)
    np.testing.assert_array_equal(gi, np.array([[d.trainIdx for d in row] for row in m], np.int32))
    np.testing.assert_array_equal(gd, np.array([[int(d.distance) for d in row] for row in m], np.int32))
