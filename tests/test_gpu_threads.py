"""-m gpu: independent handles on distinct host threads (include/orbfe_c.h: "calls on one handle are serialised by an
internal mutex; distinct handles are independent").  Two extractors (own streams, own
workspaces) run orbfe_extract_match_batch at the same time from two Python threads (ctypes drops the GIL for the call), on
the pipelined host path and on the chunked one; every result must equal the same call made alone.  Integer work: bit-exact.
This is the usage bench.py's `e2e` times: two batches in flight so that the copy-in of one overlaps the tail of the other."""
import threading

import numpy as np
import pytest

from orbslam3lib_b200 import BFMatcher, ORBextractor, scenes

pytestmark = pytest.mark.gpu


def _pairs(npairs, w, h, seed):
    base = [scenes.make("texture", w, h, seed + k) for k in range(6)]
    return np.stack([np.stack([base[k % 6], scenes.stereo_right(base[k % 6], seed + k % 6)]) for k in range(npairs)])


def _same(a, b):
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x, y)


@pytest.mark.parametrize("npairs,w,h,nf", [(160, 752, 480, 1200), (5, 640, 480, 500)])
def test_two_extractors_on_two_threads(npairs, w, h, nf):
    inputs = [_pairs(npairs, w, h, 7100), _pairs(npairs, w, h, 7200)[::-1].copy()]
    alone = []
    for k in range(2):
        ex = ORBextractor(nf, 1.2, 8, 20, 7)
        alone.append(ex.extract_match_batch(inputs[k]))
        assert int(alone[k][0].min()) > 0
    exs = [ORBextractor(nf, 1.2, 8, 20, 7) for _ in range(2)]
    got, errs = [[None] * 3, [None] * 3], []
    gate = threading.Barrier(2)

    def worker(k):
        try:
            for rep in range(3):
                gate.wait(timeout=120)    # both calls of a round start together
                got[k][rep] = exs[k].extract_match_batch(inputs[k])
        except Exception as e:    # noqa: BLE001 - reported below, in the main thread
            errs.append(e)
            gate.abort()

    ts = [threading.Thread(target=worker, args=(k,)) for k in range(2)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=600)
    assert not errs, errs
    for k in range(2):
        for rep in range(3):
            _same(got[k][rep], alone[k])


def test_matcher_and_extractor_on_two_threads():
    """a BFMatcher (map-scale tensor-core kNN, its own stream) next to an extractor on another thread"""
    q, t = scenes.map_descriptors(20000, 60000, 7300)
    bf0 = BFMatcher()
    ref_idx, ref_dist = bf0.knnMatch(q, t)
    pairs = _pairs(40, 752, 480, 7400)
    ex0 = ORBextractor(1200, 1.2, 8, 20, 7)
    ref = ex0.extract_match_batch(pairs)
    bf, ex = BFMatcher(), ORBextractor(1200, 1.2, 8, 20, 7)
    out, errs = {}, []

    def run_knn():
        try:
            for rep in range(3):
                out["knn%d" % rep] = bf.knnMatch(q, t)
        except Exception as e:    # noqa: BLE001
            errs.append(e)

    def run_ex():
        try:
            for rep in range(3):
                out["ex%d" % rep] = ex.extract_match_batch(pairs)
        except Exception as e:    # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=run_knn), threading.Thread(target=run_ex)]
    for t_ in ts:
        t_.start()
    for t_ in ts:
        t_.join(timeout=600)
    assert not errs, errs
    for rep in range(3):
        np.testing.assert_array_equal(out["knn%d" % rep][0], ref_idx)
        np.testing.assert_array_equal(out["knn%d" % rep][1], ref_dist)
        _same(out["ex%d" % rep], ref)
