"""-m gpu: MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401) on the GPU (orbfe_distinctive_descriptors) against the
CPU oracle: BestIdx, BestMedian and the chosen descriptor per MapPoint.  Integer work: bit-exact."""
import numpy as np
import pytest

from orbslam3lib_b200 import ORBmatcher, _lib
from test_search_rule import observation_blocks

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,npoints,max_n", [(9900, 200, 40), (9901, 1500, 12), (9902, 64, 300), (9903, 7, 33)])
def test_distinctive_matches_oracle(oracle, seed, npoints, max_n):
    desc, off = observation_blocks(seed, npoints, max_n)
    ri, rm = oracle.distinctive_descriptors(desc, off)
    marker = np.full((npoints, 32), 0xAB, np.uint8)
    bi, bm, od = ORBmatcher(0.6, True).ComputeDistinctiveDescriptors(desc, off, out_desc=marker.copy())
    np.testing.assert_array_equal(bi, ri)
    np.testing.assert_array_equal(bm, rm)
    has = ri >= 0
    np.testing.assert_array_equal(od[has], desc[off[:-1][has] + ri[has]])
    assert (od[~has] == 0xAB).all()    # a MapPoint without observations keeps its descriptor
    assert has.sum() > npoints // 2


def test_distinctive_capacity_and_device_buffers(oracle):
    import torch
    desc, off = observation_blocks(9910, 300, 50)
    ri, rm = oracle.distinctive_descriptors(desc, off)
    n = np.diff(off)
    # max_n below the largest block: those points are reported as -1, the others are unaffected
    bi, bm, _ = ORBmatcher(0.6, True).ComputeDistinctiveDescriptors(desc, off, max_n=20)
    assert (bi[n > 20] == -1).all() and (n > 20).any()
    np.testing.assert_array_equal(bi[n <= 20], ri[n <= 20])
    dev = torch.device("cuda", 0)
    dd, doff = torch.from_numpy(desc).to(dev), torch.from_numpy(off).to(dev)
    dbi = torch.zeros(300, dtype=torch.int32, device=dev)
    dbm = torch.zeros(300, dtype=torch.int32, device=dev)
    dod = torch.zeros((300, 32), dtype=torch.uint8, device=dev)
    m = ORBmatcher(0.6, True)
    torch.cuda.synchronize()    # the fills above run on torch's stream, the library on its own
    m._bf._check(_lib.lib().orbfe_distinctive_descriptors(m._bf._h, _lib.ptr(dd), _lib.ptr(doff), 300, len(desc), 50, _lib.ptr(dbi),
                                                          _lib.ptr(dbm), _lib.ptr(dod), _lib.IN_DEVICE | _lib.OUT_DEVICE))
    np.testing.assert_array_equal(dbi.cpu().numpy(), ri)
    np.testing.assert_array_equal(dbm.cpu().numpy(), rm)
    with pytest.raises(Exception):
        ORBmatcher(0.6, True).ComputeDistinctiveDescriptors(desc, off + 1)    # offsets[0] != 0
