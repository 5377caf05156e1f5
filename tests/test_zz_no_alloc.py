"""-m gpu: orbfe_knn2_static_scratch (the kNN entry point that allocates nothing: its scratch memory is a file-scope
`static __device__` array; no counterpart in the reference) against the
existing entry point orbfe_knn2, through the C ABI, on device buffers.  Integer work: bit-exact.  Collected last on purpose
(the file name sorts after every other test file)."""
import ctypes as C

import numpy as np
import pytest

from orbslam3lib_b200 import BFMatcher, _lib, scenes

pytestmark = pytest.mark.gpu


def _static_knn(dq, nq, dt, nt, didx, ddist, stream=None):
    err = C.create_string_buffer(256)
    rc = _lib.lib().orbfe_knn2_static_scratch(_lib.ptr(dq), int(nq), _lib.ptr(dt), int(nt), _lib.ptr(didx), _lib.ptr(ddist),
                                              stream, err, len(err))
    return rc, err.value.decode()


def _existing_knn(bf, dq, nq, dt, nt):
    import torch
    idx = torch.full((nq, 2), -7, dtype=torch.int32, device=dq.device)
    dist = torch.full((nq, 2), -7, dtype=torch.int32, device=dq.device)
    torch.cuda.synchronize()    # the fills run on torch's stream, orbfe_knn2 on the matcher's own non-blocking stream
    bf.knn2_raw(dq, nq, dt, nt, idx, dist, flags=_lib.IN_DEVICE | _lib.OUT_DEVICE)
    return idx.cpu().numpy(), dist.cpu().numpy()


def test_static_scratch_knn_equals_knn2_on_two_calls_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    bf = BFMatcher()
    # two calls in a row that share the static array: a large problem (128 query tiles x 8 train segments: the whole array
    # is written), then a smaller one with other segment counts, whose tiles must not see the first call's partials / tickets
    for nq, nt, seed in ((16384, 40000, 501), (5000, 3000, 502), (300, 777, 503), (129, 100, 504)):
        q, t = scenes.map_descriptors(nq, nt, seed)
        dq, dt = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
        ref_idx, ref_dist = _existing_knn(bf, dq, nq, dt, nt)
        for rep in range(2):
            didx = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
            ddist = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
            rc, msg = _static_knn(dq, nq, dt, nt, didx, ddist)
            assert rc == 0, msg
            torch.cuda.synchronize(dev)
            np.testing.assert_array_equal(didx.cpu().numpy(), ref_idx)
            np.testing.assert_array_equal(ddist.cpu().numpy(), ref_dist)
        assert (ref_idx[:, 0] >= 0).all() and (ref_dist[:, 0] <= ref_dist[:, 1]).all()


def test_static_scratch_knn_on_a_caller_stream_and_edges():
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    bf = BFMatcher()
    q, t = scenes.map_descriptors(2000, 5000, 511)
    dq, dt = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
    ref_idx, ref_dist = _existing_knn(bf, dq, 2000, dt, 5000)
    st = torch.cuda.Stream(device=dev)
    didx = torch.zeros((2000, 2), dtype=torch.int32, device=dev)
    ddist = torch.zeros((2000, 2), dtype=torch.int32, device=dev)
    torch.cuda.synchronize(dev)
    rc, msg = _static_knn(dq, 2000, dt, 5000, didx, ddist, C.c_void_p(st.cuda_stream))
    assert rc == 0, msg
    st.synchronize()
    np.testing.assert_array_equal(didx.cpu().numpy(), ref_idx)
    np.testing.assert_array_equal(ddist.cpu().numpy(), ref_dist)
    # one train row: the second neighbour is missing; no train rows: both are
    for nt in (1, 0):
        rc, msg = _static_knn(dq, 200, dt, nt, didx, ddist)
        assert rc == 0, msg
        torch.cuda.synchronize(dev)
        got = didx[:200].cpu().numpy()
        assert (got[:, 1] == -1).all() and (got[:, 0] == (0 if nt else -1)).all()
    # beyond the stated maximum: rejected with a clear error, nothing launched
    before = _lib.lib().orbfe_kernel_launches()
    rc, msg = _static_knn(dq, _lib.KNN2_STATIC_MAX_NQ + 1, dt, 5000, didx, ddist)
    assert rc == _lib.ERR_CAPACITY and str(_lib.KNN2_STATIC_MAX_NQ) in msg and "maximum" in msg
    rc, msg = _static_knn(dq, -1, dt, 5000, didx, ddist)
    assert rc == _lib.ERR_ARG and msg
    rc, msg = _static_knn(None, 10, dt, 5000, didx, ddist)
    assert rc == _lib.ERR_ARG and "null" in msg
    assert _lib.lib().orbfe_kernel_launches() == before
    assert _static_knn(dq, 0, dt, 5000, didx, ddist)[0] == 0
