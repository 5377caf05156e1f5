"""-m gpu: the sharded map-scale matcher through the C ABI (orbfe_knn2_sharded).  world = 1 always; world = 2 over
NCCL when the box has two GPUs (one process per GPU, torchrun)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from orbslam3lib_b200 import BFMatcher, ShardedMatcher, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_world1_matches_oracle(oracle):
    import torch
    nq, nt = 5000, 30011
    q, t = scenes.map_descriptors(nq, nt, 31)
    dev = torch.device("cuda", 0)
    d_q, d_t = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
    idx = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
    dist = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
    sm = ShardedMatcher(BFMatcher(), 0, 1, 0)
    ms = sm.knn2(d_q, nq, d_t, nt, idx, dist)
    assert ms > 0
    sel = np.arange(0, nq, 13)
    ri, rd = oracle.knn2_hamming(q[sel], t, threads=4)
    np.testing.assert_array_equal(idx.cpu().numpy()[sel], ri)
    np.testing.assert_array_equal(dist.cpu().numpy()[sel], rd)


_SEG_WORKER = r"""
import os, sys
import numpy as np, torch
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "oracle", "py"))
import oracle_lib
from orbslam3lib_b200 import BFMatcher, ShardedMatcher, scenes
dev = torch.device("cuda", 0)
sm = ShardedMatcher(BFMatcher(), 0, 1, 0)
for nq, nt, seed in ((5000, 30011, 31), (777, 16384, 32), (40000, 70001, 33), (5000, 30011, 31)):
    q, t = scenes.map_descriptors(nq, nt, seed)
    if seed == 33:          # ties across segment boundaries: copies of one row spread over the whole train set
        t[::4999] = t[7]
        q[:50] = t[7]
    d_q, d_t = torch.from_numpy(q).to(dev), torch.from_numpy(t).to(dev)
    idx = torch.full((nq, 2), -9, dtype=torch.int32, device=dev); dd = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
    sm.knn2(d_q, nq, d_t, nt, idx, dd)
    tm = sm.last_times()
    assert tm["total_ms"] > 0 and tm["kernel_ms"] > 0
    sel = np.unique(np.concatenate([np.arange(0, nq, 41), np.arange(min(nq, 64)), np.arange(max(0, nq - 64), nq)]))
    ri, rd = oracle_lib.knn2_hamming(q[sel], t, threads=4)
    assert np.array_equal(idx.cpu().numpy()[sel], ri) and np.array_equal(dd.cpu().numpy()[sel], rd), (nq, nt)
print("segments ok")
"""


@pytest.mark.parametrize("chunks", ["4", "1", "7"])
def test_sharded_multi_launch_segments_single_gpu(tmp_path, oracle, chunks):
    """The chunked multi-launch kNN of the sharded matcher (train pieces matched by separate launches that share one
    segment scratch, k_knn_tc.cu: launch_knn2_tc_segments) on ONE GPU: ORBFE_SHARD_FORCE_SEG=1 takes that path at
    world = 1, so the driver's single-GPU box covers it.  Ties across segment and launch boundaries included; the same
    matcher is reused (tickets must come back to zero)."""
    script = tmp_path / "seg_worker.py"
    script.write_text(_SEG_WORKER)
    env = dict(os.environ, ORBFE_SHARD_FORCE_SEG="1", ORBFE_SHARD_CHUNKS=chunks)
    out = subprocess.run([sys.executable, str(script), ROOT], capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0 and "segments ok" in out.stdout, out.stdout[-2000:] + out.stderr[-4000:]


_WORKER = r"""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "oracle", "py"))
import oracle_lib
from orbslam3lib_b200 import BFMatcher, ShardedMatcher, scenes, shard_range
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
nq, nt = 20001, 40000
q, t = scenes.map_descriptors(nq, nt, 77)
b, c = shard_range(nq, world, rank)
d_q = torch.from_numpy(q[b:b + c].copy()).to(dev)
d_t = torch.from_numpy(t).to(dev) if rank == 0 else torch.zeros((nt, 32), dtype=torch.uint8, device=dev)
idx = torch.full((nq, 2), -9, dtype=torch.int32, device=dev); dd = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
sm = ShardedMatcher(BFMatcher(device=local), rank, world, local, dist)
sm.knn2(d_q, nq, d_t, nt, idx, dd)
assert np.array_equal(d_t.cpu().numpy(), t), "train broadcast"
sel = np.arange(rank, nq, 37)
ri, rd = oracle_lib.knn2_hamming(q[sel], t, threads=4)
assert np.array_equal(idx.cpu().numpy()[sel], ri) and np.array_equal(dd.cpu().numpy()[sel], rd)
# large enough for the overlapped path (train pieces on a second stream, segments per piece, one all-gather); uneven
# shards; run twice on the same matcher
nq, nt = 80003, 50000
q, t = scenes.map_descriptors(nq, nt, 78)
b, c = shard_range(nq, world, rank)
d_q = torch.from_numpy(q[b:b + c].copy()).to(dev)
for rep in range(2):
    d_t = torch.from_numpy(t).to(dev) if rank == 0 else torch.zeros((nt, 32), dtype=torch.uint8, device=dev)
    idx = torch.full((nq, 2), -9, dtype=torch.int32, device=dev); dd = torch.full((nq, 2), -9, dtype=torch.int32, device=dev)
    sm.knn2(d_q, nq, d_t, nt, idx, dd)
    assert np.array_equal(d_t.cpu().numpy(), t), "train broadcast (pieces)"
    sel = np.unique(np.concatenate([np.arange(rank, nq, 97), np.arange(nq // 2 - 40, nq // 2 + 40), np.arange(nq - 40, nq)]))
    ri, rd = oracle_lib.knn2_hamming(q[sel], t, threads=4)
    assert np.array_equal(idx.cpu().numpy()[sel], ri) and np.array_equal(dd.cpu().numpy()[sel], rd)
    assert (idx.cpu().numpy() >= 0).all()
sm.close(); dist.destroy_process_group(); print("rank", rank, "ok")
"""


def test_sharded_world2_nccl(tmp_path, oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29741", str(script), ROOT],
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert out.stdout.count("ok") == 2
