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
