"""CPU (not gpu): host-side logic of the sharded map-scale matcher — the shard ranges exported by the
C library and, with two gloo processes, the partition / exchange scheme: every rank matches its query shard against
the whole train set (global indices) and the gathered result must equal the unsharded kNN, ties included.  The
per-shard arithmetic here is the CPU oracle (test infrastructure); on the GPU box the same scheme runs through
orbfe_knn2_sharded (tests/test_gpu_sharded.py)."""
import os
import subprocess
import sys

import numpy as np

from orbslam3lib_b200 import shard_range

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_partition_the_rows():
    for n in (0, 1, 7, 1000, 1000003):
        for world in (1, 2, 3, 8):
            pos = 0
            for r in range(world):
                b, c = shard_range(n, world, r)
                assert b == pos and c >= 0 and abs(c - n / world) < 1.0 + 1e-9
                pos += c
            assert pos == n


_WORKER = r"""
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "oracle", "py"))
import oracle_lib
from orbslam3lib_b200 import scenes, shard_range
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
nq, nt = 1001, 777
q, t = scenes.map_descriptors(nq, nt, 42)
train = torch.from_numpy(t.copy()) if rank == 0 else torch.zeros((nt, 32), dtype=torch.uint8)
dist.broadcast(train, src=0)                                   # train set: root -> everyone
b, c = shard_range(nq, world, rank)
idx, d = oracle_lib.knn2_hamming(q[b:b + c], train.numpy())    # shard against the FULL train set: global indices
idx_all = torch.full((nq, 2), -7, dtype=torch.int32); dist_all = torch.full((nq, 2), -7, dtype=torch.int32)
idx_all[b:b + c] = torch.from_numpy(idx); dist_all[b:b + c] = torch.from_numpy(d)
for r in range(world):                                         # one broadcast per shard (uneven shards allowed)
    rb, rc = shard_range(nq, world, r)
    if rc:
        dist.broadcast(idx_all[rb:rb + rc], src=r); dist.broadcast(dist_all[rb:rb + rc], src=r)
ri, rd = oracle_lib.knn2_hamming(q, t)
assert np.array_equal(idx_all.numpy(), ri) and np.array_equal(dist_all.numpy(), rd), "rank %d" % rank
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_sharded_scheme_two_gloo_ranks(tmp_path, oracle):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29731", str(script), ROOT],
                         capture_output=True, text=True, env=env, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert out.stdout.count("ok") == 2
