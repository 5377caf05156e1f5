"""e2e time of the one-call host->host form (orbfe_extract_match_batch: 512 pairs 752x480, pinned buffers) for the current
ORBFE_PIPE_SCHED / ORBFE_PIPE_DUAL; with ORBFE_PIPE_TRACE=1 the library prints the stage timeline of every call."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from orbslam3lib_b200 import ORBextractor
P = 512; n = 2 * P; W, H = bench.W, bench.H
imgs = bench.make_pairs(P, 1000)
pairs = torch.from_numpy(np.ascontiguousarray(np.stack([imgs[:P], imgs[P:]], 1))).pin_memory()
ex = ORBextractor(1200, 1.2, 8, 20, 7); cap = ex.max_keypoints
hk = torch.zeros((n, cap, 28), dtype=torch.uint8).pin_memory(); hd = torch.zeros((n, cap, 32), dtype=torch.uint8).pin_memory()
hc = torch.zeros(n, dtype=torch.int32).pin_memory(); hm = torch.zeros(n, dtype=torch.int32).pin_memory()
hi = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory(); hdist = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory()
def step():
    ex.extract_match_batch_raw(pairs, P, W, H, hk, hd, cap, hc, hm, hi, hdist)
reps = int(os.environ.get("REPS", "10"))
for _ in range(3): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(reps): step()
torch.cuda.synchronize()
print("%-28s dual=%s one-call e2e %.3f ms/step" % (os.environ.get("ORBFE_PIPE_SCHED", "default"), os.environ.get("ORBFE_PIPE_DUAL", "1"),
                                                   (time.perf_counter() - t0) / reps * 1e3), flush=True)
