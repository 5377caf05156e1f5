"""e2e time of one host->host extraction call (1024 images 752x480, pinned buffers) for the current ORBFE_PIPE_SCHED."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from orbslam3lib_b200 import BFMatcher, ORBextractor, _lib
P = 512; n = 2 * P; W, H = bench.W, bench.H
imgs = torch.from_numpy(bench.make_pairs(P, 1000)).pin_memory()
ex = ORBextractor(1200, 1.2, 8, 20, 7); mt = BFMatcher(); cap = ex.max_keypoints
hk = torch.zeros((n, cap, 28), dtype=torch.uint8).pin_memory(); hd = torch.zeros((n, cap, 32), dtype=torch.uint8).pin_memory()
hc = torch.zeros(n, dtype=torch.int32).pin_memory(); hm = torch.zeros(n, dtype=torch.int32).pin_memory()
hi = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory(); hdist = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory()
def step():
    ex.extract_batch_raw(imgs, n, W, H, hk, hd, cap, hc, hm, flags=0)
    _, dd, dc, dn, dcap = ex.device_outputs()
    mt.knn2_batch_raw(dd, dc, cap, dd + P * cap * 32, dc + 4 * P, cap, P, hi, hdist, flags=_lib.IN_DEVICE)
for _ in range(3): step()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): step()
torch.cuda.synchronize()
print("%-28s e2e %.3f ms/step" % (os.environ.get("ORBFE_PIPE_SCHED", "default"), (time.perf_counter() - t0) / 10 * 1e3))
