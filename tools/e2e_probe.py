"""Where the end-to-end step spends its time: PCIe copies alone, pipelined extraction alone, matcher alone."""
import sys, os, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from orbslam3lib_b200 import BFMatcher, ORBextractor, _lib
P = 512; n = 2 * P; W, H = bench.W, bench.H
dev = torch.device("cuda", 0)
imgs = torch.from_numpy(bench.make_pairs(P, 1000)).pin_memory()
ex = ORBextractor(1200, 1.2, 8, 20, 7); mt = BFMatcher()
cap = ex.max_keypoints
h_kps = torch.zeros((n, cap, 28), dtype=torch.uint8).pin_memory(); h_desc = torch.zeros((n, cap, 32), dtype=torch.uint8).pin_memory()
h_counts = torch.zeros(n, dtype=torch.int32).pin_memory(); h_monos = torch.zeros(n, dtype=torch.int32).pin_memory()
h_idx = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory(); h_dist = torch.zeros((P, cap, 2), dtype=torch.int32).pin_memory()
d = torch.empty_like(imgs, device=dev); dk = torch.zeros((n, cap, 60), dtype=torch.uint8, device=dev)
def t(fn, k=5):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(k): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / k * 1e3
print("h2d 370MB ms", t(lambda: d.copy_(imgs, non_blocking=True)))
hk = torch.zeros((n, cap, 60), dtype=torch.uint8).pin_memory()
print("d2h 78MB ms", t(lambda: hk.copy_(dk, non_blocking=True)))
def both():
    d.copy_(imgs, non_blocking=True)
    with torch.cuda.stream(s2): hk.copy_(dk, non_blocking=True)
s2 = torch.cuda.Stream()
print("h2d + d2h concurrently ms", t(both))
def ext(): ex.extract_batch_raw(imgs, n, W, H, h_kps, h_desc, cap, h_counts, h_monos, flags=0)
print("extract host->host ms", t(ext))
def knn():
    _, dd, dc, dn, dcap = ex.device_outputs()
    mt.knn2_batch_raw(dd, dc, cap, dd + P * cap * 32, dc + 4 * P, cap, P, h_idx, h_dist, flags=_lib.IN_DEVICE)
print("knn dev->host ms", t(knn), "kernel", mt.last_ms())
for sub in (64, 128):
    pass
