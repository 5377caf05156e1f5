"""FAST stage time with and without a real second pass (minThFAST = iniThFAST finds nothing new in the empty cells)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from orbslam3lib_b200 import ORBextractor, _lib
P = 256; n = 2 * P; W, H = bench.W, bench.H
dev = torch.device("cuda", 0)
d_imgs = torch.from_numpy(bench.make_pairs(P, 1000)).to(dev)
for mn in (7, 20):
    ex = ORBextractor(1200, 1.2, 8, 20, mn)
    ex.set_stream(torch.cuda.current_stream().cuda_stream)
    cap = ex.max_keypoints
    kps = torch.zeros((n, cap, 28), dtype=torch.uint8, device=dev); desc = torch.zeros((n, cap, 32), dtype=torch.uint8, device=dev)
    counts = torch.zeros(n, dtype=torch.int32, device=dev); monos = torch.zeros(n, dtype=torch.int32, device=dev)
    acc = {}
    for it in range(6):
        ex.extract_batch_raw(d_imgs, n, W, H, kps, desc, cap, counts, monos, flags=_lib.IN_DEVICE | _lib.OUT_DEVICE)
        torch.cuda.synchronize()
        if it >= 2:
            for k, v in ex.stage_ms().items(): acc[k] = acc.get(k, 0) + v / 4
    print("minTh", mn, {k: round(v, 4) for k, v in acc.items()}, "keypoints/image", float(counts.float().mean()))
