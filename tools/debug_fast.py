import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "py"))
import numpy as np
import oracle_lib
from orbslam3lib_b200 import ORBextractor, scenes
img = scenes.make("texture", 640, 480, 1000)
ref = oracle_lib.OracleExtractor(1000, 1.2, 8, 20, 7)
ref(img)
ex = ORBextractor(1000, 1.2, 8, 20, 7)
ex(img)
for l in range(2):
    cr, cg = ref.level_candidates(l), ex.level_candidates(l)
    sr = set(map(tuple, cr.tolist())); sg = set(map(tuple, cg.tolist()))
    print("level", l, len(cr), len(cg), "missing", len(sr - sg), "extra", len(sg - sr))
    miss = sorted(sr - sg, key=lambda t: (t[1], t[0]))
    print("missing sample", miss[:20])
    print("extra sample", sorted(sg - sr)[:20])
    xs = np.array([m[0] for m in miss]); ys = np.array([m[1] for m in miss])
    if len(xs):
        print("missing x//36 hist", np.bincount(xs // 36, minlength=18))
        print("missing y//35 hist", np.bincount(ys // 35, minlength=14))
        rs = np.array([m[2] for m in miss]); print("missing resp", np.bincount(rs // 10))
print("---- shift search level 0")
cr, cg = ref.level_candidates(0), ex.level_candidates(0)
sr = set((a, b) for a, b, c in cr.tolist())
best = []
for dx in range(-45, 46):
    for dy in range(-45, 46):
        ov = sum(((a + dx, b + dy) in sr) for a, b, c in cg.tolist())
        best.append((ov, dx, dy))
best.sort(reverse=True)
print(best[:5])
# per-cell oracle counts
pix = ref.level_pixels(0)
h, w = pix.shape
nC, nR, wC, hC = 17, 12, 36, 38
for i in range(2):
    for j in range(3):
        x0, y0 = 16 + j * wC, 16 + i * hC
        x1, y1 = min(x0 + wC + 6, w - 16), min(y0 + hC + 6, h - 16)
        sub = np.ascontiguousarray(pix[y0:y1, x0:x1])
        o = oracle_lib.fast9_nms(sub, 20)
        if len(o) == 0: o = oracle_lib.fast9_nms(sub, 7)
        mine = [t for t in cg.tolist() if j * wC + 3 <= t[0] < j * wC + 3 + (x1 - x0 - 6) and i * hC + 3 <= t[1] < i * hC + 3 + (y1 - y0 - 6)]
        print("cell", i, j, "oracle", [(a + j * wC, b + i * hC, c) for a, b, c in o.tolist()], "gpu-in-range", mine)
