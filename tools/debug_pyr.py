import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle", "py"))
import oracle_lib
from orbslam3lib_b200 import ORBextractor, scenes
kind, w, h, seed, nf, nl = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
img = scenes.make(kind, w, h, seed)
ref = oracle_lib.OracleExtractor(nf, 1.2, nl, 20, 7); ref(img)
ex = ORBextractor(nf, 1.2, nl, 20, 7); ex(img)
for l in range(nl):
    a, b = ex.level_padded(l), ref.level_padded(l)
    bad = np.argwhere(a != b)
    print("level", l, a.shape, "mismatches", len(bad))
    if len(bad):
        print(" rows", np.unique(bad[:, 0])[:20], "cols", np.unique(bad[:, 1])[:40])
        for (y, x) in bad[:8]:
            print("  ", y, x, a[y, x], b[y, x])
