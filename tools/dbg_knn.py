"""Debug: list the queries whose k-NN list differs from the golden fixture (optionally with the exact path forced)."""
import os, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi
G = np.load("tests/golden/golden_small.npz")
pts = G["urban"]
with capi.Pcp(0) as p:
    p.set_cloud(pts); p.build_grid(0.0, 20)
    print("grid", p.grid_info())
    idx, d2 = p.knn_neighbors(20)
ref, rd2 = G["urban_knn20_idx"], G["urban_knn20_d2"]
bad = np.nonzero((idx != ref).any(1))[0]
print("bad queries", bad.size, bad[:20])
for q in bad[:6]:
    print("q", q, pts[q])
    print(" got", idx[q]); print(" ref", ref[q]); print(" gd2", d2[q]); print(" rd2", rd2[q])
