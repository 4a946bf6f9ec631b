"""Ad-hoc stage timing on the GPU (not the bench contract)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
scales = [float(s) for s in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1.0]
pts = synth.urban(3, n, extent=200.0 * (n / 1e6) ** 0.5)
p = capi.Pcp(0)
p.set_cloud(pts)
for sc in scales:
    p.build_grid(0.0, k)
    dims, cell, nc = p.grid_info()
    if sc != 1.0:
        p.build_grid(cell * sc, k)
        dims, cell, nc = p.grid_info()
    for it in range(3):
        p.build_grid(cell, k)
        out = p.knn_pca(k)
    t = p.timings()
    print(f"n={n} k={k} cell={cell:.3f} dims={dims} cells={nc} " + " ".join(f"{a}={b:.3f}" for a, b in t.items()), flush=True)
r = 0.5
p.build_grid(r * 1.001, 0)
try:
    for it in range(2):
        out = p.radius_pca(r)
    t = p.timings()
    print(f"radius r={r} mean ptNum={out['ptNum'].mean():.1f} max={out['ptNum'].max()} " + " ".join(f"{a}={b:.3f}" for a, b in t.items()))
except capi.PcpError as e:
    print("radius:", e)
