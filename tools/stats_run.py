"""Debug build only (make NVCC_EXTRA=-DPCP_STATS): per-query event counts of the search kernel."""
import ctypes, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
pts = synth.urban(3, n, extent=200.0 * (n / 1e6) ** 0.5)
p = capi.Pcp(0)
lib = capi.load()
st = (ctypes.c_ulonglong * 8)()
p.set_cloud(pts)
for cell in [float(c) for c in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["0"])]:
    p.build_grid(cell, k)
    lib.pcp_debug_stats(st)
    p.knn_pca(k)
    lib.pcp_debug_stats(st)
    q = st[0]
    names = ["queries", "passes", "runs", "chunks", "cands", "passed", "merges_in", "flushes"]
    print(p.grid_info()[1], {nm: round(st[i] / q, 2) for i, nm in enumerate(names)}, "ms", p.timings()["search_pca"])
