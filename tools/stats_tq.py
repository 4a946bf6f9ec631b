"""Debug build only (make NVCC_EXTRA=-DPCP_STATS): per-query event counts of the thread-per-query k-NN kernel."""
import ctypes, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
pts = synth.urban(3, n, extent=200.0 * (n / 1e6) ** 0.5)
p = capi.Pcp(0)
lib = capi.load()
st = (ctypes.c_ulonglong * 16)()
p.set_cloud(pts)
names = ["queries", "warp_iters", "lane_iters", "warp_hist", "lane_hist", "hist_cands", "range_redo", "narrow", "coll_cands", "sorted_in", "slow_lanes"]
for cell in [float(c) for c in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["0"])]:
    p.build_grid(cell, k)
    lib.pcp_debug_tq_stats(st)
    p.knn_pca(k)
    lib.pcp_debug_tq_stats(st)
    q = st[0]
    d = {nm: round(st[i] / q, 3) for i, nm in enumerate(names)}
    d["warp_iters"] = round(st[1] * 32 / q, 3); d["warp_hist"] = round(st[3] * 32 / q, 3)
    print(p.grid_info()[1], d, "ms(stats build)", p.timings()["search_pca"], flush=True)
