"""Debug build only (make NVCC_EXTRA=-DPCP_STATS): counters of the k-NN tile kernel."""
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
for cell in [float(c) for c in (sys.argv[3].split(",") if len(sys.argv) > 3 else ["0"])]:
    p.build_grid(cell, k)
    lib.pcp_debug_tile6_stats(st)
    p.knn_pca(k)
    lib.pcp_debug_tile6_stats(st)
    q = st[0]
    print(f"cell {p.grid_info()[1]:.3f} queries {q} deferred {st[1] / q:.4f} staged/tile {st[2] * 32 / q:.1f} sorted_in/ok {st[3] / max(q - st[1], 1):.2f} | other-row {st[4] / q:.4f} uncovered {st[5] / q:.4f} bin-full {st[6] / q:.4f} refined {st[9] / q:.4f} unproven {st[10] / q:.5f} kept-in-pass2/tile {st[8] * 32 / q:.1f}", flush=True)
