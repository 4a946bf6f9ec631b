"""Work-list kernel: sweep of the ticket tile size and the number of CTAs (search stage time, ms per cloud)."""
import os, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
pts = synth.urban(3, n, extent=200.0 * (n / 1e6) ** 0.5)
p = capi.Pcp(0)
p.set_cloud(pts)
p.build_grid(0.0, k)
ref = None
for blocks in (444, 592, 296):
    for tile in (32, 16, 8, 4):
        os.environ["PCP_WL_BLOCKS"], os.environ["PCP_WL_TILE"] = str(blocks), str(tile)
        ts = []
        for it in range(6):
            p.build_grid(0.0, k)
            out = p.knn_pca(k)
            ts.append(p.timings()["search_pca"])
        if ref is None:
            ref = out
        same = bool(np.array_equal(out.view(np.uint8), ref.view(np.uint8)))
        print(f"blocks={blocks} tile={tile}: search {min(ts[1:]):.4f} ms (median {sorted(ts[1:])[2]:.4f})  same_bytes={same}", flush=True)
