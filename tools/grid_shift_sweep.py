"""Diagnostic: search time against the grid origin's offset (PCP_GRID_SHIFT, fraction of a cell) at fixed cell sizes."""
import os, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
k = int(sys.argv[1]) if len(sys.argv) > 1 else 15
cells = [float(c) for c in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1.094, 1.204, 1.313]
pts = synth.urban(3, 1_000_000, extent=200.0)
ref = None
axes = [int(a) for a in sys.argv[3].split(",")] if len(sys.argv) > 3 else [-1]
for cell in [c for c in cells for _ in axes]:
    axis = axes[0]
    axes = axes[1:] + axes[:1]
    os.environ["PCP_GRID_SHIFT_AXIS"] = str(axis)
    row = [f"axis {axis}"]
    for sh in (0.0, 0.2, 0.4, 0.6, 0.8):
        os.environ["PCP_GRID_SHIFT"] = str(sh)
        ts = []
        with capi.Pcp(0) as p:  # the knobs are read when a context is created
            p.set_cloud(pts)
            for _ in range(4):
                p.build_grid(cell, k)
                out = p.knn_pca(k)
                ts.append(p.timings()["search_pca"])
        if ref is None:
            ref = out
        assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
        row.append(f"{sh:.1f}:{min(ts[1:]):.3f}")
    print(f"k={k} cell={cell}: " + "  ".join(row), flush=True)
