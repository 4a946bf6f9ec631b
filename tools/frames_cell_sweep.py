"""Search-stage time of one batch of VLP-16 frames (segmented grid) against the cell size."""
import sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 128
k = int(sys.argv[2]) if len(sys.argv) > 2 else 10
fr = [synth.vlp16_frame(1000 + f, pose_jitter=0.5) for f in range(16)]
mine = [fr[f % 16] for f in range(nb)]
cloud = np.concatenate(mine)
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in mine])]).astype(np.int64)
with capi.Pcp(0) as p:
    p.set_cloud(cloud)
    for cell in (0.0, 0.2, 0.3, 0.4, 0.5, 0.7, 1.0, 1.4, 2.0):
        for it in range(3):
            p.set_frames(offs)
            p.build_grid(cell, k)
            p.knn_pca(k)
            t = p.timings()
        dims, c, nc = p.grid_info()
        grid = t["bbox"] + t["keys"] + t["sort"] + t["table"]
        print(f"cell={c:.3f} dims={list(dims)} cells={nc} grid={grid:.3f} ms search={t['search_pca']:.3f} ms "
              f"-> {cloud.shape[0] / (grid + t['search_pca']) / 1e3:.1f} M points/s", flush=True)
