"""Debug build only (make NVCC_EXTRA=-DPCP_STATS): tile-kernel counters on a batch of VLP-16 frames."""
import ctypes, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
nb, k = 64, int(sys.argv[1]) if len(sys.argv) > 1 else 10
fr = [synth.vlp16_frame(1000 + f, pose_jitter=0.5) for f in range(16)]
mine = [fr[f % 16] for f in range(nb)]
cloud = np.concatenate(mine)
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in mine])]).astype(np.int64)
p = capi.Pcp(0)
lib = capi.load()
st = (ctypes.c_ulonglong * 16)()
p.set_cloud(cloud)
for cell in (0.0, 0.3, 0.5, 1.5, 2.5):
    p.set_frames(offs)
    p.build_grid(cell, k)
    lib.pcp_debug_tile6_stats(st)
    p.knn_pca(k)
    lib.pcp_debug_tile6_stats(st)
    q = st[0]
    t = p.timings()
    print(f"cell {p.grid_info()[1]:.3f} search {t['search_pca']:.2f} ms queries {q} deferred {st[1] / q:.4f} staged/tile {st[2] * 32 / q:.1f} sorted_in/ok {st[3] / max(q - st[1], 1):.2f} | other-row {st[4] / q:.4f} uncovered {st[5] / q:.4f} bin-full {st[6] / q:.4f} refined {st[9] / q:.4f} unproven {st[10] / q:.5f} kept-in-pass2/tile {st[8] * 32 / q:.1f}", flush=True)
