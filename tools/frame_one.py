"""One VLP-16 frame through the C-ABI a few times (for ncu launch lists)."""
import sys
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
fr = synth.vlp16_frame(300, pose_jitter=0.5)
with capi.Pcp(0) as p:
    for it in range(4):
        p.set_cloud(fr)
        p.ring_project()
        p.build_grid(float(sys.argv[1]) if len(sys.argv) > 1 else 0.0, 10)
        p.knn_pca(10)
    print(fr.shape, p.grid_info(), p.timings())
