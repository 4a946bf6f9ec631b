"""Small pass over every entry point, sized for compute-sanitizer --tool memcheck."""
import sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth

pts = synth.urban(1, 6000, extent=20.0)
fr = synth.vlp16_frame(2)[::4].copy()
fb, offs = synth.vlp16_batch(2, seed0=5)
fb, offs = np.ascontiguousarray(fb[: offs[2]]), offs[:3]
with capi.Pcp(0) as p:
    p.set_cloud(pts)
    for k in (10, 20, 32, 50):
        p.build_grid(0.0, k)
        p.knn_pca(k, 0, want_normals=True)
        p.knn_neighbors(k)
    p.build_grid(0.5, 0)
    p.radius_pca(0.5)
    p.radius_neighbors(0.4)
    p.radius_count(0.5, 6)
    flat = pts.copy()
    flat[:, 2] = 0
    p.set_cloud(flat)
    p.build_grid(0.3, 0)
    lab, nc = p.dbscan(0.3, 6)
    p.set_cloud(pts)
    p.cluster_boxes(lab, nc)
    p.build_grid(1.0, 0)
    p.mean_shift(1.0, 10, 2)
    p.set_cloud(fr)
    p.ring_project()
    p.ring_segment()
    p.build_grid(3.0, 0)
    p.odom_range_image()
    p.odom_smoothness(n_frames=1)
    p.set_cloud(fb)
    p.set_frames(offs)
    p.build_grid(0.0, 10)
    p.knn_pca(10)
    p.odom_range_image(frame_offsets=offs)
    p.odom_smoothness(n_frames=2)
    got = p.set_cloud_txt(b"1 2 3 4\n5 6 7\n\n8.5 abc 9 1\n1e-40 3.5e38 -2.5e-3 7\n3")
    gold = np.load("tests/golden/decoder_sweep.npz")
    p.decode_packets(gold["packets"])
print("memcheck pass done", got.shape)
