"""The whole VLP-16 chain on the device: packets -> points -> ring projection + ground removal + segmentation ->
segmented grid -> k-NN PCA (k = 10), 128 revolutions per launch.  Prints stage times; also the ncu target for
the kernels of the 'next' rows (profiles/)."""
import os, sys, time
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, os.path.join("oracle", "ref_decoder"))
from make_packets import make_packets
from pointcloudprocessing_b200 import capi
import torch

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
n_pk = 75 * 128
pk = make_packets(1, n_pk, az0=0, az_step=40, zero_frac=0.1)
# plausible geometry instead of random ranges: a 12 m x 9 m room seen from its middle
blocks = pk[:, :1200].reshape(-1, 100)
az = (blocks[:, 2].astype(np.int64) | (blocks[:, 3].astype(np.int64) << 8)) / 100.0 * np.pi / 180.0
vert = np.deg2rad(np.array([-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15] * 2, np.float64))
wall = np.minimum(6.0 / np.maximum(np.abs(np.sin(az)), 1e-3), 4.5 / np.maximum(np.abs(np.cos(az)), 1e-3))[:, None]
floor = np.where(vert < 0, 1.8 / np.maximum(-np.sin(vert), 1e-3), 1e9)[None, :]
dist = np.minimum(wall / np.cos(vert)[None, :], floor)
raw = np.clip(dist / 0.002, 1, 65535).astype(np.uint16)
ret = blocks[:, 4:].reshape(-1, 32, 3)
ret[:, :, 0] = raw & 0xFF
ret[:, :, 1] = raw >> 8
pk[:, :1200] = blocks.reshape(-1, 1200)
d_pk = torch.from_numpy(pk).cuda()
calib, prm = capi.ring_calib_default(), capi.ring_seg_params_default()
with capi.Pcp(0) as p:
    t = {}
    for it in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dec = p.decode_packets(None, want_arrays=False, packets_device=d_pk.data_ptr(), n_packets=n_pk)
        t1 = time.perf_counter()
        offs = dec["frame_offsets"]
        nf = offs.shape[0] - 1
        if it == 0:
            seg = p.ring_segment(calib, prm, frame_offsets=offs)  # host outputs once
        else:
            p.ring_segment(calib, prm, frame_offsets=offs)
        p.synchronize()
        t2 = time.perf_counter()
        p.build_grid(0.0, 10)
        out = torch.empty(dec["n_points"] * 68, dtype=torch.uint8, device="cuda")
        p.knn_pca(10, capi.CENTER_QUERY, out_device=out.data_ptr())
        p.synchronize()
        t3 = time.perf_counter()
        t = {"decode_ms": (t1 - t0) * 1e3, "grid_knn_pca_ms": (t3 - t2) * 1e3}
    print(f"packets {n_pk} -> {dec['n_points']} points in {nf} frames ({dec['n_complete']} complete); "
          f"decode {t['decode_ms']:.3f} ms, grid + k-NN PCA {t['grid_knn_pca_ms']:.3f} ms; "
          f"segments in frame 1: {int(seg['n_labels'][1]) if seg else '-'}")
