"""Device time of the lidarOdometry row (capped radius count r = 3 m, range/point-id images, smoothness) for a
batch of VLP-16 frames, next to the CPU restatement of the same calls on one frame."""
import ctypes as C
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from pointcloudprocessing_b200 import capi, synth
from oracle import oracle_py as O

nf = int(sys.argv[1]) if len(sys.argv) > 1 else 128
distinct = [synth.vlp16_frame(300 + f, pose_jitter=0.5) for f in range(16)]
frames = [distinct[f % 16] for f in range(nf)]
offs = np.zeros(nf + 1, np.int64)
np.cumsum([f.shape[0] for f in frames], out=offs[1:])
pts = np.ascontiguousarray(np.concatenate(frames))
prm = capi.odom_params_default()
cells = nf * 16 * 900
with capi.Pcp(0) as p:
    lib = p.lib
    stream = torch.cuda.ExternalStream(p.stream)
    p.set_cloud(pts)
    p.set_frames(offs)
    d_rng = torch.empty(cells, dtype=torch.float32, device="cuda")
    d_sm = torch.empty(cells, dtype=torch.float32, device="cuda")
    d_ty = torch.empty(cells, dtype=torch.float32, device="cuda")
    d_cnt = torch.empty(pts.shape[0], dtype=torch.int32, device="cuda")
    for cell in (3.0, 1.5, 0.75):
        p.build_grid(cell, 0)
        res = {}
        for name, fn in (
            ("count cap 6", lambda: lib.pcp_radius_count(p.h, 3.0, 6, d_cnt.data_ptr(), 1)),
            ("range+ptid images (incl. count)", lambda: lib.pcp_odom_range_image(p.h, C.byref(prm), offs.ctypes.data, nf, d_rng.data_ptr(), None, None, None, None, 1)),
            ("smoothness", lambda: lib.pcp_odom_smoothness(p.h, C.byref(prm), d_rng.data_ptr(), 1, nf, d_sm.data_ptr(), d_ty.data_ptr(), 1)),
        ):
            for _ in range(3):
                assert fn() == 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                for _ in range(10):
                    fn()
                e1.record(stream)
            e1.synchronize()
            res[name] = e0.elapsed_time(e1) / 10
        print(f"cell {cell}: " + ", ".join(f"{k} {v:.3f} ms" for k, v in res.items()) + f"  per {nf} frames ({pts.shape[0]} points)", flush=True)
fr = frames[0]
t0 = time.perf_counter()
cnt = O.KdTree(fr).radius_counts(3.0, nthreads=1)
t1 = time.perf_counter()
rng = O.odom_range_image(fr, cnt)[0]
t2 = time.perf_counter()
O.odom_smoothness(rng)
t3 = time.perf_counter()
print(f"CPU restatement, one frame, 1 thread: kd-tree + radiusSearch(3.0) counts {1e3*(t1-t0):.1f} ms, images {1e3*(t2-t1):.2f} ms, smoothness {1e3*(t3-t2):.2f} ms")
