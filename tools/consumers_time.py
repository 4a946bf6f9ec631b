"""Device time of DBSCAN (planar, as clusterDBSCAN runs it) and of one mean-shift sweep on an urban cloud, next to
the CPU restatement on a sample.  Usage: consumers_time.py [n_points] [--no-cpu]"""
import ctypes as C
import sys, time
import numpy as np
sys.path.insert(0, ".")
import torch
from pointcloudprocessing_b200 import capi, synth
from oracle import oracle_py as O

n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 1000000
cpu = "--no-cpu" not in sys.argv
pts = synth.urban(3, n, extent=200.0 * (n / 1e6) ** 0.5)
flat = pts.copy()
flat[:, 2] = 0.0


def timed(stream, fn, reps=5):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    e1.synchronize()
    return e0.elapsed_time(e1) / reps


with capi.Pcp(0) as p:
    stream = torch.cuda.ExternalStream(p.stream)
    lab = torch.empty(n, dtype=torch.int32, device="cuda")
    nc = C.c_int32()
    out = torch.empty((n, 4), dtype=torch.float32, device="cuda")
    p.set_cloud(flat)
    for r, mp in ((0.3, 8), (0.5, 10)):
        p.build_grid(r, 0)
        ms = timed(stream, lambda: p._ck(p.lib.pcp_dbscan(p.h, r, mp, lab.data_ptr(), C.addressof(nc), 1)))
        print(f"DBSCAN r={r} minpts={mp}: {ms:.3f} ms per {n} points ({n / ms / 1e3:.1f} M points/s), {nc.value} clusters", flush=True)
    p.set_cloud(pts)
    for r, cell in ((2.0, 2.0), (2.0, 1.0), (0.5, 0.5)):
        p.build_grid(cell, 0)
        ms = timed(stream, lambda: p._ck(p.lib.pcp_mean_shift(p.h, r, 15, 1, None, 0, out.data_ptr(), None, 1)))
        cnt = p.radius_count(r, 0)
        print(f"mean shift r={r} cell={cell}: {ms:.3f} ms per sweep of {n} points ({n / ms / 1e3:.1f} M points/s), mean {cnt.mean():.0f} neighbours",
              flush=True)
if cpu:
    m = min(n, 100000)
    order = np.argsort(pts[:, 0], kind="stable")[:m]  # a strip of the cloud at its full density
    flat, pts = flat[order], pts[order]
    sub = np.ascontiguousarray(flat[:m])
    t0 = time.perf_counter()
    O.dbscan(sub, 0.3, 8)
    t1 = time.perf_counter()
    sub3 = np.ascontiguousarray(pts[:m])
    O.mean_shift_step(sub3, sub3, 2.0, 15)
    t2 = time.perf_counter()
    print(f"CPU restatement on a {m}-point strip of the same cloud: DBSCAN (1 thread, as the reference) {t1 - t0:.2f} s = {m / (t1 - t0) / 1e6:.3f} M points/s; "
          f"mean-shift sweep ({O.num_threads()} threads) {t2 - t1:.2f} s = {m / (t2 - t1) / 1e6:.3f} M points/s")
