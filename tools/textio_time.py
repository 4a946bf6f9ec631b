"""Device time of the text parser (pcp_set_cloud_txt) on a cloud file of the playground shape, next to the CPU
restatement (the reference's own stream statements, minus its per-line std::cout print)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
from oracle import oracle_py as O

n = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
pts = synth.urban(1, n, extent=60.0).astype(np.float64)
text = "".join("%.3f %.3f %.3f %.0f\n" % tuple(p) for p in pts).encode()
with capi.Pcp(0) as p:
    p.set_cloud_txt(text)
    t0 = time.perf_counter()
    for _ in range(5):
        got = p.set_cloud_txt(text)
    dt = (time.perf_counter() - t0) / 5
t0 = time.perf_counter()
ref = O.txt2pc(text)
dc = time.perf_counter() - t0
assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
print(f"text cloud {n} lines, {len(text) / 1e6:.1f} MB: device parse incl. upload of the text and download of the points {dt * 1e3:.2f} ms "
      f"({n / dt / 1e6:.1f} M points/s, {len(text) / dt / 1e9:.2f} GB/s); CPU stream parse, 1 thread {dc * 1e3:.0f} ms ({n / dc / 1e6:.2f} M points/s)")
