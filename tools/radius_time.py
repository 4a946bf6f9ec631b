"""Ad-hoc timing of the radius + PCA path at configs[3] density (100 pts/m^2, r = 0.5 m)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
r = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
ext = (n / 100.0) ** 0.5
pts = synth.urban(4, n, extent=ext)
p = capi.Pcp(0)
p.set_cloud(pts)
for cellf in (1.001, 1.3):
    p.build_grid(r * cellf, 0)
    try:
        for it in range(3):
            p.build_grid(r * cellf, 0)
            out = p.radius_pca(r)
        t = p.timings()
        print(f"n={n} extent={ext:.0f} r={r} cell={r*cellf:.3f} mean ptNum={out['ptNum'].mean():.1f} max={out['ptNum'].max()} valid={np.mean(out['ptNum']>=13):.3f} "
              + " ".join(f"{a}={b:.3f}" for a, b in t.items()), flush=True)
    except capi.PcpError as e:
        print("radius:", e, flush=True)
