"""A/B timing of the k-NN + PCA step on one GPU (not the bench contract): every variant runs in a fresh context (the
PCP_* knobs are read at pcp_create), records are compared bitwise with the first variant's.
usage: ab_time.py [--n N] [--k K[,K..]] [--impl a,b,..] [--cells s1,s2,..(scale of the automatic cell)] [--cloud urban|playground|frames]"""
import argparse, os, sys
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, synth

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1_000_000)
ap.add_argument("--k", default="20")
ap.add_argument("--impl", default="default,tile4")
ap.add_argument("--cells", default="1.0")
ap.add_argument("--cloud", default="urban")
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--seed", type=int, default=3)
args = ap.parse_args()
offs = None
if args.cloud == "urban":
    pts = synth.urban(args.seed, args.n, extent=200.0 * (args.n / 1e6) ** 0.5)
elif args.cloud == "playground":
    pts = synth.playground(1, args.n)
else:
    pts, offs = synth.vlp16_batch(128, seed0=5)
for k in [int(x) for x in args.k.split(",")]:
    ref = None
    for impl in args.impl.split(","):
        os.environ["PCP_KNN_IMPL"] = "" if impl == "default" else impl
        with capi.Pcp(0) as p:
            p.set_cloud(pts)
            if offs is not None:
                p.set_frames(offs)
            p.build_grid(0.0, k)
            cell0 = p.grid_info()[1]
            for sc in [float(s) for s in args.cells.split(",")]:
                ts, tg = [], []
                for _ in range(args.reps):
                    p.build_grid(cell0 * sc if sc != 1.0 else cell0, k)
                    out = p.knn_pca(k)
                    t = p.timings()
                    ts.append(t["search_pca"])
                    tg.append(t["bbox"] + t["keys"] + t["sort"] + t["table"])
                same = "ref" if ref is None else ("same bytes" if out.tobytes() == ref.tobytes() else "DIFFERENT RECORDS")
                if ref is None:
                    ref = out
                print(f"{args.cloud} n={pts.shape[0]} k={k} impl={impl:8s} cell={p.grid_info()[1]:.3f} (x{sc}) search min {min(ts[1:]):.3f} med {sorted(ts[1:])[len(ts[1:]) // 2]:.3f} ms, "
                      f"grid {min(tg[1:]):.3f} ms  [{same}]", flush=True)
