"""configs[3] at full size on ONE GPU: synthetic urban cloud of N points at 100 pts/m^2, radiusSearch r = 0.5 m + PCA.

Size-independent parity property: a spatial crop of the cloud, handed to a second context as its own
shard (own points + halo of width r, global indices for the tie-break), must reproduce the full run's
records for its own points BIT FOR BIT, and must match the CPU oracle run on the same crop.  Also times
the k-NN path on the full cloud.   usage: python tools/fullsize_check.py [N] [crops] > profiles/...json"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from pointcloudprocessing_b200 import capi, synth
from oracle import oracle_py as O
from parity import check_features

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
n_crops = int(sys.argv[2]) if len(sys.argv) > 2 else 6
r, win = 0.5, 30.0
ext = (n / 100.0) ** 0.5
t0 = time.time()
pts = synth.urban(4, n, extent=ext, order="asis")
gen_s = time.time() - t0
res = {"n": n, "extent_m": ext, "radius": r, "generate_s": round(gen_s, 1)}
p = capi.Pcp(0)
d_pts = torch.from_numpy(pts).cuda()
d_out = torch.empty(n * 68, dtype=torch.uint8, device="cuda")
p.set_cloud_device(d_pts.data_ptr(), n)
for it in range(2):
    p.build_grid(r * 1.001, 0)
    p.radius_pca(r, 13, capi.CENTER_QUERY, False, out_device=d_out.data_ptr())
t = p.timings()
build_ms = t["bbox"] + t["keys"] + t["sort"] + t["table"]
res["radius_pca"] = {"grid_ms": round(build_ms, 2), "search_pca_ms": round(t["search_pca"], 2), "points_per_s": n / ((build_ms + t["search_pca"]) * 1e-3),
                     "grid": p.grid_info()[0], "cells": p.grid_info()[2]}
full = np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=capi.FEATURE_DTYPE)
res["radius_pca"]["mean_ptNum"] = float(full["ptNum"].mean()); res["radius_pca"]["max_ptNum"] = int(full["ptNum"].max())
res["radius_pca"]["valid_fraction"] = float((full["ptNum"] >= 13).mean())
assert np.array_equal(full["ptID"], np.arange(n, dtype=np.int32)), "ptID != index"
rng = np.random.default_rng(0)
crops = []
q = capi.Pcp(0)
for c in range(n_crops):
    x0, y0 = rng.uniform(0, ext - win, 2)
    if c == 0: x0, y0 = 0.0, 0.0  # a corner of the cloud
    inx = (pts[:, 0] >= x0 - r - 1e-3) & (pts[:, 0] < x0 + win + r + 1e-3) & (pts[:, 1] >= y0 - r - 1e-3) & (pts[:, 1] < y0 + win + r + 1e-3)
    cand = np.nonzero(inx)[0]
    own_m = (pts[cand, 0] >= x0) & (pts[cand, 0] < x0 + win) & (pts[cand, 1] >= y0) & (pts[cand, 1] < y0 + win)
    gid = np.concatenate([cand[own_m], cand[~own_m]]).astype(np.int32)
    n_own = int(own_m.sum())
    sh = torch.from_numpy(np.ascontiguousarray(pts[gid])).cuda(); sg = torch.from_numpy(gid).cuda()
    so = torch.empty(n_own * 68, dtype=torch.uint8, device="cuda")
    q.set_cloud_device(sh.data_ptr(), gid.size, sg.data_ptr(), n_own, keep=(sh, sg))
    q.build_grid(r * 1.001, 0)
    q.radius_pca(r, 13, capi.CENTER_QUERY, False, out_device=so.data_ptr())
    q.synchronize()
    got = np.frombuffer(so.cpu().numpy().tobytes(), dtype=capi.FEATURE_DTYPE)
    ref = full[gid[:n_own]]
    same = got.tobytes() == ref.tobytes()
    # CPU oracle on the crop (own + halo as one cloud), compared on the own points
    sub = np.ascontiguousarray(pts[gid])
    oref, ovalid = O.radius_pca_cloud(sub, r, 13, False)
    o_own = oref[:n_own].copy(); o_own["ptID"] = gid[:n_own]
    stats = check_features(got, o_own, ovalid[:n_own], f"crop {c}")
    crops.append({"x0": float(x0), "y0": float(y0), "own": n_own, "halo": int(gid.size - n_own), "bitwise_equal_to_full_run": bool(same),
                  "oracle_checked_records": int(ovalid[:n_own].sum())})
    assert same, f"crop {c}: shard records differ from the full run"
res["crops"] = crops
# k-NN + PCA (k = 20) on the full cloud, timing only (parity of the k-NN path is covered at small sizes)
for it in range(2):
    p.build_grid(0.0, 20)
    p.knn_pca(20, capi.CENTER_QUERY, out_device=d_out.data_ptr())
t = p.timings()
build_ms = t["bbox"] + t["keys"] + t["sort"] + t["table"]
res["knn20_pca"] = {"grid_ms": round(build_ms, 2), "search_pca_ms": round(t["search_pca"], 2), "points_per_s": n / ((build_ms + t["search_pca"]) * 1e-3), "cell": p.grid_info()[1]}
print(json.dumps(res, indent=1))
