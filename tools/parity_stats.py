"""UNGATED error distributions of the PCA records, GPU vs the CPU oracle and both vs the fp64 eigen-solve of the same
fp32 moment, for the stated configs.  Writes one JSON object; commit it under profiles/.
usage: parity_stats.py [out.json] [--small]"""
import json, sys
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from oracle import oracle_py as O
from pointcloudprocessing_b200 import capi, synth
import parity

small = "--small" in sys.argv
out_path = next((a for a in sys.argv[1:] if not a.startswith("--")), "gpurun_out/parity_stats.json")


def q(x):
    x = np.asarray(x, np.float64).ravel()
    x = x[np.isfinite(x)]
    if x.size == 0:
        return None
    return {"n": int(x.size), "median": float(np.median(x)), "p99": float(np.quantile(x, 0.99)), "p999": float(np.quantile(x, 0.999)), "max": float(x.max())}


def moment_truth(pts, idx_rows, centre):
    """fp64 eigenvalues of the fp32 moment in PCL accumulation order (bitwise the matrix both sides solve)."""
    k = idx_rows.shape[1]
    P = pts[idx_rows, :3].astype(np.float32) - centre[:, None, :].astype(np.float32)
    acc = [np.zeros(P.shape[0], np.float32) for _ in range(6)]  # 00 01 02 11 12 22
    for j in range(k):
        px, py, pz = P[:, j, 0], P[:, j, 1], P[:, j, 2]
        acc[3] += py * py; acc[4] += py * pz; acc[5] += pz * pz
        acc[0] += px * px; acc[1] += py * px; acc[2] += pz * px
    n32 = np.float32(k)
    C = np.zeros((P.shape[0], 3, 3), np.float64)
    C[:, 0, 0] = acc[0] / n32; C[:, 0, 1] = C[:, 1, 0] = acc[1] / n32; C[:, 0, 2] = C[:, 2, 0] = acc[2] / n32
    C[:, 1, 1] = acc[3] / n32; C[:, 1, 2] = C[:, 2, 1] = acc[4] / n32; C[:, 2, 2] = acc[5] / n32
    w, v = np.linalg.eigh(C)
    return w, v


def stats(name, got, ref, truth=None, valid=None):
    m = np.ones(got.shape[0], bool) if valid is None else valid.astype(bool)
    lg, lr = got["la"][m].astype(np.float64), ref["la"][m].astype(np.float64)
    lmax = np.abs(lr).max(1, keepdims=True) + 1e-300
    d = np.abs(lg - lr)
    r = {"records": int(m.sum()),
         "ptID_ptNum_equal": bool(np.array_equal(got["ptID"], ref["ptID"]) and np.array_equal(got["ptNum"], ref["ptNum"])),
         "dlambda_over_lmax": q(d / lmax),
         "dlambda_over_self": q(d / np.maximum(np.abs(lr), 1e-300)),
         "frac_lambda_outside_1e-4_of_self": float((d > 1e-4 * np.abs(lr)).mean()),
         "frac_lambda_outside_survey_bar_floor_1e-3": float((d > 1e-4 * np.maximum(np.abs(lr), 1e-3 * lmax)).mean()),
         "frac_lambda_outside_test_bar_floor_1e-2": float((d > 1e-4 * np.maximum(np.abs(lr), 1e-2 * lmax)).mean())}
    gaps = np.stack([lr[:, 1] - lr[:, 0], np.minimum(lr[:, 1] - lr[:, 0], lr[:, 2] - lr[:, 1]), lr[:, 2] - lr[:, 1]], 1)
    for c, nm in enumerate(("v1", "v2", "v3")):
        e = 1.0 - np.abs((got[nm][m].astype(np.float64) * ref[nm][m].astype(np.float64)).sum(1))
        ok = gaps[:, c] >= parity.GAP_MIN * lmax[:, 0]
        r[nm + "_1_minus_absdot_ungated"] = q(e)
        r[nm + "_1_minus_absdot_gap_ge_1e-3"] = q(e[ok])
        r[nm + "_frac_gated_out"] = float(1.0 - ok.mean())
    for nm in ("lineStruc", "planeStruc", "scatterStruc"):
        fg, fr = got[nm][m].astype(np.float64), ref[nm][m].astype(np.float64)
        fin = np.isfinite(fg) & np.isfinite(fr)
        e = np.abs(fg - fr)[fin] / np.maximum(1.0, np.abs(fr[fin]))
        r[nm + "_relerr_ungated"] = q(e)
        r[nm + "_frac_outside_1e-4"] = float((e > 1e-4).mean())
        r[nm + "_frac_outside_3e-4"] = float((e > 3e-4).mean())
        r[nm + "_nonfinite_mismatch"] = int((np.isfinite(fg) != np.isfinite(fr)).sum())
    if truth is not None:
        w = truth[m]
        wl = np.abs(w).max(1, keepdims=True) + 1e-300
        r["gpu_vs_fp64_truth_over_lmax"] = q(np.abs(lg - w) / wl)
        r["oracle_vs_fp64_truth_over_lmax"] = q(np.abs(lr - w) / wl)
    print(name, json.dumps(r)[:400], "...", flush=True)
    return r


res = {"bars": {"LAMBDA_RTOL": parity.LAMBDA_RTOL, "LAMBDA_FLOOR": parity.LAMBDA_FLOOR, "NORMAL_TOL": parity.NORMAL_TOL,
                "GAP_MIN": parity.GAP_MIN, "FEATURE_RTOL": parity.FEATURE_RTOL}}
nthr = O.num_threads()
with capi.Pcp(0) as p:
    cases = [("configs[2] urban 1M k=20 (bench cloud)", synth.urban(3, 50000 if small else 1_000_000, extent=200.0 * ((50000 if small else 1e6) / 1e6) ** 0.5), 20, 0),
             ("configs[2] urban 1M k=30", synth.urban(3, 50000 if small else 1_000_000, extent=200.0 * ((50000 if small else 1e6) / 1e6) ** 0.5), 30, 0),
             ("configs[0] playground 500k k=20 query-centred", synth.playground(1, 30000 if small else 500_000), 20, 0),
             ("configs[0] playground 500k k=20 mean-centred", synth.playground(1, 30000 if small else 500_000), 20, 1),
             ("configs[1] one VLP-16 frame k=10", synth.vlp16_frame(2), 10, 0)]
    for name, pts, k, mode in cases:
        p.set_cloud(pts)
        p.build_grid(0.0, k)
        got = p.knn_pca(k, mode)
        ref = O.knn_pca_cloud(pts, k, mode, nthreads=nthr)
        truth = None
        if mode == 0:
            idx, _ = O.KdTree(pts).knn_all(k, nthr)
            sel = np.arange(0, pts.shape[0], max(1, pts.shape[0] // 100000))
            w, _ = moment_truth(pts, idx[sel], pts[sel, :3])
            r = stats(name, got[sel], ref[sel], w)
            r["note"] = f"fp64-truth statistics on {sel.size} sampled records"
            r_all = stats(name + " (all records)", got, ref)
            r["all_records"] = {a: r_all[a] for a in ("records", "ptID_ptNum_equal", "dlambda_over_lmax", "frac_lambda_outside_survey_bar_floor_1e-3",
                                                      "frac_lambda_outside_test_bar_floor_1e-2", "v1_1_minus_absdot_gap_ge_1e-3")}
        else:
            r = stats(name, got, ref)
        res[name] = r
    # configs[3] density: radius 0.5 m, 13-neighbour gate, stale records
    n3 = 40000 if small else 1_000_000
    pts = synth.urban(4, n3, extent=100.0 * (n3 / 1e6) ** 0.5)  # 100 pts/m^2 as the 100 M-point cloud over 1 km x 1 km
    p.set_cloud(pts)
    p.build_grid(0.5005, 0)
    got = p.radius_pca(0.5, 13, 0, True)
    ref, valid = O.radius_pca_cloud(pts, 0.5, 13, True, nthreads=nthr)
    res["configs[3] density tile r=0.5"] = stats("configs[3] density tile r=0.5", got, ref, None, valid)
json.dump(res, open(out_path, "w"), indent=1)
print("written", out_path)
