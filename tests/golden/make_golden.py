"""Generates tests/golden/golden_small.npz from the CPU oracle (oracle/ — the restated reference path).

The reference ships no fixtures for this path (parity unpinned) and cannot be built
here (no PCL/FLANN/Eigen), so these vectors pin the ORACLE, not the reference: they freeze its
outputs at the commit that passed the brute-force and analytic checks of tests/test_oracle_kat.py,
and both the oracle (CPU suite) and the CUDA path (GPU suite) are compared against them.
Inputs are stored too, so the fixtures do not depend on numpy's generator.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from pointcloudprocessing_b200 import synth  # noqa: E402


def main():
    urban = synth.urban(101, 3000, extent=12.0)
    frame = synth.vlp16_frame(102)[::4][:4000].copy()
    out = {"urban": urban, "frame": frame}
    t = O.KdTree(urban)
    out["urban_knn20_idx"], out["urban_knn20_d2"] = t.knn_all(20)
    off, idx, d2 = t.radius_all(0.5)
    out["urban_r05_off"], out["urban_r05_idx"], out["urban_r05_d2"] = off, idx, d2
    f, valid = O.radius_pca_cloud(urban, 0.5, 13, True, nthreads=1)
    out["urban_r05_pca"], out["urban_r05_valid"] = f.view(np.uint8).reshape(-1, 68), valid
    out["urban_knn20_pca"] = O.knn_pca_cloud(urban, 20, 0).view(np.uint8).reshape(-1, 68)
    out["frame_knn10_idx"], _ = O.KdTree(frame).knn_all(10)
    out["frame_knn10_pca"] = O.knn_pca_cloud(frame, 10, 0).view(np.uint8).reshape(-1, 68)
    full = synth.vlp16_frame(103)
    out["ring_in"] = full
    rng, pti, tag, lr = O.ring_project(full)
    out["ring_range"], out["ring_index"], out["ring_long"] = rng, pti, np.int32(lr)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_small.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
