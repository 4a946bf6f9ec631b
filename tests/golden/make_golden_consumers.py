"""Generates tests/golden/golden_consumers.npz from the CPU oracle: the rows of SURVEY §8f added last (lidarOdometry
images, DBSCAN, mean shift, text clouds).  Same status as golden_small.npz: the vectors freeze the ORACLE's outputs
(the reference ships no fixtures); for the text row the oracle IS the reference's own stream statements run by this
image's C++ library.  Inputs are stored too.

    python tests/golden/make_golden_consumers.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from pointcloudprocessing_b200 import synth  # noqa: E402

TEXT = (b"1 2 3 4\n5 6 7\n\n8.5 abc 9 1\n 1e3 .5 -.25e-2 +7\r\n1e 2 3 4\n16777217 16777217.0000000000000001 1e39 5\n"
        b"-0.0 +.0e5 00012.500 1E+2\n1.4e-45 7.1e-46 3.4028235e38 0.1\n12.375 -7.0625 1.6777215e7 255\n3")


def main():
    out = {}
    fr = synth.vlp16_frame(104)[::2].copy()
    out["odom_in"] = fr
    cnt = O.KdTree(fr).radius_counts(3.0)
    out["odom_counts_capped"] = np.minimum(cnt, 6).astype(np.int32)
    rng, pid, pti, stats, mm = O.odom_range_image(fr, cnt)
    sm, ty = O.odom_smoothness(rng)
    out.update(odom_range=rng, odom_ptid=pid, odom_index=pti, odom_stats=stats, odom_minmax=mm, odom_smooth=sm, odom_type=ty)
    urban = synth.urban(105, 6000, extent=18.0)
    flat = urban.copy()
    flat[:, 2] = 0.0
    lab, nc = O.dbscan(flat, 0.3, 8)
    cen, lo, hi, cc = O.cluster_boxes(urban, lab, nc)
    out.update(db_in=urban, db_labels=lab, db_n=np.int32(nc), db_centroid=cen, db_min=lo, db_max=hi, db_count=cc)
    nxt, shift, mcnt = O.mean_shift_step(urban, urban, 2.0, 15)
    nxt2, _, mcnt2 = O.mean_shift_step(urban, nxt, 2.0, 15)
    out.update(ms_next=nxt, ms_counts=mcnt, ms_next2=nxt2, ms_counts2=mcnt2)
    out["txt_in"] = np.frombuffer(TEXT, np.uint8)
    out["txt_points"] = O.txt2pc(TEXT)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_consumers.npz"), **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()})


if __name__ == "__main__":
    main()
