"""GPU: edge cases of the rows added last (odometry images, DBSCAN, mean shift, text clouds, k > 32) against the oracle."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import synth

pytestmark = pytest.mark.gpu


def test_tiny_and_degenerate_clouds(pcp, oracle):
    one = np.array([[1.0, 2.0, 3.0, 4.0]], np.float32)
    same = np.tile(one, (40, 1))
    for pts in (one, same):
        pcp.set_cloud(pts)
        pcp.build_grid(1.0, 0)
        np.testing.assert_array_equal(pcp.radius_count(1.0, 0), oracle.KdTree(pts).radius_counts(1.0))
        for mp in (1, 2, 41):
            lab, nc = pcp.dbscan(1.0, mp)
            rlab, rnc = oracle.dbscan(pts, 1.0, mp)
            assert nc == rnc
            np.testing.assert_array_equal(lab, rlab)
        got, cnt = pcp.mean_shift(1.0, 15, 4)
        ref = pts
        for _ in range(4):
            ref, _, rcnt = oracle.mean_shift_step(pts, ref, 1.0, 15)
        np.testing.assert_array_equal(cnt, rcnt)
        np.testing.assert_allclose(got[:, :3], ref[:, :3], atol=1e-5)
        od = pcp.odom_range_image()
        rng, pid, pti, stats, mm = oracle.odom_range_image(pts, oracle.KdTree(pts).radius_counts(3.0))
        np.testing.assert_array_equal(od["range"][0].view(np.uint32), rng.view(np.uint32))
        np.testing.assert_array_equal(od["pt_index"][0], pti)
        np.testing.assert_array_equal(od["stats"][0], stats)
        idx, d2 = pcp.knn_neighbors(40)
        ridx, rd2 = oracle.KdTree(pts).knn_all(40)
        np.testing.assert_array_equal(idx, ridx)


def test_wide_k_in_frame_batches(pcp, oracle):
    """k = 40 on a segmented grid: the cover-radius search stays inside its frame."""
    pts, offs = synth.vlp16_batch(3, seed0=90)
    keep = np.concatenate([np.arange(offs[f], offs[f + 1])[::6] for f in range(3)])
    sub = np.ascontiguousarray(pts[keep])
    so = np.zeros(4, np.int64)
    np.cumsum([len(np.arange(offs[f], offs[f + 1])[::6]) for f in range(3)], out=so[1:])
    pcp.set_cloud(sub)
    pcp.set_frames(so)
    pcp.build_grid(0.0, 40)
    idx, d2 = pcp.knn_neighbors(40)
    for f in range(3):
        fr = sub[so[f]:so[f + 1]]
        ridx, rd2 = oracle.KdTree(fr).knn_all(40)
        np.testing.assert_array_equal(idx[so[f]:so[f + 1]], ridx + so[f])
        np.testing.assert_array_equal(d2[so[f]:so[f + 1]].view(np.uint32), rd2.view(np.uint32))
    pcp.set_frames(np.array([0, sub.shape[0]], np.int64))


def test_txt_odd_inputs(pcp, oracle):
    rs = np.random.default_rng(8)
    cases = [
        b"\n", b" ", b"\r\n\r\n", b"1e999999999999999999 2 3 4\n1e-999999999999999999 5\n",
        b"0" * 400 + b"1.5 " + b"9" * 300 + b" 0." + b"0" * 200 + b"123 7\n",
        b"1 2 3 4" + b" " * 5000 + b"\n" + b"7\t8\v9\f10\n",
        bytes(rs.integers(0, 256, 20000, dtype=np.uint8)),                                   # binary noise
        bytes(rs.choice(np.frombuffer(b"0123456789.eE+- \n\t", np.uint8), 60000)),          # number-ish noise
        b"+ 1\n+1 -\n-.e1 2\n.e1\n1.e1 2.E-1 3.e+1 4\n1e1.5 2\n",
    ]
    for text in cases:
        ref = oracle.txt2pc(text)
        got = pcp.set_cloud_txt(text)
        assert got.shape == ref.shape, text[:40]
        np.testing.assert_array_equal(got.view(np.uint32), ref.view(np.uint32), err_msg=repr(text[:60]))
