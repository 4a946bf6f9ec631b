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


def test_error_behaviour_of_new_entry_points():
    """Status codes instead of crashes: no cloud / no grid / bad arguments / sharded or batched clouds where undefined."""
    import ctypes as C

    from pointcloudprocessing_b200 import capi

    with capi.Pcp(0) as p:
        lib, h = p.lib, p.h
        buf = np.zeros(64, np.int32)
        nc = C.c_int32()
        prm = capi.odom_params_default()
        img = np.zeros((16, 900), np.float32)
        assert lib.pcp_radius_count(h, 1.0, 0, buf.ctypes.data, 0) == -4          # PCP_ERR_NO_CLOUD
        assert lib.pcp_dbscan(h, 1.0, 5, buf.ctypes.data, C.addressof(nc), 0) == -4
        pts = synth.urban(2, 64, extent=5.0)
        p.set_cloud(pts)
        assert lib.pcp_radius_count(h, 1.0, 0, buf.ctypes.data, 0) == -5          # PCP_ERR_NO_GRID
        assert lib.pcp_mean_shift(h, 1.0, 5, 1, None, 0, img.ctypes.data, None, 0) == -5
        assert lib.pcp_odom_range_image(h, C.byref(prm), None, 1, img.ctypes.data, None, None, None, None, 0) == -5
        p.build_grid(1.0, 0)
        assert lib.pcp_radius_count(h, -1.0, 0, buf.ctypes.data, 0) == -1         # PCP_ERR_INVALID_ARG
        assert lib.pcp_dbscan(h, 1.0, 0, buf.ctypes.data, C.addressof(nc), 0) == -1
        assert lib.pcp_odom_range_image(h, C.byref(prm), None, 2, img.ctypes.data, None, None, None, None, 0) == -1  # offsets required
        bad = capi.odom_params_default()
        bad.n_blocks = 0
        assert lib.pcp_odom_smoothness(h, C.byref(bad), img.ctypes.data, 0, 1, None, img.ctypes.data, 0) == -1
        assert lib.pcp_odom_smoothness(h, C.byref(prm), None, 0, 3, None, img.ctypes.data, 0) == -1   # no images of 3 frames here
        assert lib.pcp_knn_pca(h, 65, 0, buf.ctypes.data, None, 0) == -7          # PCP_ERR_K_UNSUPPORTED
        n = C.c_int64(-1)
        assert lib.pcp_set_cloud_txt(h, None, 0, 0, C.addressof(n), None, 0, 0) == 0 and n.value == 0
        assert lib.pcp_set_cloud_txt(h, None, 5, 0, C.addressof(n), None, 0, 0) == -1
        assert b"pcp_set_cloud_txt" in lib.pcp_last_error(h)
