"""Golden vectors (tests/golden/golden_small.npz, made by tests/golden/make_golden.py from the oracle):
the oracle must keep reproducing them (CPU suite) and the CUDA path must match them (GPU suite)."""
import os

import numpy as np
import pytest

from parity import check_features

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_small.npz"))


def _feat(oracle, raw):
    return np.ascontiguousarray(raw).view(oracle.FEATURE_DTYPE).reshape(-1)


def test_oracle_reproduces_golden(oracle):
    t = oracle.KdTree(G["urban"])
    idx, d2 = t.knn_all(20)
    np.testing.assert_array_equal(idx, G["urban_knn20_idx"])
    np.testing.assert_array_equal(d2.view(np.uint32), G["urban_knn20_d2"].view(np.uint32))
    off, ridx, _ = t.radius_all(0.5)
    np.testing.assert_array_equal(off, G["urban_r05_off"])
    np.testing.assert_array_equal(ridx, G["urban_r05_idx"])
    f, valid = oracle.radius_pca_cloud(G["urban"], 0.5, 13, True, nthreads=1)
    assert f.tobytes() == G["urban_r05_pca"].tobytes()
    assert oracle.knn_pca_cloud(G["frame"], 10, 0).tobytes() == G["frame_knn10_pca"].tobytes()
    rng, pti, _, lr = oracle.ring_project(G["ring_in"])
    np.testing.assert_array_equal(pti, G["ring_index"])
    assert lr == int(G["ring_long"])


@pytest.mark.gpu
def test_cuda_matches_golden(pcp, oracle):
    pcp.set_cloud(G["urban"])
    pcp.build_grid(0.0, 20)
    idx, d2 = pcp.knn_neighbors(20)
    np.testing.assert_array_equal(idx, G["urban_knn20_idx"])
    np.testing.assert_array_equal(d2.view(np.uint32), G["urban_knn20_d2"].view(np.uint32))
    check_features(pcp.knn_pca(20), _feat(oracle, G["urban_knn20_pca"]), None, "golden knn20")
    pcp.build_grid(0.5001, 0)
    off, ridx, rd2 = pcp.radius_neighbors(0.5)
    np.testing.assert_array_equal(off, G["urban_r05_off"])
    np.testing.assert_array_equal(ridx, G["urban_r05_idx"])
    np.testing.assert_array_equal(rd2.view(np.uint32), G["urban_r05_d2"].view(np.uint32))
    got = pcp.radius_pca(0.5, 13, 0, True)
    ref, valid = _feat(oracle, G["urban_r05_pca"]), G["urban_r05_valid"]
    check_features(got, ref, valid, "golden radius")
    first = int(np.argmax(valid))
    stale = (~valid.astype(bool)) & (np.arange(valid.size) > first)
    check_features(got[stale], ref[stale], None, "golden radius stale")
    pcp.set_cloud(G["frame"])
    pcp.build_grid(0.0, 10)
    idx, _ = pcp.knn_neighbors(10)
    np.testing.assert_array_equal(idx, G["frame_knn10_idx"])
    check_features(pcp.knn_pca(10), _feat(oracle, G["frame_knn10_pca"]), None, "golden frame knn10")
    pcp.set_cloud(G["ring_in"])
    rng, pti, lr = pcp.ring_project()
    np.testing.assert_array_equal(pti[0], G["ring_index"])
    np.testing.assert_array_equal(rng[0].view(np.uint32), G["ring_range"].view(np.uint32))
    assert int(lr[0]) == int(G["ring_long"])


VERT_16DB = np.array(([-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15] * 2) + [0] * 32, np.float32)  # 16db.xml:813-1249


@pytest.mark.parametrize("name", ["sweep", "ragged", "upper_bank"])
def test_decode_oracle_is_pinned_to_the_reference_binary(oracle, name):
    """tests/golden/decoder_*.npz hold what the reference's own lib/libPacketDecoder.so produced for these
    packets (oracle/ref_decoder/make_golden.py).  The restated decoder must agree bit for bit, doubles included,
    and split the frames at the same points."""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", f"decoder_{name}.npz"))
    got = oracle.decode_packets(g["packets"], VERT_16DB)
    nc = got["n_complete"]
    assert nc == g["frame_offsets"].shape[0] - 1
    np.testing.assert_array_equal(got["frame_offsets"][: nc + 1], g["frame_offsets"])
    n = int(g["frame_offsets"][-1])
    for key in ("x", "y", "z", "distance"):
        np.testing.assert_array_equal(got[key][:n].view(np.uint64), g[key].view(np.uint64), err_msg=key)
    for key in ("intensity", "laser_id", "azimuth", "ms"):
        np.testing.assert_array_equal(got[key][:n], g[key], err_msg=key)


# ---- the rows added last (tests/golden/golden_consumers.npz, made by tests/golden/make_golden_consumers.py) ----
GC = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_consumers.npz"))


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def test_oracle_reproduces_consumer_golden(oracle):
    fr = GC["odom_in"]
    cnt = oracle.KdTree(fr).radius_counts(3.0)
    np.testing.assert_array_equal(np.minimum(cnt, 6), GC["odom_counts_capped"])
    rng, pid, pti, stats, mm = oracle.odom_range_image(fr, cnt)
    np.testing.assert_array_equal(_bits(rng), _bits(GC["odom_range"]))
    np.testing.assert_array_equal(pti, GC["odom_index"])
    np.testing.assert_array_equal(stats, GC["odom_stats"])
    np.testing.assert_array_equal(mm, GC["odom_minmax"])
    sm, ty = oracle.odom_smoothness(rng)
    np.testing.assert_array_equal(_bits(sm), _bits(GC["odom_smooth"]))
    np.testing.assert_array_equal(ty, GC["odom_type"])
    urban = GC["db_in"]
    flat = urban.copy()
    flat[:, 2] = 0.0
    lab, nc = oracle.dbscan(flat, 0.3, 8)
    assert nc == int(GC["db_n"])
    np.testing.assert_array_equal(lab, GC["db_labels"])
    nxt, _, mcnt = oracle.mean_shift_step(urban, urban, 2.0, 15)
    np.testing.assert_array_equal(_bits(nxt), _bits(GC["ms_next"]))
    np.testing.assert_array_equal(mcnt, GC["ms_counts"])
    np.testing.assert_array_equal(_bits(oracle.txt2pc(GC["txt_in"].tobytes())), _bits(GC["txt_points"]))


@pytest.mark.gpu
def test_cuda_matches_consumer_golden(pcp):
    fr = GC["odom_in"]
    pcp.set_cloud(fr)
    pcp.build_grid(3.0, 0)
    np.testing.assert_array_equal(pcp.radius_count(3.0, 6), GC["odom_counts_capped"])
    od = pcp.odom_range_image()
    np.testing.assert_array_equal(_bits(od["range"][0]), _bits(GC["odom_range"]))
    np.testing.assert_array_equal(_bits(od["ptid"][0]), _bits(GC["odom_ptid"]))
    np.testing.assert_array_equal(od["pt_index"][0], GC["odom_index"])
    np.testing.assert_array_equal(od["stats"][0], GC["odom_stats"])
    np.testing.assert_array_equal(od["minmax"][0], GC["odom_minmax"])
    sm, ty = pcp.odom_smoothness(n_frames=1)
    np.testing.assert_array_equal(_bits(sm[0]), _bits(GC["odom_smooth"]))
    np.testing.assert_array_equal(ty[0], GC["odom_type"])
    urban = GC["db_in"]
    flat = urban.copy()
    flat[:, 2] = 0.0
    pcp.set_cloud(flat)
    pcp.build_grid(0.3, 0)
    lab, nc = pcp.dbscan(0.3, 8)
    assert nc == int(GC["db_n"])
    np.testing.assert_array_equal(lab, GC["db_labels"])
    pcp.set_cloud(urban)
    cen, lo, hi, cc = pcp.cluster_boxes(lab, nc)
    np.testing.assert_array_equal(cc, GC["db_count"])
    np.testing.assert_array_equal(_bits(lo), _bits(GC["db_min"]))
    np.testing.assert_array_equal(_bits(hi), _bits(GC["db_max"]))
    assert np.abs(cen - GC["db_centroid"]).max() <= 2e-4
    pcp.build_grid(2.0, 0)
    got, cnt = pcp.mean_shift(2.0, 15, 1)
    np.testing.assert_array_equal(cnt, GC["ms_counts"])
    assert np.abs(got[:, :3] - GC["ms_next"][:, :3]).max() <= 2e-5
    got2, cnt2 = pcp.mean_shift(2.0, 15, 1, start=GC["ms_next"])
    np.testing.assert_array_equal(cnt2, GC["ms_counts2"])
    assert np.abs(got2[:, :3] - GC["ms_next2"][:, :3]).max() <= 2e-5
    np.testing.assert_array_equal(_bits(pcp.set_cloud_txt(GC["txt_in"].tobytes())), _bits(GC["txt_points"]))
