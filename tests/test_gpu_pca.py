"""GPU parity: fused search + PCA records against the oracle (reference path restated) — C-ABI."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import capi, synth
from pointcloudprocessing_b200.capi import CENTER_MEAN, CENTER_QUERY
from parity import check_features

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k,seed", [(20, 1), (10, 2), (30, 3)])
def test_knn_pca_query_centre(pcp, oracle, k, seed):
    pts = synth.urban(seed, 30000, extent=40.0)
    pcp.set_cloud(pts)
    pcp.build_grid(0.0, k)
    got = pcp.knn_pca(k, CENTER_QUERY)
    ref = oracle.knn_pca_cloud(pts, k, 0)
    stats = check_features(got, ref, None, f"knn k={k}")
    assert stats["v1"][0] > 1000


def test_knn_pca_mean_centre_and_normals(pcp, oracle):
    pts = synth.urban(4, 20000, extent=30.0)
    pts[:, :3] += np.float32(3.0)  # keep the un-centred PCL formula well conditioned
    k = 15  # src/smoothing.cpp:245
    pcp.set_cloud(pts)
    pcp.build_grid(0.0, k)
    got, nrm = pcp.knn_pca(k, CENTER_MEAN, want_normals=True)
    ref, rnrm = oracle.knn_pca_cloud(pts, k, 1, want_normals=True)
    check_features(got, ref, None, "knn mean")
    # the PCL twin solves in fp32 with the trigonometric closed form and an un-centred moment: its
    # own error is ~1e-5·λmax, so normals are compared where the plane is well defined
    la = ref["la"].astype(np.float64)
    ok = (la[:, 1] - la[:, 0]) > 0.05 * la[:, 2]
    dots = np.abs((nrm[:, :3].astype(np.float64) * rnrm[:, :3]).sum(1))
    assert ok.sum() > 1000
    assert (1 - dots[ok]).max() < 2e-3
    # orientation: toward the origin
    assert ((nrm[:, :3] * (-pts[:, :3])).sum(1) >= -1e-6).all()
    curv_ref = la[:, 0] / la.sum(1)
    assert np.abs(nrm[ok, 3] - curv_ref[ok]).max() < 1e-4


@pytest.mark.parametrize("carry", [True, False])
def test_radius_pca_cloud(pcp, oracle, carry):
    """calculatePCAofPointCloud (src/pcaAnalysis.cpp:11-42) incl. the <13 gate and stale records."""
    pts = synth.urban(6, 30000, extent=30.0)
    r = 0.5
    pcp.set_cloud(pts)
    pcp.build_grid(r * 1.001, 0)
    got = pcp.radius_pca(r, 13, CENTER_QUERY, carry)
    ref, valid = oracle.radius_pca_cloud(pts, r, 13, carry, nthreads=1)
    assert 0.2 < valid.mean() < 0.98
    check_features(got, ref, valid, "radius valid")
    inv = ~valid.astype(bool)
    if carry:
        first = int(np.argmax(valid))
        lead = np.arange(pts.shape[0]) < first
        check_features(got[inv & ~lead], ref[inv & ~lead], None, "radius stale")
        for name in ("la", "v1", "lineStruc"):
            assert not got[name][lead].any()
    else:
        for name in ("la", "v1", "v2", "v3", "lineStruc", "planeStruc", "scatterStruc"):
            assert not got[name][inv].any()


def test_radius_pca_moment_bitwise(pcp, oracle):
    """With ≥13 neighbours on an exact plane z=0 the fp32 moment is reproduced bit for bit, so the
    in-plane eigenvalues match the oracle's to fp32 round-off and λmin is ~0."""
    g = np.arange(40, dtype=np.float32) * np.float32(0.1)
    x, y = np.meshgrid(g, g, indexing="ij")
    pts = np.stack([x.ravel(), y.ravel(), np.zeros(x.size, np.float32), np.zeros(x.size, np.float32)], 1)
    pcp.set_cloud(pts)
    pcp.build_grid(0.36, 0)
    got = pcp.radius_pca(0.35, 13, CENTER_QUERY, False)
    ref, valid = oracle.radius_pca_cloud(pts, 0.35, 13, False)
    np.testing.assert_array_equal(got["ptNum"], ref["ptNum"])
    v = valid.astype(bool)
    assert np.abs(got["la"][v, 0]).max() < 1e-9
    np.testing.assert_allclose(got["la"][v, 1:], ref["la"][v, 1:], rtol=2e-6)
    assert np.abs(np.abs(got["v1"][v, 2]) - 1).max() < 1e-6


def test_segment_pca(pcp, oracle):
    """calculatePCAofPoint(cloud, centro, idx, fea) (src/pcaAnalysis.cpp:89-121): voxel segments."""
    pts = synth.urban(7, 20000, extent=30.0)
    vox = np.floor(pts[:, :3] / 3.0).astype(np.int64)
    key = (vox[:, 0] * 1000 + vox[:, 1]) * 1000 + vox[:, 2]
    order = np.argsort(key, kind="stable")
    uk, start, counts = np.unique(key[order], return_index=True, return_counts=True)
    keep = counts >= 20  # src/buildingFacadeExtraction.cpp:172
    offs = np.zeros(keep.sum() + 1, np.int64)
    np.cumsum(counts[keep], out=offs[1:])
    idx = np.concatenate([order[s:s + c] for s, c in zip(start[keep], counts[keep])]).astype(np.int32)
    centres = (np.stack([uk // 1000000, (uk // 1000) % 1000, uk % 1000], 1)[keep] + 0.5).astype(np.float32) * 3.0
    pcp.set_cloud(pts)
    got = pcp.segment_pca(centres, offs, idx)
    ref = oracle.segment_pca(pts, centres, offs, idx)
    check_features(got, ref, None, "segment")


def test_host_stride_32(pcp, oracle):
    """pcl::PointXYZI host layout: 32-byte stride, xyz + pad + intensity + pad."""
    pts = synth.urban(8, 5000, extent=15.0)
    aos = np.zeros((pts.shape[0], 8), np.float32)
    aos[:, :3] = pts[:, :3]
    aos[:, 3] = 1.0
    aos[:, 4] = pts[:, 3]
    pcp.set_cloud(aos)
    pcp.build_grid(0.0, 10)
    idx, _ = pcp.knn_neighbors(10)
    ridx, _ = oracle.KdTree(pts).knn_all(10)
    np.testing.assert_array_equal(idx, ridx)


def test_eigen_closer_to_fp64_truth(pcp, oracle):
    """The in-kernel fp64 Jacobi solve of the (bitwise-equal) fp32 moment lands on the exact
    eigenvalues; the oracle's fp32 QR is the side that carries the few-ulp error."""
    pts = synth.urban(9, 20000, extent=30.0)
    k = 20
    pcp.set_cloud(pts)
    pcp.build_grid(0.0, k)
    got = pcp.knn_pca(k, CENTER_QUERY)
    ref = oracle.knn_pca_cloud(pts, k, 0)
    idx, _ = oracle.KdTree(pts).knn_all(k)
    eg, eo = 0.0, 0.0
    for i in range(0, pts.shape[0], 40):
        Cm = oracle.covariance_about(pts, idx[i], pts[i, :3]).astype(np.float64)
        w = np.linalg.eigvalsh(Cm)
        eg = max(eg, np.abs(got["la"][i] - w).max() / w.max())
        eo = max(eo, np.abs(ref["la"][i] - w).max() / w.max())
    assert eg < 1.2e-7  # fp32 rounding of the exact value
    assert eg <= eo


def test_radius_pca_heavy_neighbourhoods(pcp, oracle):
    """r = 3 m on a dense scan (the radius src/lidarOdometry.cpp:76 uses): thousands of hits per query, far
    beyond the 512-entry warp buffer — records come from the unordered double-precision accumulation, counts
    stay exact, eigen outputs within the parity bar (lists of that size: test_radius_neighbors_heavy_lists)."""
    pts = synth.urban(8, 40000, extent=30.0)
    r = 3.0
    pcp.set_cloud(pts)
    pcp.build_grid(r * 1.001, 0)
    got = pcp.radius_pca(r, 13, CENTER_QUERY, True)
    ref, valid = oracle.radius_pca_cloud(pts, r, 13, True, nthreads=0)
    assert got["ptNum"].max() > 2000 and (got["ptNum"] > 512).mean() > 0.5
    assert np.array_equal(got["ptNum"], ref["ptNum"])
    check_features(got, ref, valid, "radius heavy")
    # a mix: sparse clutter (short lists, bitwise path) next to dense ground (heavy path) in one launch
    r = 1.2
    pcp.build_grid(r * 1.001, 0)
    got = pcp.radius_pca(r, 13, CENTER_QUERY, True)
    ref, valid = oracle.radius_pca_cloud(pts, r, 13, True, nthreads=0)
    assert (got["ptNum"] > 512).any() and (got["ptNum"] < 200).any()
    assert np.array_equal(got["ptNum"], ref["ptNum"])
    check_features(got, ref, valid, "radius mixed")


@pytest.mark.parametrize("div", [3.5, 5.0])
def test_radius_pca_small_cell(pcp, oracle, div):
    """A cell smaller than r / 3 (stencil half-width 4-6: more rows than the tile kernel's run tables hold) must
    take the warp-per-query path and still match the oracle (such grids arise from
    build_grid(0.0, k) followed by radius_pca, or from a small cell chosen by the caller on a sparse cloud)."""
    pts = synth.urban(9, 20000, extent=60.0)  # sparse: ~5 points/m^2, so the density guess stays on the tile side
    r = 1.0
    pcp.set_cloud(pts)
    pcp.build_grid(r / div, 0)
    got = pcp.radius_pca(r, 13, CENTER_QUERY, True)
    ref, valid = oracle.radius_pca_cloud(pts, r, 13, True, nthreads=0)
    assert np.array_equal(got["ptNum"], ref["ptNum"])
    check_features(got, ref, valid, f"radius cell=r/{div}")


def test_bbox_pca(pcp, oracle):
    """getMiniBoundingBox (src/pcaAnalysis.cpp:123-148) as device reductions: against the statement-by-statement restatement
    (whose fp32 running sums carry ~1e-4 on a cloud of this size) and against an fp64 evaluation."""
    pts = synth.urban(5, 200000, extent=60.0)
    pts[::1000, 1] = np.nan  # non-finite points do not enter the sums
    fin = np.isfinite(pts[:, :3]).all(1)
    pcp.set_cloud(pts)
    got = pcp.bbox_pca()
    assert got["n_points"] == int(fin.sum())
    ref = oracle.bbox_pca(pts[fin])
    x = pts[fin, :3].astype(np.float64)
    c = x.mean(0)
    w, v = np.linalg.eigh(np.cov((x - c).T, bias=True))
    assert np.abs(got["centroid"] - c).max() < 1e-5 * 60
    assert np.abs(got["eigval"] - w).max() < 1e-5 * w.max()
    assert np.abs(got["centroid"] - ref["centroid"]).max() < 2e-3 and np.abs(got["eigval"] - ref["eigval"]).max() < 2e-3 * w.max()
    R = got["axes"].astype(np.float64)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-5) and np.linalg.det(R) > 0.999  # right-handed: row 2 = row 0 x row 1
    for kk in range(2):
        assert 1 - abs(R[kk] @ v[:, kk]) < 1e-5
    t = (x - c) @ R.T
    assert np.abs(got["min_pt"] - t.min(0)).max() < 1e-3 and np.abs(got["max_pt"] - t.max(0)).max() < 1e-3
    assert np.allclose(got["mean_diag"], 0.5 * (got["min_pt"] + got["max_pt"]), atol=1e-5)
    # the restatement agrees up to eigenvector signs and its fp32 sums
    s = np.sign((got["axes"] * ref["axes"]).sum(1))
    assert np.abs(got["axes"] - s[:, None] * ref["axes"]).max() < 2e-3



def test_radius_neighbors_heavy_lists(pcp, oracle):
    """Neighbour LISTS beyond the 512-entry on-chip buffer (radiusSearch(i, 3.0): src/lidarOdometry.cpp:76, main.cpp:117;
    r = 2.0: src/buildingFacadeExtraction.cpp:83): count-then-fill through global memory + segmented sort — indices and
    distances bit-exact in (d2, index) order, heavy and light queries mixed in one call."""
    pts = synth.urban(8, 14000, extent=18.0)
    for r in (3.0, 2.2):
        pcp.set_cloud(pts)
        pcp.build_grid(r * 1.001, 0)
        off, idx, d2 = pcp.radius_neighbors(r)
        roff, ridx, rd2 = oracle.KdTree(pts).radius_all(r)
        cnt = np.diff(roff)
        assert (cnt > 512).any() and (r > 2 or (cnt <= 512).any())
        assert np.array_equal(off, roff)
        assert np.array_equal(idx, ridx)
        assert np.array_equal(d2.view(np.uint32), rd2.view(np.uint32))
    # duplicates: more than 512 points at distance 0 of each other (ties ordered by index)
    dup = synth.urban(9, 3000, extent=10.0)
    dup[500:1300] = dup[7]
    pcp.set_cloud(dup)
    pcp.build_grid(0.3, 0)
    off, idx, d2 = pcp.radius_neighbors(0.25)
    roff, ridx, rd2 = oracle.KdTree(dup).radius_all(0.25)
    assert np.diff(roff).max() > 800
    assert np.array_equal(off, roff) and np.array_equal(idx, ridx) and np.array_equal(d2.view(np.uint32), rd2.view(np.uint32))
