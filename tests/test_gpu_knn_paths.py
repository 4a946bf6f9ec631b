"""GPU: the k-NN kernels (tile kernel of knn_tile6.cu + work-list kernel = the product path; PCP_KNN_IMPL=tile4 / v3 =
the round-1 tile kernel and the warp-per-query kernel kept for A/B timing; the variable is read when a context is
created) must give bit-identical neighbour lists, equal to the oracle's, on
the inputs that stress the hand-over between them: exact ties, duplicates, sparse and tiny clouds, k
larger than the cloud, clouds thinner than one cell, non-finite points."""
import os

import numpy as np
import pytest

from pointcloudprocessing_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _knn(pts, k, impl=None, cell=0.0, env=None):
    old = {v: os.environ.get(v) for v in ("PCP_KNN_IMPL", "PCP_KNN_GROUPS")}
    try:
        if impl:
            os.environ["PCP_KNN_IMPL"] = impl
        for a, b in (env or {}).items():
            os.environ[a] = b
        with capi.Pcp(0) as p:
            p.set_cloud(pts)
            p.build_grid(cell, k)
            idx, d2 = p.knn_neighbors(k)
            feat = p.knn_pca(k)
        return idx, d2, feat
    finally:
        for v, x in old.items():
            if x is None:
                os.environ.pop(v, None)
            else:
                os.environ[v] = x


def _clouds():
    rng = np.random.default_rng(7)
    lattice = np.stack(np.meshgrid(np.arange(24), np.arange(24), np.arange(3), indexing="ij"), -1).reshape(-1, 3)
    lat = np.zeros((lattice.shape[0], 4), np.float32)
    lat[:, :3] = lattice * 0.25  # exact ties everywhere
    dup = synth.urban(5, 4000, extent=15.0)
    dup[1000:1400] = dup[17]  # 401 copies of one point: more ties than a lane can hold
    sparse = np.zeros((3000, 4), np.float32)
    sparse[:, :3] = rng.uniform(0, 200, (3000, 3))  # ~1 point per 2600 m^3: every query expands
    line = np.zeros((500, 4), np.float32)
    line[:, 0] = np.linspace(0, 50, 500)  # thinner than one cell in two axes
    mixed = synth.urban(9, 20000, extent=35.0)
    mixed[::997, 2] = np.nan  # non-finite points are nobody's neighbour
    return {"lattice": lat, "duplicates": dup, "sparse": sparse, "line": line, "mixed": mixed,
            "frame": synth.vlp16_frame(4)[:12000].copy()}


@pytest.mark.parametrize("name", ["lattice", "duplicates", "sparse", "line", "mixed", "frame"])
def test_knn_matches_oracle_and_other_kernels(oracle, name):
    pts = _clouds()[name]
    k = 20
    fin = np.isfinite(pts[:, :3]).all(1)
    ridx, rd2 = oracle.KdTree(pts[fin]).knn_all(k)
    remap = np.nonzero(fin)[0].astype(np.int32)
    ridx = np.where(ridx >= 0, remap[np.maximum(ridx, 0)], -1)
    idx, d2, feat = _knn(pts, k)
    np.testing.assert_array_equal(idx[fin], ridx)
    np.testing.assert_array_equal(d2[fin].view(np.uint32), rd2.view(np.uint32))
    assert (idx[~fin] == -1).all()
    for impl, env in (("v3", None), ("tile4", None), ("tile4", {"PCP_KNN_GROUPS": "4"}), (None, {"PCP_KNN_GROUPS": "4"})):
        i2, e2, f2 = _knn(pts, k, impl, env=env)
        np.testing.assert_array_equal(i2, idx)
        np.testing.assert_array_equal(e2.view(np.uint32), d2.view(np.uint32))
        assert f2.tobytes() == feat.tobytes(), f"records differ between kernels ({impl}, {env})"


@pytest.mark.parametrize("n,k", [(1, 5), (7, 20), (19, 20), (33, 32), (100, 1)])
def test_knn_tiny_clouds(oracle, n, k):
    pts = synth.urban(11, 200, extent=5.0)[:n].copy()
    ridx, rd2 = oracle.KdTree(pts).knn_all(k)
    idx, d2, _ = _knn(pts, k)
    np.testing.assert_array_equal(idx, ridx)
    np.testing.assert_array_equal(d2.view(np.uint32), rd2.view(np.uint32))


@pytest.mark.parametrize("cell", [0.3, 0.8, 3.0])
def test_knn_any_cell_size(oracle, cell):
    pts = synth.urban(13, 15000, extent=30.0)
    ridx, _ = oracle.KdTree(pts).knn_all(30)
    idx, _, _ = _knn(pts, 30, cell=cell)
    np.testing.assert_array_equal(idx, ridx)


def test_host_async_pipelining(oracle):
    """Two contexts in pcp_set_host_async mode, clouds in flight on both: same records as synchronous calls."""
    k = 20
    clouds = [synth.urban(20 + i, 25000, extent=40.0) for i in range(4)]
    with capi.Pcp(0) as p:
        want = []
        for c in clouds:
            p.set_cloud(c)
            p.build_grid(0.0, k)
            want.append(p.knn_pca(k))
    with capi.Pcp(0) as a, capi.Pcp(0) as b:
        ctxs = [a, b]
        outs = [np.zeros(c.shape[0], capi.FEATURE_DTYPE) for c in clouds]
        for c in ctxs:
            c.set_host_async(True)
        for i, c in enumerate(clouds):
            x = ctxs[i & 1]
            x.synchronize()
            x.set_cloud_host_ptr(c.ctypes.data, c.shape[0], 16)
            x.build_grid(0.0, k)
            x.knn_pca(k, out_host_ptr=outs[i].ctypes.data)
        for c in ctxs:
            c.synchronize()
    for got, ref in zip(outs, want):
        assert got.tobytes() == ref.tobytes()


@pytest.mark.parametrize("impl", ["tile", "tile64", "v3"])
def test_radius_pca_both_kernels(oracle, impl):
    """radius + PCA records: tile kernel (+ work list for neighbourhoods above 32) and the warp-per-query
    kernel give the same bytes, equal to the oracle's within the PCA tolerances."""
    import sys
    from parity import check_features
    pts = synth.urban(17, 40000, extent=22.0)  # ~83 pts/m^2: many neighbourhoods above and below 32
    old = os.environ.get("PCP_RADIUS_IMPL")
    try:
        res = {}
        for which in (impl, "v3"):  # the variable is read when a context is created
            os.environ["PCP_RADIUS_IMPL"] = which
            with capi.Pcp(0) as p:
                p.set_cloud(pts)
                p.build_grid(0.5005, 0)
                res[which] = p.radius_pca(0.5, 13, capi.CENTER_QUERY, True)
        got, ref_gpu = res[impl], res["v3"]
    finally:
        if old is None:
            os.environ.pop("PCP_RADIUS_IMPL", None)
        else:
            os.environ["PCP_RADIUS_IMPL"] = old
    assert got.tobytes() == ref_gpu.tobytes()
    ref, valid = oracle.radius_pca_cloud(pts, 0.5, 13, True, nthreads=1)
    assert (got["ptNum"] > 32).mean() > 0.05 and (got["ptNum"] <= 32).mean() > 0.05
    check_features(got, ref, valid, f"radius {impl}")
