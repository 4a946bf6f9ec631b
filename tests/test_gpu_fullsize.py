"""GPU parity at the sizes BASELINE.json states: the bench cloud and configs[2] as written
(1 M-point urban scan, k = 20 and k = 30), configs[0]'s stand-in (synth.playground, 500 k points: k = 20 in both
centre modes with normals, and r = 3 m on a 50 k subset — main.cpp:115-117, src/smoothing.cpp:236-248), a
configs[3]-density tile at r = 0.5 m, and a 128-frame batch of configs[4] with 16 frames checked.  Neighbour
indices and distances bit-exact, records within tests/parity.py's bars, all through the C-ABI."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import capi, frames, synth
from pointcloudprocessing_b200.capi import CENTER_MEAN, CENTER_QUERY
from parity import check_features

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def urban_1m():
    return synth.urban(3, 1_000_000, extent=200.0, order="random")  # bench.py's cloud


@pytest.mark.parametrize("k", [20, 30])
def test_urban_1m_knn(pcp, oracle, urban_1m, k):
    """configs[2] (k = 30) and the metric's own workload (k = 20) at full size."""
    pts = urban_1m
    pcp.set_cloud(pts)
    pcp.build_grid(0.0, k)
    idx, d2 = pcp.knn_neighbors(k)
    ridx, rd2 = oracle.KdTree(pts).knn_all(k)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(d2.view(np.uint32), rd2.view(np.uint32))
    got = pcp.knn_pca(k, CENTER_QUERY)
    stats = check_features(got, oracle.knn_pca_cloud(pts, k, 0), None, f"urban 1M k={k}")
    assert stats["v1"][0] > 500_000


@pytest.fixture(scope="module")
def playground_500k():
    return synth.playground(1, 500_000)


def test_playground_500k_knn_both_centres(pcp, oracle, playground_500k):
    """configs[0] stand-in: pcaAnalysis records (query-centred) and the PCL twin (mean-centred + normals, curvature)."""
    pts = playground_500k
    k = 20
    pcp.set_cloud(pts)
    pcp.build_grid(0.0, k)
    idx, d2 = pcp.knn_neighbors(k)
    ridx, rd2 = oracle.KdTree(pts).knn_all(k)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(d2.view(np.uint32), rd2.view(np.uint32))
    check_features(pcp.knn_pca(k, CENTER_QUERY), oracle.knn_pca_cloud(pts, k, 0), None, "playground query-centred")
    got, nrm = pcp.knn_pca(k, CENTER_MEAN, want_normals=True)
    ref, rnrm = oracle.knn_pca_cloud(pts, k, 1, want_normals=True)
    check_features(got, ref, None, "playground mean-centred")
    la = ref["la"].astype(np.float64)
    # the PCL twin's un-centred fp32 single pass needs a defined plane AND a spectral gap above its own
    # cancellation noise, ~|p|^2 ulps: 60 m from the origin 18 k of these 500 k reference normals are off by more
    # than 2e-3 against an fp64 evaluation of the same neighbourhoods (measured), so the comparison is gated on
    # gap > 1e-5 |p|^2, where the restated PCL arithmetic itself is within 2e-4 of fp64
    r2 = (pts[:, :3].astype(np.float64) ** 2).sum(1)
    ok = ((la[:, 1] - la[:, 0]) > 0.05 * la[:, 2]) & ((la[:, 1] - la[:, 0]) > 1e-5 * r2)
    assert ok.sum() > 50_000
    dots = np.abs((nrm[:, :3].astype(np.float64) * rnrm[:, :3]).sum(1))
    assert (1 - dots[ok]).max() < 2e-3
    assert np.abs(nrm[ok, 3] - la[ok, 0] / la[ok].sum(1)).max() < 1e-4


def test_playground_radius_3m_subset(pcp, oracle, playground_500k):
    """main.cpp:115-117: calculatePCAofPointCloud(keyCloud, 3, ...) — r = 3 m on a 50 k subset."""
    pts = np.ascontiguousarray(playground_500k[::10])
    r = 3.0
    pcp.set_cloud(pts)
    pcp.build_grid(r * 1.001, 0)
    got = pcp.radius_pca(r, 13, CENTER_QUERY, True)
    ref, valid = oracle.radius_pca_cloud(pts, r, 13, True, nthreads=0)
    assert np.array_equal(got["ptNum"], ref["ptNum"]) and np.array_equal(got["ptID"], ref["ptID"])
    assert valid.mean() > 0.99 and got["ptNum"].max() > 512
    check_features(got, ref, valid, "playground r=3")


def test_cfg3_density_tile_radius(pcp, oracle):
    """configs[3]'s density (100 points / m^2 over the ground plane) on a 1 M-point tile: radiusSearch r = 0.5 m + PCA
    with the 13-neighbour gate and the stale records of src/pcaAnalysis.cpp:16,26,47-51; neighbour lists of a
    60 k subset bit-exact."""
    pts = synth.urban(4, 1_000_000, extent=100.0)
    r = 0.5
    pcp.set_cloud(pts)
    pcp.build_grid(r * 1.001, 0)
    got = pcp.radius_pca(r, 13, CENTER_QUERY, True)
    ref, valid = oracle.radius_pca_cloud(pts, r, 13, True, nthreads=0)
    assert np.array_equal(got["ptNum"], ref["ptNum"])
    assert 20 < got["ptNum"].mean() < 120
    check_features(got, ref, valid, "cfg3 tile")
    sub = np.ascontiguousarray(pts[pts[:, 0] < 6.0])
    pcp.set_cloud(sub)
    pcp.build_grid(r * 1.001, 0)
    off, idx, d2 = pcp.radius_neighbors(r)
    roff, ridx, rd2 = oracle.KdTree(sub).radius_all(r)
    assert np.array_equal(off, roff) and np.array_equal(idx, ridx)
    assert np.array_equal(d2.view(np.uint32), rd2.view(np.uint32))


def test_frame_batch_128(oracle):
    """configs[4] shape: 128 VLP-16 frames in one launch (segmented grid), ring projection + k = 10 PCA; 16 of the
    frames compared with the oracle (src/dynaObjTest.cpp:103-106 per frame)."""
    fr = [synth.vlp16_frame(1000 + f, pose_jitter=0.5) for f in range(128)]
    with frames.FrameBatcher(0, n_ctx=1, batch=128) as bat:
        got = bat.run(fr, k=10)
    for f in range(0, 128, 8):
        rrng, rpti, _, rlong = oracle.ring_project(fr[f])
        assert np.array_equal(got[f]["index"], rpti)
        assert np.array_equal(got[f]["range"].view(np.uint32), rrng.view(np.uint32))
        assert got[f]["long_range"] == rlong
        check_features(got[f]["features"], oracle.knn_pca_cloud(fr[f], 10, 0), None, f"frame {f} of 128")
