"""GPU parity: the radius-search consumers — DynaTools::clusterDBSCAN (include/dbscanCLuster.h:256-345) and
buildingFacadeExtractor::meanShiftClustering (src/buildingFacadeExtraction.cpp:59-132) — through the C-ABI."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import synth

pytestmark = pytest.mark.gpu


def _flat(pts):
    f = np.ascontiguousarray(pts.copy())
    f[:, 2] = 0.0  # cloud2d (:264-267)
    return f


@pytest.mark.parametrize("seed,n,r,min_pts", [(3, 20000, 0.3, 8), (4, 20000, 0.5, 20), (5, 30000, 0.2, 4), (6, 8000, 0.4, 12)])
def test_dbscan_labels_bit_exact(pcp, oracle, seed, n, r, min_pts):
    """Labels equal the literal sequential walk: cluster numbering by seed order, border points by the first-seeded
    cluster, border points the seed loop reaches first stay noise."""
    pts = synth.urban(seed, n, extent=30.0)
    pts = np.ascontiguousarray(pts[np.random.default_rng(seed).permutation(n)])  # index order unrelated to space
    flat = _flat(pts)
    ref, ref_nc = oracle.dbscan(flat, r, min_pts)
    pcp.set_cloud(flat)
    for cell in (r, 0.6 * r):
        pcp.build_grid(cell, 0)
        lab, nc = pcp.dbscan(r, min_pts)
        assert nc == ref_nc
        np.testing.assert_array_equal(lab, ref)
    assert ref_nc > 3 and (ref == -2).any()
    # border points that stay noise although a core point is in reach exist in these clouds (the quirk is exercised)
    pcp.set_cloud(pts)  # the original 3-D cloud for the boxes (:312-325)
    cen, lo, hi, cnt = pcp.cluster_boxes(lab, nc)
    rcen, rlo, rhi, rcnt = oracle.cluster_boxes(pts, ref, ref_nc)
    np.testing.assert_array_equal(cnt, rcnt)
    np.testing.assert_array_equal(lo.view(np.uint32), rlo.view(np.uint32))
    np.testing.assert_array_equal(hi.view(np.uint32), rhi.view(np.uint32))
    # centroid: double mean vs the reference's fp32 running sum — within fp32 rounding of a sum of |x| <= 30 over cnt terms
    assert np.abs(cen - rcen).max() <= 1e-4 * 30.0


def test_dbscan_double_radius_as_the_reference_passes_it(pcp, oracle):
    """pcp_dbscan_d: the radius stays the `double r` of DynaTools::clusterDBSCAN (include/dbscanCLuster.h:256; call site 0.2,
    src/dynaObjTest.cpp:133), so r2 = (float)(r * r) in double = 0x3d23d70a, one ulp below the square of 0.2f: a pair at
    exactly that squared distance is no neighbour pair for the reference; a cloud at the call site's parameters gets the
    labels of the literal walk with that r2."""
    dx, dy = np.float32(0.19999999), np.float32(6.1e-5)
    pair = np.array([[0, 0, 0, 0], [dx, dy, 0, 0]], np.float32)
    pcp.set_cloud(pair)
    pcp.build_grid(0.2 * 1.0001, 0)
    assert pcp.dbscan_d(0.2, 2)[1] == 0 == oracle.dbscan_d(pair, 0.2, 2)[1]
    assert pcp.dbscan(0.2, 2)[1] == 1 == oracle.dbscan(pair, 0.2, 2)[1]
    pts = synth.urban(7, 40000, extent=25.0)
    # coordinates on a 4 cm lattice: many pairs sit five steps apart, at squared distances one ulp around 0.04
    flat = _flat(pts)
    flat[:, :2] = (np.round(flat[:, :2] / 0.04) * np.float32(0.04)).astype(np.float32)
    ref, ref_nc = oracle.dbscan_d(flat, 0.2, 8)
    narrowed, _ = oracle.dbscan(flat, 0.2, 8)
    assert (ref != narrowed).sum() > 1000  # the two radii label this cloud differently: the test can tell them apart
    pcp.set_cloud(flat)
    pcp.build_grid(0.2 * 1.0001, 0)
    lab, nc = pcp.dbscan_d(0.2, 8)
    assert nc == ref_nc and ref_nc > 0
    np.testing.assert_array_equal(lab, ref)
    lab_f, _ = pcp.dbscan(0.2, 8)
    np.testing.assert_array_equal(lab_f, narrowed)


def test_dbscan_3d_duplicates_and_degenerate(pcp, oracle):
    rs = np.random.default_rng(9)
    base = rs.uniform(0, 4, (3000, 4)).astype(np.float32)
    pts = np.concatenate([base, base[:500], base[:100]], 0)  # exact duplicates (distance 0 ties)
    pts = np.ascontiguousarray(pts)
    pcp.set_cloud(pts)
    pcp.build_grid(0.35, 0)
    for mp in (1, 5, 40, 100000):
        ref, ref_nc = oracle.dbscan(pts, 0.35, mp)
        lab, nc = pcp.dbscan(0.35, mp)
        assert nc == ref_nc
        np.testing.assert_array_equal(lab, ref)
    # min_pts = 1: every point is core; huge min_pts: everything noise
    assert (lab == -2).all()


def test_mean_shift_sweeps(pcp, oracle):
    """Each sweep from the oracle's state: neighbour counts exact, positions within 1e-5 of the radius; the
    tree indexes the ORIGINAL positions while queries and differences use the current ones (:63-64, :79, :96-98)."""
    pts = synth.urban(8, 20000, extent=30.0)
    r, mp = 2.0, 15
    pcp.set_cloud(pts)
    pcp.build_grid(r, 0)
    cur = pts
    for sweep in range(3):
        ref_next, ref_shift, ref_cnt = oracle.mean_shift_step(pts, cur, r, mp)
        got, cnt = pcp.mean_shift(r, mp, 1, start=cur)
        np.testing.assert_array_equal(cnt, ref_cnt)
        assert np.abs(got[:, :3] - ref_next[:, :3]).max() <= 1e-5 * r
        np.testing.assert_array_equal(got[:, 3], pts[:, 3])
        moved = np.abs(ref_shift).max(1) > 0
        assert moved.sum() > 1000 and (~moved).sum() >= 0
        cur = ref_next
    # three sweeps in one call, from the cloud itself: same end state up to the accumulated rounding (a point
    # whose neighbour sits within rounding of the radius may flip: allow a handful)
    got3, _ = pcp.mean_shift(r, mp, 3)
    err = np.abs(got3[:, :3] - cur[:, :3]).max(1)
    assert (err > 1e-4 * r).sum() <= 20
    # a finer grid gives the same sweep
    pcp.build_grid(0.7, 0)
    got_f, cnt_f = pcp.mean_shift(r, mp, 1)
    got_c, cnt_c = pcp.mean_shift(r, mp, 0)  # do-while: zero iterations still sweeps once
    np.testing.assert_array_equal(cnt_f, cnt_c)
    ref_next, _, ref_cnt = oracle.mean_shift_step(pts, pts, r, mp)
    np.testing.assert_array_equal(cnt_f, ref_cnt)
    assert np.abs(got_f[:, :3] - ref_next[:, :3]).max() <= 1e-5 * r


def test_mean_shift_sparse_and_outside(pcp, oracle):
    """Fewer than min_pts neighbours: the point stays (:81-85); start positions outside the box of the cloud."""
    rs = np.random.default_rng(2)
    pts = rs.uniform(0, 10, (4000, 4)).astype(np.float32)
    pts[:50, :3] += 40.0  # isolated
    pts = np.ascontiguousarray(pts)
    start = pts.copy()
    start[100:140, :3] += np.float32(1.5)
    start[140:150, :3] = np.float32(-25.0)  # far outside
    start[150, :3] = np.float32(12.0)       # just outside the upper corner
    pcp.set_cloud(pts)
    pcp.build_grid(2.0, 0)
    ref_next, _, ref_cnt = oracle.mean_shift_step(pts, start, 2.0, 15)
    got, cnt = pcp.mean_shift(2.0, 15, 1, start=start)
    np.testing.assert_array_equal(cnt, ref_cnt)
    assert np.abs(got[:, :3] - ref_next[:, :3]).max() <= 2e-5
    assert (ref_cnt < 15).sum() >= 50
