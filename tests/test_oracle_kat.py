"""CPU: pins of the oracle (parity is otherwise unpinned — the reference ships no fixtures):
brute force vs kd-tree, analytic known-answer cases, layout sizes, reference quirks Q3-Q7."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import synth


def test_layout_sizes(oracle):
    from pointcloudprocessing_b200 import capi

    assert oracle.FEATURE_DTYPE.itemsize == 68 == capi.FEATURE_DTYPE.itemsize  # include/pcaAnalysis.h:34-45
    assert [oracle.FEATURE_DTYPE.fields[f][1] for f in ("la", "v1", "ptID", "ptNum", "lineStruc")] == [0, 12, 48, 52, 56]


@pytest.mark.parametrize("k", [1, 10, 20, 30])
def test_kdtree_knn_equals_brute(oracle, k):
    pts = synth.urban(21, 6000, extent=20.0)
    idx, d2 = oracle.KdTree(pts).knn_all(k)
    for q in range(0, 6000, 97):
        bi, bd = oracle.knn_brute(pts, q, k)
        np.testing.assert_array_equal(idx[q], bi)
        np.testing.assert_array_equal(d2[q].view(np.uint32), bd.view(np.uint32))
    assert (idx[:, 0] == np.arange(6000)).all()  # the query is its own nearest neighbour (Q5)


@pytest.mark.parametrize("r", [0.3, 0.5, 2.0])
def test_kdtree_radius_equals_brute(oracle, r):
    pts = synth.urban(22, 5000, extent=20.0)
    off, idx, d2 = oracle.KdTree(pts).radius_all(r)
    for q in range(0, 5000, 83):
        bi, bd = oracle.radius_brute(pts, q, r)
        np.testing.assert_array_equal(idx[off[q]:off[q + 1]], bi)
        np.testing.assert_array_equal(d2[off[q]:off[q + 1]].view(np.uint32), bd.view(np.uint32))


def test_ties_broken_by_index(oracle):
    """Duplicates and lattice points: (d2, index) order, first k of it."""
    pts = np.zeros((10, 4), np.float32)
    pts[5:, 0] = 1.0
    idx, d2 = oracle.knn_brute(pts, 7, 7)
    assert idx.tolist() == [5, 6, 7, 8, 9, 0, 1]
    assert d2.tolist() == [0, 0, 0, 0, 0, 1, 1]
    tidx, _ = oracle.KdTree(pts).knn_all(7)
    assert tidx[7].tolist() == idx.tolist()


def test_radius_is_strict_and_squared_in_double(oracle):
    """Q6: r2 = (float)((double)r*(double)r); membership d2 < r2."""
    r = np.float32(0.1)
    assert oracle.radius2(float(r)) == float(np.float32(np.float64(r) * np.float64(r)))
    pts = np.zeros((3, 4), np.float32)
    pts[1, 0] = 0.5
    pts[2, 0] = np.nextafter(np.float32(0.5), np.float32(0))
    idx, _ = oracle.radius_brute(pts, 0, 0.5)
    assert idx.tolist() == [0, 2]


def test_covariance_about_query_not_mean(oracle):
    """Q2: second moment about the given centre, divisor n (not n−1), PCL accumulation order."""
    pts = np.array([[1, 2, 3, 0], [2, 2, 3, 0], [1, 4, 3, 0], [1, 2, 7, 0]], np.float32)
    Cm = oracle.covariance_about(pts, [0, 1, 2, 3], pts[0, :3])
    np.testing.assert_array_equal(Cm, np.diag([1 / 4, 4 / 4, 16 / 4]).astype(np.float32))


def test_eigen_known_answers(oracle):
    ev, Q, _ = oracle.eigen_sym3_qr(np.diag([3.0, 1.0, 2.0]))
    assert ev.tolist() == [1.0, 2.0, 3.0]  # ascending (Q3)
    assert np.allclose(np.abs(Q), np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0]]))
    ev, Q, _ = oracle.eigen_sym3_qr(np.zeros((3, 3)))
    assert ev.tolist() == [0, 0, 0] and np.allclose(Q, np.eye(3))
    rng = np.random.default_rng(0)
    for _ in range(300):
        A = rng.normal(size=(3, 3)) * rng.choice([1e-3, 1.0, 1e3])
        S = (A @ A.T).astype(np.float32)
        ev, Q, it = oracle.eigen_sym3_qr(S)
        w = np.linalg.eigvalsh(S.astype(np.float64))
        assert np.abs(ev - w).max() <= 1.5e-6 * np.abs(w).max()  # fp32 QR: a few ulps of λmax
        assert np.allclose(Q.T @ Q, np.eye(3), atol=2e-6) and it <= 90
        assert np.abs(S.astype(np.float64) @ Q - Q * ev).max() <= 2e-6 * np.abs(w).max()


def test_planar_patch_and_features(oracle):
    """Plane z = 0 lattice, spacing s: λ = {0, Σx²/n, Σy²/n}; v1 = ±z; the reference's features divide
    by sqrt of the SMALLEST λ (Q3), so they are inf/nan on an exact plane."""
    g = (np.arange(-3, 4) * 0.1).astype(np.float32)
    x, y = np.meshgrid(g, g, indexing="ij")
    pts = np.stack([x.ravel(), y.ravel(), 0 * x.ravel(), 0 * x.ravel()], 1).astype(np.float32)
    f, valid = oracle.radius_pca_cloud(pts, 10.0, 13, True)
    c = f[24]  # centre point: all 49 points are neighbours
    assert valid.all() and c["ptNum"] == 49 and c["ptID"] == 24
    m = float((g.astype(np.float64) ** 2).sum() * 7 / 49)
    assert abs(c["la"][0]) < 1e-8 and np.allclose(c["la"][1:], m, rtol=1e-5)
    assert abs(abs(c["v1"][2]) - 1) < 1e-6
    s = oracle.structure_features(1.0, 4.0, 9.0)
    assert s.tolist() == [-1.0, -1.0, 3.0]  # (√λ1−√λ2)/√λ1, (√λ2−√λ3)/√λ1, √λ3/√λ1


def test_min_pts_gate_and_stale_records(oracle):
    """Q4: < 13 neighbours keeps only ptID/ptNum; with the carry the payload is the last valid one's."""
    rng = np.random.default_rng(3)
    dense = rng.normal(0, 0.05, (40, 3))
    sparse = rng.uniform(5, 50, (10, 3))
    xyz = np.concatenate([sparse[:3], dense, sparse[3:]]).astype(np.float32)
    pts = np.column_stack([xyz, np.zeros(len(xyz), np.float32)]).astype(np.float32)
    f, valid = oracle.radius_pca_cloud(pts, 0.5, 13, True)
    assert not valid[:3].any() and valid[3:43].all() and not valid[43:].any()
    assert (f["ptID"] == np.arange(len(pts))).all() and (f["ptNum"][43:] == 1).all()
    assert not f["la"][:3].any()  # leading records: zero-filled
    for i in range(43, len(pts)):
        assert f["la"][i].tobytes() == f["la"][42].tobytes() and f["scatterStruc"][i] == f["scatterStruc"][42]
    f12, v12 = oracle.radius_pca_cloud(pts[3:15], 10.0, 13, False)
    f13, v13 = oracle.radius_pca_cloud(pts[3:16], 10.0, 13, False)
    assert not v12.any() and v13.all()  # exactly 12 vs 13 neighbours
    fa, _ = oracle.radius_pca_cloud(pts, 0.5, 13, True, nthreads=1)
    fb, _ = oracle.radius_pca_cloud(pts, 0.5, 13, True, nthreads=4)
    assert fa.tobytes() == fb.tobytes()  # the parallel baseline equals the faithful sequential loop


def test_pcl_eigen33_smallest_pair(oracle):
    rng = np.random.default_rng(1)
    for _ in range(100):
        A = rng.normal(size=(3, 3))
        S = (A @ A.T + 0.1 * np.eye(3)).astype(np.float32)
        ev, v = oracle.pcl_eigen33(S)
        w, V = np.linalg.eigh(S.astype(np.float64))
        assert abs(ev - w[0]) < 2e-5 * w[2] and abs(abs(v @ V[:, 0]) - 1) < 1e-3


def test_ring_projection_known_cells(oracle):
    """Beam −15° lands in row 0, +15° in row 15; azimuth along +y is column 900 - round(-90/0.2)... =
    the reference formula col = −round((h−90)/0.2) + 900 with h = atan2(x, y) in degrees."""
    d = 10.0
    pts = []
    for v in (-15, 1, 15):
        for az in (0.0, 90.0, 180.0, 270.0):
            a, e = np.deg2rad(az), np.deg2rad(v)
            pts.append([d * np.cos(e) * np.sin(a), d * np.cos(e) * np.cos(a), d * np.sin(e), 0])
    pts = np.array(pts, np.float32)
    rng, pti, tag, lr = oracle.ring_project(pts)
    rows = {-15: 0, 1: 8, 15: 15}
    cols = {0.0: 1350, 90.0: 900, 180.0: 450, 270.0: 0}
    i = 0
    for v in (-15, 1, 15):
        for az in (0.0, 90.0, 180.0, 270.0):
            assert pti[rows[v], cols[az]] == i, (v, az)
            assert abs(rng[rows[v], cols[az]] - d) < 1e-4
            i += 1
    assert (pti >= 0).sum() == 12 and lr == 0
    assert rng[3, 3] == np.finfo(np.float32).max


def _ring_image(pixels, n_scan=16, horizon=1800):
    """pixels: {(row, col): range}.  Builds the projection outputs directly: one point per filled pixel
    (coordinates only matter for the ground test; rows >= 8 never take part in it)."""
    rng = np.full((n_scan, horizon), np.finfo(np.float32).max, np.float32)
    pti = np.full((n_scan, horizon), -1, np.int32)
    pts = np.zeros((len(pixels), 4), np.float32)
    for n, ((r, c), d) in enumerate(sorted(pixels.items())):
        rng[r, c] = d
        pti[r, c] = n
        pts[n] = (d, 0.0, 0.0, 0.0)
    return pts, rng, pti


def test_ring_segment_quirks_by_hand(oracle):
    """Hand-derived from src/ringProjection.cpp:199-289 with the unsigned neighbour table
    (include/ringProjection.h:69,119-122): the flood grows right, +255 columns and down, never left or up."""
    d = np.float32(10.0)
    # (a) a 6-pixel run on ring 9 plus the pixel below its first pixel: 7 pixels, pushed rows {9, 10} -> only
    #     two rings counted (the seed's ring 9 counts because OTHER ring-9 pixels were pushed) -> rejected
    px = {(9, c): d for c in range(100, 106)}
    px[(10, 100)] = d
    # (b) the same shape over three rings -> valid (>= 5 points, 3 rings)
    for c in range(300, 304):
        px[(9, c)] = d
    px[(10, 300)] = d
    px[(11, 300)] = d
    # (c) far link: (12, 500) reaches (12, 755) although nothing lies between them; 30 more pixels to the
    #     right of 755 make the segment valid by size
    px[(12, 500)] = d
    for c in range(755, 785):
        px[(12, c)] = d
    # (d) wrap: (13, 1799) links right to (13, 0), and (13, 1600) links far to (13, 0) — but (13, 0) was
    #     flooded first as its own seed, so all three stay separate single pixels -> rejected
    px[(13, 0)] = d
    px[(13, 1799)] = d
    px[(13, 1600)] = d
    # (e) no upward growth: (15, 900) cannot reach (14, 900); raster order visits (14, 900) first and IT grows down
    px[(14, 900)] = d
    px[(15, 900)] = d
    pts, rng, pti = _ring_image(px)
    g, lab, cm, above, nl = oracle.ring_segment(pts, rng, pti)
    assert (g == 0).all() or set(np.unique(g)) <= {-1, 0}  # rows 0..7 are empty: no ground anywhere
    assert (lab[9, 100:106] == 999999).all() and lab[10, 100] == 999999
    assert lab[9, 300] == 1 and (lab[9, 300:304] == 1).all() and lab[10, 300] == 1 and lab[11, 300] == 1
    assert lab[12, 500] == 2 and (lab[12, 755:785] == 2).all()
    assert lab[13, 0] == 999999 and lab[13, 1799] == 999999 and lab[13, 1600] == 999999
    assert lab[14, 900] == 999999 and lab[15, 900] == 999999  # one 2-pixel segment, rejected
    assert nl == 2
    assert (lab[rng == np.finfo(np.float32).max] == -1).all()
    assert above.shape[0] == 0  # no ground -> maxGroundrangeOfcol stays 0 -> nothing is "above ground"


def test_ring_segment_ground_pairs_by_hand(oracle):
    """groundRemoval (src/ringProjection.cpp:88-114): flat pairs are ground; a missing upper pixel writes -1 over
    the 1 the previous pair gave; the column's farthest ground range gates cloudaboveGround (:186-189)."""
    n_scan, horizon = 16, 1800
    rng = np.full((n_scan, horizon), np.finfo(np.float32).max, np.float32)
    pti = np.full((n_scan, horizon), -1, np.int32)
    pts = []

    def put(r, c, x, z):
        pti[r, c] = len(pts)
        rng[r, c] = np.float32(np.sqrt(np.float32(x * x + z * z)))
        pts.append((x, 0.0, z, 0.0))

    put(0, 7, 5.0, -1.8); put(1, 7, 7.0, -1.8); put(2, 7, 9.0, -1.8)      # flat: rows 0,1,2 ground; row 3 empty
    put(9, 7, 4.0, 1.0)                                                     # an object nearer than the ground: above
    put(9, 8, 4.0, 1.0)                                                     # no ground in column 8: not above
    put(0, 20, 5.0, -1.8); put(1, 20, 5.2, 0.5)                             # steep pair: not ground
    pts = np.asarray(pts, np.float32)
    g, lab, cm, above, nl = oracle.ring_segment(pts, rng, pti)
    assert g[0, 7] == 1 and g[1, 7] == 1
    assert g[2, 7] == -1  # pair (2,3): upper empty -> -1 written over the 1 from pair (1,2)
    assert g[3, 7] == -1 and g[6, 7] == -1 and g[7, 7] == 0
    assert g[0, 20] == 0 and g[1, 20] == -1
    assert lab[0, 7] == -1 and lab[1, 7] == -1 and lab[2, 7] == 999999  # row 2 is free again: a 1-pixel segment
    assert cm[7] == rng[1, 7] and cm[8] == 0.0
    assert above.tolist() == [int(pti[9, 7])]


# ---- lidarOdometry images (src/lidarOdometry.cpp:21-187) ----

def test_atanf_restatement_equals_libm(oracle):
    """The CUDA path evaluates atanf with the fdlibm algorithm (odom.cu); the oracle's copy of that algorithm
    equals this C library's atanf on every sampled bit pattern (all 2^32 were checked once, stride 1)."""
    assert oracle.load().orc_atanf_mismatches(0, 257) == 0
    assert oracle.load().orc_atanf_mismatches(0x3E000000, 31) == 0  # a second lattice from the reduction thresholds on


def test_odom_range_image_known_cells(oracle):
    """Four returns on known beams: rows from the vertical angle above the minimum in 2-degree steps, columns
    from atan(y / x) (NOT atan2: the back half folds onto the front half) above the minimum in 0.2-degree steps."""
    def pt(v_deg, h_deg, d):
        v, h = np.radians(v_deg), np.radians(h_deg)
        return [d * np.cos(v) * np.cos(h), d * np.cos(v) * np.sin(h), d * np.sin(v), 0.0]
    pts = np.array([pt(-15, -60, 10), pt(-15 + 4.5, -60 + 10.1, 12), pt(1, 0, 5), pt(-15 + 4.5, -60 + 10.1 + 180, 7)], np.float32)
    cnt = np.full(4, 6, np.int32)
    rng, pid, pti, stats, mm = oracle.odom_range_image(pts, cnt)
    assert pti[0, 0] == 0 and pti[8, 300] == 2
    assert pti[2, 50] == 3          # the point behind the sensor lands in the cell of point 1 and overwrites it
    assert stats.tolist() == [1, 3]
    assert abs(mm[0] + 15) < 1e-5 and abs(mm[2] + 60) < 1e-5
    assert rng[2, 50] == np.float32(7.0) or abs(rng[2, 50] - 7.0) < 1e-5
    cnt[3] = 5                      # fewer than 6 neighbours: left out, point 1 shows again
    rng, pid, pti, stats, mm = oracle.odom_range_image(pts, cnt)
    assert pti[2, 50] == 1 and stats.tolist() == [0, 3] and pid[2, 50] == 1.0


def test_odom_smoothness_known_answers(oracle):
    img = np.full((16, 900), 10.0, np.float32)
    sm, ty = oracle.odom_smoothness(img)
    assert (sm == 0).all()
    # every smoothness ties at 0: `>= sorted[size-3]` holds for all, `< sorted[3]` for none (:167-172);
    # the first nine columns of each 180-column block are never visited (:133)
    exp = np.ones((16, 900), np.float32)
    for b in range(5):
        exp[:, 180 * b:180 * b + 9] = 0
    assert np.array_equal(ty, exp)
    # one spike: the window of 18 pixels [col-9, col+9) that contains it has sum = 17*10 + 30
    img[4, 100] = 30.0
    sm, ty = oracle.odom_smoothness(img)
    assert sm[4, 100] == np.float32(abs((17 * 10.0 + 30.0) - 30.0 * 18) / (18 * 30.0))
    assert sm[4, 95] == np.float32(abs((17 * 10.0 + 30.0) - 10.0 * 18) / (18 * 10.0))
    # the window is asymmetric: pixel 100 is inside [col-9, col+9) for col = 92 .. 109
    assert sm[4, 92] == np.float32(20.0 / 180.0) and sm[4, 109] == np.float32(20.0 / 180.0) and sm[4, 110] == 0 and sm[4, 91] == 0
    assert ty[4, 100] == 1  # among the three roughest
    # fewer than ten valid pixels in a block: no types at all there
    img2 = np.zeros((16, 900), np.float32)
    img2[0, 20:29] = 5.0   # nine pixels, each sees exactly nine valid ones: nine smoothness values, not > 9 (:162)
    sm, ty = oracle.odom_smoothness(img2)
    assert (ty == 0).all() and (sm == 0).all()
    img2[0, 29] = 5.0      # ten values, all 0: `0 >= sorted[size-3] = 0` marks the whole block, empty pixels too
    sm, ty = oracle.odom_smoothness(img2)
    assert (ty[0, 9:180] == 1).all() and (ty[0, :9] == 0).all() and (ty[1:] == 0).all() and (ty[0, 180:] == 0).all()


def test_xyz2pc_known_answers(oracle):
    """preprocess::xyz2pc (src/smoothing.cpp:83-125) by hand: intensity = first byte of the 4th token, fields a line does
    not supply keep the previous line's values, the eof() loop repeats the last point after a final newline."""
    got = oracle.xyz2pc(b"1 2 3 a\n4.5 5 6 zz\n7 8\n")
    assert got.tolist() == [[1, 2, 3, 97], [4.5, 5, 6, 122], [7, 8, 6, 122], [7, 8, 6, 122]]
    assert oracle.xyz2pc(b"1 2 3 7").tolist() == [[1, 2, 3, 55]]  # '7' is the character, not the number
    assert oracle.xyz2pc(b"").tolist() == [[0, 0, 0, 0]]
    assert oracle.xyz2pc(b"x" * 300) is None  # getline(line, 256) sets failbit: the reference loops forever
    # the boundary: 255 characters are stored and the delimiter (or the end of the file) is tested before the count
    # ([istream.unformatted]: eof, delimiter, then n - 1 characters stored), so a 255-character line reads and 256 fail
    assert oracle.xyz2pc(b"5" * 255 + b"\n").shape == (2, 4) and oracle.xyz2pc(b"5" * 255).shape == (1, 4)
    assert oracle.xyz2pc(b"5" * 256 + b"\n") is None and oracle.xyz2pc(b"5" * 256) is None


def test_dbscan_radius_is_the_references_double(oracle):
    """clusterDBSCAN takes `double r` (include/dbscanCLuster.h:256; call site: 0.2, src/dynaObjTest.cpp:133) and PCL compares
    squared distances with (float)(r * r) evaluated in double = 0x3d23d70a; the float 0.2f squares to 0x3d23d70b.  Two points
    whose squared distance IS 0x3d23d70a are therefore no neighbours for the reference (strict <) — orc_dbscan_d — and
    neighbours when the radius is narrowed to float first."""
    dx, dy = np.float32(0.19999999), np.float32(6.1e-5)
    d2 = np.float32(np.float32(dx * dx) + np.float32(dy * dy))
    assert d2.view(np.uint32) == 0x3D23D70A == np.float32(0.2 * 0.2).view(np.uint32)
    assert np.float32(np.float64(np.float32(0.2)) ** 2).view(np.uint32) == 0x3D23D70B
    pts = np.array([[0, 0, 0, 0], [dx, dy, 0, 0]], np.float32)
    assert oracle.dbscan_d(pts, 0.2, 2)[1] == 0       # the reference: each point alone, below minpts
    assert oracle.dbscan(pts, 0.2, 2)[1] == 1         # radius narrowed to float: one cluster
    st, nc = oracle.dbscan_d(pts, float(np.float32(0.2)), 2)  # the double that equals 0.2f behaves like the float path
    assert nc == 1 and st.tolist() == [1, 1]
