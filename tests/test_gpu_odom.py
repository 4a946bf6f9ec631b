"""GPU parity: lidarOdometryClass::getRangeAndptIdImage / calculateSmoothness (src/lidarOdometry.cpp:21-187)
and the capped radius count behind them — through the C-ABI, bit-exact against the oracle."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _check_frame(got, f, fr, oracle, counts=None):
    cnt = oracle.KdTree(fr).radius_counts(3.0) if counts is None else counts
    rng, pid, pti, stats, mm = oracle.odom_range_image(fr, cnt)
    np.testing.assert_array_equal(got["pt_index"][f], pti)
    np.testing.assert_array_equal(got["range"][f].view(np.uint32), rng.view(np.uint32))
    np.testing.assert_array_equal(got["ptid"][f].view(np.uint32), pid.view(np.uint32))
    np.testing.assert_array_equal(got["stats"][f], stats)
    np.testing.assert_array_equal(got["minmax"][f], mm)
    return rng


def test_radius_count_exact_and_capped(pcp, oracle):
    fr = synth.vlp16_frame(2)[::3].copy()
    ref = oracle.KdTree(fr).radius_counts(3.0)
    pcp.set_cloud(fr)
    for cell in (3.0, 1.1, 0.0):  # cell = radius, a finer grid, the automatic k-NN grid
        pcp.build_grid(cell, 10 if cell == 0.0 else 0)
        np.testing.assert_array_equal(pcp.radius_count(3.0, 0), ref)
        np.testing.assert_array_equal(pcp.radius_count(3.0, 6), np.minimum(ref, 6))
    urban = synth.urban(5, 30000, extent=40.0)
    ref = oracle.KdTree(urban).radius_counts(0.7)
    pcp.set_cloud(urban)
    pcp.build_grid(0.7, 0)
    np.testing.assert_array_equal(pcp.radius_count(0.7, 0), ref)
    np.testing.assert_array_equal(pcp.radius_count(0.7, 13), np.minimum(ref, 13))


def test_odom_single_frame(pcp, oracle):
    fr = synth.vlp16_frame(2)
    pcp.set_cloud(fr)
    pcp.build_grid(3.0, 0)
    got = pcp.odom_range_image()
    rng = _check_frame(got, 0, fr, oracle)
    assert (rng > 0).sum() > 5000
    sm, ty = pcp.odom_smoothness(n_frames=1)  # the images still on the device
    rsm, rty = oracle.odom_smoothness(rng)
    np.testing.assert_array_equal(sm[0].view(np.uint32), rsm.view(np.uint32))
    np.testing.assert_array_equal(ty[0], rty)
    assert (rty == 1).sum() > 50 and (rty == 2).sum() > 50
    sm2, ty2 = pcp.odom_smoothness(rng)  # the same from a host image
    np.testing.assert_array_equal(sm2, sm)
    np.testing.assert_array_equal(ty2, ty)


def test_odom_sparse_gate_and_rejects(pcp, oracle):
    """Returns with fewer than six neighbours inside 3 m are left out (:76-78); NaN angles and cells outside
    the 16 x 900 matrix are skipped; duplicates: the last writer wins."""
    fr = synth.vlp16_frame(7)
    far = fr[::50].copy()
    far[:, :3] *= np.float32(9.0)  # spread out: most of these lose their neighbours
    pts = np.concatenate([fr[::2], far, fr[:200]], 0)
    pts[5, :2] = 0.0              # x = y = 0: horizontal angle NaN
    pts[6, :3] = 0.0              # the origin
    pts[7, 0] = 0.0               # x = 0: atan(+-inf) = +-90 degrees
    pts[8, 2] = 80.0              # steep: widens the vertical span past 16 rows for the rest
    pts = np.ascontiguousarray(pts)
    cnt = oracle.KdTree(pts).radius_counts(3.0)
    assert (cnt < 6).sum() > 20
    pcp.set_cloud(pts)
    pcp.build_grid(3.0, 0)
    got = pcp.odom_range_image()
    _check_frame(got, 0, pts, oracle, cnt)


def test_odom_frame_batch(pcp, oracle):
    pts, offs = synth.vlp16_batch(5, seed0=70)
    # an empty frame and a tiny one in the middle
    offs = np.array([offs[0], offs[1], offs[1], offs[2], offs[2] + 40, offs[3], offs[5]], np.int64)
    pcp.set_cloud(pts)
    pcp.set_frames(offs)
    pcp.build_grid(3.0, 0)
    got = pcp.odom_range_image(frame_offsets=offs)
    nf = offs.shape[0] - 1
    sm, ty = pcp.odom_smoothness(n_frames=nf)
    for f in range(nf):
        fr = pts[offs[f]:offs[f + 1]]
        if fr.shape[0] == 0:
            assert (got["range"][f] == 0).all() and (got["pt_index"][f] == -1).all() and (got["stats"][f] == 0).all()
            assert (ty[f] == 0).all()
            continue
        rng = _check_frame(got, f, fr, oracle)
        rsm, rty = oracle.odom_smoothness(rng)
        np.testing.assert_array_equal(sm[f].view(np.uint32), rsm.view(np.uint32))
        np.testing.assert_array_equal(ty[f], rty)
    pcp.set_frames(np.array([0, pts.shape[0]], np.int64))


def test_smoothness_synthetic_images(pcp, oracle):
    """Ties, constant rows, near-range pixels, blocks with too few valid pixels, non-default parameters."""
    rs = np.random.default_rng(11)
    imgs = np.zeros((6, 16, 900), np.float32)
    imgs[0] = 10.0                                             # constant: every smoothness is 0 (all ties)
    imgs[1] = rs.uniform(0.0, 30.0, (16, 900))                 # noise, some below the 1 m near range
    imgs[2] = np.where(rs.random((16, 900)) < 0.6, 0.0, rs.uniform(2, 50, (16, 900)))  # sparse: few valid windows
    imgs[3] = np.round(rs.uniform(1.0, 4.0, (16, 900)))        # many exact ties, values exactly 1.0
    imgs[4, :, ::2] = 5.0                                      # every second pixel: exactly 9 valid per window
    imgs[5] = 20.0 + np.sin(np.arange(900) / 7.0)[None, :] * rs.uniform(0, 3, (16, 1))
    imgs[5, 3, 400:440] = 0.5
    sm, ty = pcp.odom_smoothness(imgs)
    for f in range(imgs.shape[0]):
        rsm, rty = oracle.odom_smoothness(imgs[f])
        np.testing.assert_array_equal(sm[f].view(np.uint32), rsm.view(np.uint32), err_msg=f"image {f}")
        np.testing.assert_array_equal(ty[f], rty, err_msg=f"image {f}")
    prm = capi.odom_params_default()
    prm.n_rows, prm.n_cols, prm.n_blocks, prm.half_window, prm.min_valid, prm.pick, prm.min_block_valid = 8, 1024, 4, 5, 4, 2, 6
    img = rs.uniform(0.0, 30.0, (3, 8, 1024)).astype(np.float32)
    sm, ty = pcp.odom_smoothness(img, prm)
    for f in range(3):
        rsm, rty = oracle.odom_smoothness(img[f], n_blocks=4, half_window=5, min_valid=4, pick=2, min_block_valid=6)
        np.testing.assert_array_equal(sm[f].view(np.uint32), rsm.view(np.uint32))
        np.testing.assert_array_equal(ty[f], rty)
