"""GPU parity: RingProjector::projectPointCloud (src/ringProjection.cpp:24-81) — C-ABI."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import capi, synth

pytestmark = pytest.mark.gpu


def test_ring_single_frame(pcp, oracle):
    fr = synth.vlp16_frame(2)
    pcp.set_cloud(fr)
    rng, pti, lr = pcp.ring_project()
    rrng, rpti, _, rlr = oracle.ring_project(fr)
    np.testing.assert_array_equal(pti[0], rpti)
    np.testing.assert_array_equal(rng[0].view(np.uint32), rrng.view(np.uint32))
    assert int(lr[0]) == rlr
    assert (pti >= 0).sum() > 20000


def test_ring_last_writer_wins_and_rejects(pcp, oracle):
    """Duplicates in one cell: the highest input index wins (Q11); out-of-fov points are dropped."""
    fr = synth.vlp16_frame(3)[:4000]
    dup = np.concatenate([fr, fr[::-1] * np.float32(1.0)], 0)
    dup[-5:, 2] = 50.0  # far above the 15° beam -> row >= 16 rejected
    dup[10:14, :3] *= np.float32(40.0)  # long range (> 60 m)
    pcp.set_cloud(np.ascontiguousarray(dup))
    rng, pti, lr = pcp.ring_project()
    rrng, rpti, _, rlr = oracle.ring_project(dup)
    np.testing.assert_array_equal(pti[0], rpti)
    np.testing.assert_array_equal(rng[0].view(np.uint32), rrng.view(np.uint32))
    assert int(lr[0]) == rlr


def test_ring_batch(pcp, oracle):
    pts, offs = synth.vlp16_batch(4, seed0=50)
    pcp.set_cloud(pts)
    rng, pti, lr = pcp.ring_project(frame_offsets=offs)
    for f in range(4):
        rrng, rpti, _, rlr = oracle.ring_project(pts[offs[f]:offs[f + 1]])
        np.testing.assert_array_equal(pti[f], rpti)
        np.testing.assert_array_equal(rng[f].view(np.uint32), rrng.view(np.uint32))
        assert int(lr[f]) == rlr


def _check_segmentation(got, f, fr, oracle, prm=None):
    rrng, rpti, _, _ = oracle.ring_project(fr)
    kw = {} if prm is None else dict(ground_scan_ind=prm.ground_scan_ind, sensor_mount_angle=prm.sensor_mount_angle,
                                     segment_theta=prm.segment_theta, valid_point_num=prm.segment_valid_point_num,
                                     valid_line_num=prm.segment_valid_line_num)
    g, lab, cm, above, nl = oracle.ring_segment(fr, rrng, rpti, **kw)
    np.testing.assert_array_equal(got["index"][f], rpti)
    np.testing.assert_array_equal(got["ground"][f], g)
    np.testing.assert_array_equal(got["label"][f], lab)
    np.testing.assert_array_equal(got["col_max_ground"][f].view(np.uint32), cm.view(np.uint32))
    assert int(got["n_labels"][f]) == nl
    na = int(got["above_count"][f])
    assert na == above.shape[0]
    np.testing.assert_array_equal(got["above_idx"][f][:na], above)
    assert (got["above_idx"][f][na:] == -1).all()
    return lab, nl


def test_ring_full_segmentation_matches_reference_flood(pcp, oracle):
    """groundRemoval + cloudSegmentation (src/ringProjection.cpp:83-289): ground image, label image (valid
    segments numbered in seed raster order, 999999 for rejected ones), per-column ground range and the
    cloud above ground equal the literal sequential restatement, for a batch of frames in one launch."""
    pts, offs = synth.vlp16_batch(5, seed0=70)
    pcp.set_cloud(pts)
    got = pcp.ring_segment(frame_offsets=offs)
    seen_rejected = seen_valid = 0
    for f in range(5):
        lab, nl = _check_segmentation(got, f, pts[offs[f]:offs[f + 1]], oracle)
        seen_rejected += int((lab == 999999).sum())
        seen_valid += nl
    assert seen_rejected > 0 and seen_valid > 10


@pytest.mark.parametrize("keep", [0.6, 0.25, 0.05])
def test_ring_segmentation_sparse_and_ragged(pcp, oracle, keep):
    """Thinned frames break the image into many small components: the size / ring-count validity rule,
    the far (+255 column) links and the wrap to column 0 all decide labels here."""
    rng = np.random.default_rng(int(keep * 100))
    fr = synth.vlp16_frame(91, pose_jitter=0.5)
    fr = np.ascontiguousarray(fr[rng.random(fr.shape[0]) < keep])
    fr[:, :3] *= (1.0 + 0.02 * rng.standard_normal((fr.shape[0], 1))).astype(np.float32)  # ragged ranges
    pcp.set_cloud(fr)
    got = pcp.ring_segment()
    _check_segmentation(got, 0, fr, oracle)
    prm = capi.ring_seg_params_default()
    prm.ground_scan_ind, prm.segment_theta, prm.segment_valid_point_num, prm.segment_valid_line_num = 5, 0.5, 3, 2
    prm.sensor_mount_angle = 2.0
    got = pcp.ring_segment(prm=prm)
    _check_segmentation(got, 0, fr, oracle, prm)


def test_ring_segmentation_empty_and_tiny(pcp, oracle):
    fr = synth.vlp16_frame(92)[:3].copy()
    pcp.set_cloud(fr)
    got = pcp.ring_segment()
    _check_segmentation(got, 0, fr, oracle)
    assert int(got["n_labels"][0]) == 0
