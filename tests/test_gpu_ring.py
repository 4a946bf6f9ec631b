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
