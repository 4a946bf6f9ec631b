"""GPU: frame batches (configs[1]/[4]) through the multi-context pipeline — every frame's ring images and
k = 10 PCA records must equal the oracle's, whatever the number of contexts in flight."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import frames, synth
from parity import check_features

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_ctx", [1, 3])
def test_frame_pipeline_matches_oracle(oracle, n_ctx):
    batch = [synth.vlp16_frame(200 + f, pose_jitter=0.5) for f in range(7)]
    with frames.FramePipeline(0, n_ctx=n_ctx) as pipe:
        got = pipe.run(batch, k=10)
        again = pipe.run(batch[:3], k=10)  # buffers are reused across runs
    for f, fr in enumerate(batch):
        rrng, rpti, _, rlong = oracle.ring_project(fr)
        assert np.array_equal(got[f]["index"], rpti)
        assert np.array_equal(got[f]["range"].view(np.uint32), rrng.view(np.uint32))
        assert got[f]["long_range"] == rlong
        check_features(got[f]["features"], oracle.knn_pca_cloud(fr, 10, 0), None, f"frame {f}")
    for f in range(3):
        assert again[f]["features"].tobytes() == got[f]["features"].tobytes()


@pytest.mark.parametrize("n_ctx,batch", [(1, 4), (2, 3), (1, 16)])
def test_frame_batcher_equals_per_frame_runs(oracle, n_ctx, batch):
    """One segmented grid over many frames gives, per frame, the bytes of a run on that frame alone."""
    fr = [synth.vlp16_frame(400 + f, pose_jitter=0.5) for f in range(9)]
    with frames.FramePipeline(0, n_ctx=1) as pipe:
        want = pipe.run(fr, k=10, cell=0.7)
    with frames.FrameBatcher(0, n_ctx=n_ctx, batch=batch) as bat:
        got = bat.run(fr, k=10, cell=0.7)
        auto = bat.run(fr, k=10)  # auto cell differs from the per-frame one; the records must not
    for f in range(len(fr)):
        for key in ("features", "range", "index"):
            assert got[f][key].tobytes() == want[f][key].tobytes(), (f, key)
        assert got[f]["long_range"] == want[f]["long_range"]
        assert auto[f]["features"].tobytes() == want[f]["features"].tobytes(), f
    check_features(got[2]["features"], oracle.knn_pca_cloud(fr[2], 10, 0), None, "batched frame 2")


def test_segmented_grid_edge_cases(oracle):
    """Empty frames, frames smaller than k, overlapping frames that must not see each other, and every
    k-NN kernel path (tile, work list, v3) on a segmented cloud."""
    import os
    from pointcloudprocessing_b200 import capi
    rng = np.random.default_rng(5)
    base = synth.urban(9, 6000, extent=20.0)
    parts = [base[:2500], base[:0], base[2500:2507], base[2000:6000], base[:0], base[100:101], base[:2500].copy()]
    offs = np.concatenate([[0], np.cumsum([p.shape[0] for p in parts])]).astype(np.int64)
    cloud = np.concatenate(parts)
    k = 12
    want = []
    with capi.Pcp(0) as p:
        for part in parts:
            if part.shape[0] == 0:
                want.append(None)
                continue
            p.set_cloud(part)
            p.build_grid(0.0, k)
            want.append((p.knn_pca(k), p.knn_neighbors(k)))
    # PCP_KNN_IMPL is read when a context is created: "" = automatic (round-1 tile kernel for frame batches), tile6 = round-2 tile kernel, tile4 = round-1 tile kernel,
    # v3 = warp per query for every point
    for impl in ("", "tile6", "tile4", "v3"):
        os.environ["PCP_KNN_IMPL"] = impl
        try:
            with capi.Pcp(0) as q:
                for cell in (0.0, 0.9, 40.0):
                    q.set_cloud(cloud)
                    q.set_frames(offs)
                    q.build_grid(cell, k)
                    feat = q.knn_pca(k)
                    idx, d2 = q.knn_neighbors(k)
                    for f, part in enumerate(parts):
                        if want[f] is None:
                            continue
                        a, b = int(offs[f]), int(offs[f + 1])
                        wf, (wi, wd) = want[f]
                        gf = feat[a:b].copy()
                        gf["ptID"] -= np.int32(a)
                        assert gf.tobytes() == wf.tobytes(), (impl, cell, f)
                        gi = idx[a:b].copy()
                        gi[gi >= 0] -= a
                        assert np.array_equal(gi, wi), (impl, cell, f)
                        assert np.array_equal(d2[a:b].view(np.uint32), wd.view(np.uint32)), (impl, cell, f)
        finally:
            os.environ.pop("PCP_KNN_IMPL", None)
    with capi.Pcp(0) as p:
        # radius mode: the stale-record carry restarts at every frame
        r = 0.6
        p.set_cloud(cloud)
        p.set_frames(offs)
        p.build_grid(r * 1.001, 0)
        feat = p.radius_pca(r)
        for f, part in enumerate(parts):
            if part.shape[0] == 0:
                continue
            a, b = int(offs[f]), int(offs[f + 1])
            p.set_cloud(part)
            p.build_grid(r * 1.001, 0)
            wf = p.radius_pca(r)
            gf = feat[a:b].copy()
            gf["ptID"] -= np.int32(a)
            assert gf.tobytes() == wf.tobytes(), ("radius", f)
        # a new cloud resets the segmentation
        p.set_cloud(parts[0])
        p.build_grid(0.0, k)
        assert p.knn_pca(k).tobytes() == want[0][0].tobytes()
    check_features(want[3][0], oracle.knn_pca_cloud(parts[3], k, 0), None, "part 3")


def test_frame_batcher_full_segmentation(oracle):
    """configs[4] end to end per frame: projection + ground removal + segmentation + k-NN PCA in one batch."""
    fr = [synth.vlp16_frame(500 + f, pose_jitter=0.5) for f in range(5)]
    with frames.FrameBatcher(0, n_ctx=2, batch=3, segment=True) as bat:
        got = bat.run(fr, k=10)
    for f, cloud in enumerate(fr):
        rrng, rpti, _, _ = oracle.ring_project(cloud)
        g, lab, cm, above, nl = oracle.ring_segment(cloud, rrng, rpti)
        assert np.array_equal(got[f]["ground"], g) and np.array_equal(got[f]["label"], lab)
        assert np.array_equal(got[f]["above_idx"], above) and got[f]["n_labels"] == nl
    check_features(got[4]["features"], oracle.knn_pca_cloud(fr[4], 10, 0), None, "frame 4")


def test_frame_batcher_odometry_images(oracle):
    """FrameBatcher(odom=True): per frame the lidarOdometry range image and point types of a run on that frame alone."""
    from pointcloudprocessing_b200 import frames as F
    fr = [synth.vlp16_frame(400 + f, pose_jitter=0.5) for f in range(5)]
    with F.FrameBatcher(0, n_ctx=2, batch=3, odom=True) as bat:
        res = bat.run(fr, k=10)
    for f, r in zip(fr, res):
        rng = oracle.odom_range_image(f, oracle.KdTree(f).radius_counts(3.0))[0]
        np.testing.assert_array_equal(r["odom_range"].view(np.uint32), rng.view(np.uint32))
        np.testing.assert_array_equal(r["odom_type"], oracle.odom_smoothness(rng)[1])
