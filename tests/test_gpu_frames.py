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
