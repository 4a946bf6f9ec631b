"""GPU: the C++ drop-in class `pcaAnalysist` (host/pcaAnalysis.cpp over libpcp.so) driven like the
reference's callers (main.cpp:115-117, src/buildingFacadeExtraction.cpp:212-219)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from pointcloudprocessing_b200 import synth
from parity import check_features

pytestmark = pytest.mark.gpu
PKG = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pointcloudprocessing_b200")


def test_class_dropin(tmp_path, oracle):
    exe = os.path.join(PKG, "dropin_test")
    assert os.path.exists(exe), "build the host library first (__graft_entry__.build())"
    pts = synth.urban(41, 20000, extent=25.0)
    pts.tofile(tmp_path / "cloud.bin")
    # voxel segments with >= 20 points, centre = voxel centre (constructVoxels / makeSuperVoxels)
    vox = np.floor(pts[:, :3] / 3.0).astype(np.int64)
    key = (vox[:, 0] * 1000 + vox[:, 1]) * 1000 + vox[:, 2]
    order = np.argsort(key, kind="stable")
    uk, start, counts = np.unique(key[order], return_index=True, return_counts=True)
    keep = np.nonzero(counts >= 20)[0][:60]
    centres = (np.stack([uk // 1000000, (uk // 1000) % 1000, uk % 1000], 1)[keep] + 0.5).astype(np.float32) * 3.0
    segs = [order[start[j]:start[j] + counts[j]].astype(np.int32) for j in keep]
    with open(tmp_path / "seg.bin", "wb") as f:
        f.write(struct.pack("<q", len(segs)))
        for c, s in zip(centres, segs):
            f.write(c.tobytes() + struct.pack("<q", len(s)) + s.tobytes())
    r = subprocess.run([exe, str(tmp_path / "cloud.bin"), "0.5", str(tmp_path / "out.bin"), str(tmp_path / "seg.bin"),
                        str(tmp_path / "sout.bin")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    got = np.fromfile(tmp_path / "out.bin", dtype=oracle.FEATURE_DTYPE)
    ref, valid = oracle.radius_pca_cloud(pts, 0.5, 13, True, nthreads=1)
    check_features(got, ref, valid, "dropin cloud")
    first = int(np.argmax(valid))
    stale = (~valid.astype(bool)) & (np.arange(valid.size) > first)
    check_features(got[stale], ref[stale], None, "dropin stale")
    sgot = np.fromfile(tmp_path / "sout.bin", dtype=oracle.FEATURE_DTYPE)
    offs = np.zeros(len(segs) + 1, np.int64)
    np.cumsum([len(s) for s in segs], out=offs[1:])
    sref = oracle.segment_pca(pts, centres, offs, np.concatenate(segs))
    assert (sgot["ptID"] == -7).all()  # the centro overload leaves ptID alone (src/pcaAnalysis.cpp:92)
    sref["ptID"] = -7
    check_features(sgot, sref, None, "dropin segments")
