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


def _run(args, tmp_path):
    exe = os.path.join(PKG, "consumers_test")
    assert os.path.exists(exe), "build the host library first (__graft_entry__.build())"
    r = subprocess.run([exe] + [str(a) for a in args], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr


def test_consumers_dropin_odometry_images(tmp_path, oracle):
    """lidarOdometryClass::getRangeAndptIdImage + calculateSmoothness through the C++ signatures
    (include/lidarOdometry.h:31-33), Eigen-style column-major matrices on the caller's side."""
    fr = synth.vlp16_frame(12)
    fr.tofile(tmp_path / "cloud.bin")
    _run(["odom", tmp_path / "cloud.bin", tmp_path / "r.bin", tmp_path / "p.bin", tmp_path / "t.bin"], tmp_path)
    cnt = oracle.KdTree(fr).radius_counts(3.0)
    rng, pid, _, _, _ = oracle.odom_range_image(fr, cnt)
    _, ty = oracle.odom_smoothness(rng)
    np.testing.assert_array_equal(np.fromfile(tmp_path / "r.bin", np.float32).reshape(16, 900).view(np.uint32), rng.view(np.uint32))
    np.testing.assert_array_equal(np.fromfile(tmp_path / "p.bin", np.float32).reshape(16, 900), pid)
    np.testing.assert_array_equal(np.fromfile(tmp_path / "t.bin", np.float32).reshape(16, 900), ty)


def test_consumers_dropin_dbscan_and_mean_shift(tmp_path, oracle):
    """DynaTools::clusterDBSCAN (cluster count + the centroid / min / max lists that pass its size filter,
    include/dbscanCLuster.h:308-325) and buildingFacadeExtractor::meanShiftClustering through the C++ signatures."""
    rs = np.random.default_rng(5)
    blobs = [rs.normal(c, 0.25, (400, 3)) for c in ((2, 2, 1), (8, 3, 1), (5, 9, 1), (12, 12, 1), (15, 4, 1))]
    wall = np.stack([rs.uniform(0, 18, 3000), np.full(3000, 16.0), rs.uniform(0, 3, 3000)], 1)  # 18 m long: filtered out
    xyz = np.concatenate(blobs + [wall, rs.uniform(0, 18, (300, 3))], 0)
    pts = np.ascontiguousarray(np.concatenate([xyz, np.zeros((xyz.shape[0], 1))], 1).astype(np.float32))
    pts = np.ascontiguousarray(pts[rs.permutation(pts.shape[0])])
    pts.tofile(tmp_path / "cloud.bin")
    _run(["dbscan", tmp_path / "cloud.bin", 0.4, 10, tmp_path / "d.bin"], tmp_path)
    raw = open(tmp_path / "d.bin", "rb").read()
    cluster_id, kept = struct.unpack("<ii", raw[:8])
    got = np.frombuffer(raw[8:], np.float32).reshape(kept, 9)
    flat = pts.copy()
    flat[:, 2] = 0
    state, nc = oracle.dbscan(flat, 0.4, 10)
    assert cluster_id == nc and nc > 3
    cen, lo, hi, _ = oracle.cluster_boxes(pts, state, nc)
    ext = hi - lo
    keep = ~((ext[:, 0] > 4) | (ext[:, 1] > 4) | (ext[:, 2].astype(np.float64) < 0.3))
    assert kept == int(keep.sum()) and 0 < kept < nc
    np.testing.assert_array_equal(got[:, 3:6], lo[keep])
    np.testing.assert_array_equal(got[:, 6:9], hi[keep])
    assert np.abs(got[:, :3] - cen[keep]).max() < 1e-4
    _run(["meanshift", tmp_path / "cloud.bin", 2, tmp_path / "m.bin"], tmp_path)
    moved = np.fromfile(tmp_path / "m.bin", np.float32).reshape(-1, 3)
    cur = pts
    for _ in range(2):
        cur, _, _ = oracle.mean_shift_step(pts, cur, 2.0, 15)
    err = np.abs(moved - cur[:, :3]).max(1)
    assert (err > 2e-4).sum() <= 5


def test_consumers_dropin_txt2pc(tmp_path, oracle):
    """preprocess::txt2pc(path, cloud) through the C++ signature: appends to the caller's cloud."""
    pts = synth.urban(29, 5000, extent=20.0)
    text = "".join("%.4f %.4f %.4f %.0f\n" % tuple(p) for p in pts.astype(np.float64)).encode() + b"1 2\n\n3"
    (tmp_path / "cloud.txt").write_bytes(text)
    _run(["txt", tmp_path / "cloud.txt", tmp_path / "o.bin"], tmp_path)
    got = np.fromfile(tmp_path / "o.bin", np.float32).reshape(-1, 4)
    ref = oracle.txt2pc(text)
    np.testing.assert_array_equal(got[0], np.array([-1, -2, -3, -4], np.float32))
    np.testing.assert_array_equal(got[1:].view(np.uint32), ref.view(np.uint32))


def test_consumers_dropin_xyz2pc(tmp_path, oracle):
    """preprocess::xyz2pc(path, cloud) through the C++ signature (src/smoothing.cpp:83-125): appends to the caller's cloud."""
    pts = synth.urban(31, 5000, extent=20.0)
    text = "".join("%.4f %.4f %.4f %s\n" % (p[0], p[1], p[2], "abcxyz"[i % 6]) for i, p in enumerate(pts.astype(np.float64))).encode() + b"1 2\n\n3\n"
    (tmp_path / "cloud.xyz").write_bytes(text)
    _run(["xyz", tmp_path / "cloud.xyz", tmp_path / "o.bin"], tmp_path)
    got = np.fromfile(tmp_path / "o.bin", np.float32).reshape(-1, 4)
    ref = oracle.xyz2pc(text)
    np.testing.assert_array_equal(got[0], np.array([-1, -2, -3, -4], np.float32))
    np.testing.assert_array_equal(got[1:].view(np.uint32), ref.view(np.uint32))


def test_consumers_dropin_make_super_voxels(tmp_path, oracle):
    """buildingFacadeExtractor::makeSuperVoxels (src/buildingFacadeExtraction.cpp:207-223) as ONE batched segment-PCA call:
    every voxel's record equals the oracle's calculatePCAofPoint(cloud, centro, idx) (:89-121) and, bitwise, the record
    the per-voxel class method gives; planeStruc > 0.6 selects nothing, like upstream."""
    from parity import check_features
    from pointcloudprocessing_b200 import capi

    pts = synth.urban(37, 60000, extent=40.0)
    pts.tofile(tmp_path / "cloud.bin")
    _run(["supervoxels", tmp_path / "cloud.bin", 3.0, tmp_path / "o.bin"], tmp_path)
    raw = (tmp_path / "o.bin").read_bytes()
    nv, ns = np.frombuffer(raw[:8], np.int32)
    assert nv > 150 and ns == 0
    pos, batched, single, centres, offs, idx = 8, [], [], [], [0], []
    for v in range(nv):
        batched.append(np.frombuffer(raw[pos:pos + 68], capi.FEATURE_DTYPE)[0]); pos += 68
        single.append(np.frombuffer(raw[pos:pos + 68], capi.FEATURE_DTYPE)[0]); pos += 68
        centres.append(np.frombuffer(raw[pos:pos + 12], np.float32)); pos += 12
        m = int(np.frombuffer(raw[pos:pos + 4], np.int32)[0]); pos += 4
        idx.append(np.frombuffer(raw[pos:pos + 4 * m], np.int32)); pos += 4 * m
        offs.append(offs[-1] + m)
    assert pos == len(raw)
    batched = np.array(batched, capi.FEATURE_DTYPE)
    single = np.array(single, capi.FEATURE_DTYPE)
    ref = oracle.segment_pca(pts, np.array(centres), np.array(offs, np.int64), np.concatenate(idx))
    ref["ptID"] = batched["ptID"]  # not set by this overload (src/pcaAnalysis.cpp:92)
    check_features(batched, ref, None, "makeSuperVoxels")
    for name in ("la", "v1", "v2", "v3", "ptNum", "lineStruc", "planeStruc", "scatterStruc"):
        assert batched[name][:64].tobytes() == single[name][:64].tobytes(), name
    assert (batched["planeStruc"] <= 0.6).all()
