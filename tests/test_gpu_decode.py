"""GPU parity: Velodyne packet decode on the device (pcp_decode_packets) against (1) the golden fixtures the
reference's own prebuilt lib/libPacketDecoder.so produced (tests/golden/decoder_*.npz), (2) the pinned oracle
on longer streams, and (3) the stages that follow on the device-resident cloud."""
import os
import sys

import numpy as np
import pytest

from pointcloudprocessing_b200 import capi

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle", "ref_decoder"))
from make_packets import make_packets  # noqa: E402  (packet synthesis only; the reference binary is not needed here)

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", ["sweep", "ragged", "upper_bank"])
def test_decode_equals_reference_binary(pcp, name):
    g = np.load(os.path.join(GOLD, f"decoder_{name}.npz"))
    got = pcp.decode_packets(g["packets"])
    nc = got["n_complete"]
    assert nc == g["frame_offsets"].shape[0] - 1
    np.testing.assert_array_equal(got["frame_offsets"][: nc + 1], g["frame_offsets"])
    n = int(g["frame_offsets"][-1])
    for key in ("x", "y", "z", "distance"):
        np.testing.assert_array_equal(got[key][:n].view(np.uint64), g[key].view(np.uint64), err_msg=key)  # doubles, bit for bit
    for key in ("intensity", "laser_id", "azimuth", "ms"):
        np.testing.assert_array_equal(got[key][:n], g[key], err_msg=key)


@pytest.mark.parametrize("n_packets,az_step,zero_frac", [(1, 40, 0.0), (377, 40, 0.1), (64, 1, 0.9), (50, 3000, 1.0)])
def test_decode_equals_oracle_on_streams(pcp, oracle, n_packets, az_step, zero_frac):
    calib = capi.ring_calib_default()
    pk = make_packets(n_packets + az_step, n_packets, az0=35990, az_step=az_step, zero_frac=zero_frac)
    want = oracle.decode_packets(pk, np.ctypeslib.as_array(calib.vert_deg))
    if want["x"].shape[0] == 0:
        with pytest.raises(capi.PcpError):
            pcp.decode_packets(pk)
            pcp.build_grid(0.0, 10)  # an empty cloud cannot be searched
        return
    got = pcp.decode_packets(pk)
    assert got["n_complete"] == want["n_complete"] and got["n_points"] == want["x"].shape[0]
    np.testing.assert_array_equal(got["frame_offsets"], want["frame_offsets"])
    for key in ("x", "y", "z", "distance"):
        np.testing.assert_array_equal(got[key].view(np.uint64), want[key].view(np.uint64), err_msg=key)
    for key in ("intensity", "laser_id", "azimuth", "ms"):
        np.testing.assert_array_equal(got[key], want[key], err_msg=key)


def test_decoded_cloud_feeds_the_frame_pipeline(pcp, oracle):
    """Packets in, per-frame range images + k-NN PCA records out, the points never visiting the host: equal to
    running the same stages on the host-decoded frames one by one."""
    calib = capi.ring_calib_default()
    pk = make_packets(5, 160, az0=0, az_step=40, zero_frac=0.05)
    got = pcp.decode_packets(pk, want_arrays=False)
    offs, nc = got["frame_offsets"], got["n_complete"]
    assert nc >= 2
    rng, pti, lr = pcp.ring_project(frame_offsets=offs)
    pcp.build_grid(0.0, 10)
    feat = pcp.knn_pca(10)
    host = oracle.decode_packets(pk, np.ctypeslib.as_array(calib.vert_deg))
    cloud = np.stack([host["x"], host["y"], host["z"], host["intensity"].astype(np.float64)], 1).astype(np.float32)
    with capi.Pcp(0) as one:
        for f in range(nc):
            a, b = int(offs[f]), int(offs[f + 1])
            fr = np.ascontiguousarray(cloud[a:b])
            one.set_cloud(fr)
            r1, p1, l1 = one.ring_project()
            assert np.array_equal(r1[0].view(np.uint32), rng[f].view(np.uint32)) and np.array_equal(p1[0], pti[f])
            one.build_grid(0.0, 10)
            w = one.knn_pca(10)
            gf = feat[a:b].copy()
            gf["ptID"] -= np.int32(a)
            assert gf.tobytes() == w.tobytes(), f
