"""GPU: pcp_carry_forward_static — the stale-record carry as an entry point that allocates nothing (its scratch is a
static device array inside libpcp.so, no counterpart in the reference).
Called through the C-ABI on DEVICE records and compared with the existing entry point, pcp_radius_pca(carry_forward = 1),
on two calls in a row (the second on another cloud, so it finds the first call's contents in the shared scratch).  The
carry only copies record words, so the comparison is bitwise — inside every tolerance of tests/parity.py.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import capi, synth

pytestmark = pytest.mark.gpu


def _run(p, torch, pts, r, min_pts):
    n = pts.shape[0]
    p.set_cloud(pts)
    p.build_grid(r * 1.001, 0)
    want = p.radius_pca(r, min_pts, capi.CENTER_QUERY, True)  # existing entry point: carry applied inside
    d_raw = torch.empty(n * 68, dtype=torch.uint8, device="cuda:0")
    p.radius_pca(r, min_pts, capi.CENTER_QUERY, False, out_device=d_raw.data_ptr())  # records without the carry, on the device
    p.synchronize()
    raw = d_raw.cpu().numpy().view(capi.FEATURE_DTYPE)
    d_out = torch.empty_like(d_raw)
    p.carry_forward_static(d_raw.data_ptr(), d_out.data_ptr(), n, min_pts)  # separate output
    p.synchronize()
    got = d_out.cpu().numpy().view(capi.FEATURE_DTYPE)
    p.carry_forward_static(d_raw.data_ptr(), d_raw.data_ptr(), n, min_pts)  # in place
    p.synchronize()
    inplace = d_raw.cpu().numpy().view(capi.FEATURE_DTYPE)
    return want, raw, got, inplace


def test_carry_forward_static_matches_radius_pca():
    import torch

    with capi.Pcp(0) as p:
        assert capi.load().pcp_carry_forward_static_max() == 1 << 20
        first = synth.urban(6, 300_000, extent=95.0)       # 293 blocks of the scan, many records below the gate
        second = synth.urban(7, 41_003, extent=33.0)[::-1].copy()  # smaller: most of the scratch still holds the first call's values
        for pts, r in ((first, 0.5), (second, 0.4)):
            want, raw, got, inplace = _run(p, torch, pts, r, 13)
            invalid = want["ptNum"] < 13
            assert 0.02 < invalid.mean() < 0.98
            assert raw.tobytes() != want.tobytes()  # the carry does change records
            assert got.tobytes() == want.tobytes(), "separate output differs from pcp_radius_pca(carry_forward = 1)"
            assert inplace.tobytes() == want.tobytes(), "in-place result differs from pcp_radius_pca(carry_forward = 1)"
        # the size limit is an error, not a silent truncation; nothing is launched
        with pytest.raises(capi.PcpError) as e:
            p.carry_forward_static(1, 1, (1 << 20) + 1, 13)
        assert e.value.status == -1 and "1048576" in str(e.value)
        p.carry_forward_static(0, 0, 0, 13)  # empty: fine
