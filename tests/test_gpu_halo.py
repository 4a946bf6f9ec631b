"""GPU: fixed-capacity halo classification (pcp_halo_pack) — two slab ranks emulated on one GPU (their
send buffers copied into each other's halo regions, as the NCCL exchange does) must reproduce the
unsharded run bit for bit, padding included."""
import numpy as np
import pytest
import torch

from pointcloudprocessing_b200 import capi, sharding, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", ["knn", "radius"])
def test_two_slabs_equal_unsharded(pcp, mode):
    pts = synth.urban(31, 60000, extent=50.0)
    with capi.Pcp(0) as full:
        full.set_cloud(pts)
        if mode == "knn":
            full.build_grid(0.0, 20)
            want = full.knn_pca(20)
        else:
            full.build_grid(0.5005, 0)
            want = full.radius_pca(0.5, 13, capi.CENTER_QUERY, False)
    split, h, cap = 25.0, (3.0 if mode == "knn" else 0.5), 8192
    gid = np.arange(pts.shape[0], dtype=np.int32)
    ranks = []
    for r, m in enumerate((pts[:, 0] < split, pts[:, 0] >= split)):
        ctx = capi.Pcp(0)
        ex = sharding.SlabExchange(ctx, int(m.sum()), cap, "cuda", r, 2, 0, 0.0 if r == 0 else split, split if r == 0 else 50.0, h)
        ex.own_pts.copy_(torch.from_numpy(pts[m]))
        ex.own_gid.copy_(torch.from_numpy(gid[m]))
        ex.pack()
        ctx.synchronize()
        ranks.append((ctx, ex, gid[m]))
    counts = [ex.check() for _, ex, _ in ranks]
    assert counts[0][0] > 0 and counts[0][1] > 0  # strips at both faces are packed, whoever receives them
    # rank 0's high strip -> rank 1's low halo region, rank 1's low strip -> rank 0's high halo region
    for (dst, side_dst), (src, side_src) in (((1, 0), (0, 1)), ((0, 1), (1, 0))):
        rp, rg = ranks[dst][1].halo_region(side_dst)
        rp.copy_(ranks[src][1].send_pts[side_src])
        rg.copy_(ranks[src][1].send_gid[side_src])
    torch.cuda.synchronize()
    got = np.zeros(pts.shape[0], capi.FEATURE_DTYPE)
    for ctx, ex, g in ranks:
        ctx.set_cloud_device(ex.buf_pts.data_ptr(), ex.buf_pts.shape[0], ex.buf_gid.data_ptr(), ex.n_own)
        if mode == "knn":
            ctx.build_grid(0.0, 20)
            out = torch.empty(ex.n_own * 68, dtype=torch.uint8, device="cuda")
            ctx.knn_pca(20, capi.CENTER_QUERY, out_device=out.data_ptr())
        else:
            ctx.build_grid(0.5005, 0)
            out = torch.empty(ex.n_own * 68, dtype=torch.uint8, device="cuda")
            ctx.radius_pca(0.5, 13, capi.CENTER_QUERY, False, out_device=out.data_ptr())
        ctx.synchronize()
        got[g] = np.frombuffer(out.cpu().numpy().tobytes(), dtype=capi.FEATURE_DTYPE)
        ctx.close()
    assert got.tobytes() == want.tobytes()
