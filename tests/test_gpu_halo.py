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


def test_peer_store_exchange_three_slabs():
    """pcp_halo_push / pcp_halo_wait: three slab ranks emulated on one GPU (one context and stream each; a
    'peer' pointer is then plain device memory) exchange their strips by direct stores + flags over
    several epochs and reproduce the unsharded k-NN records bit for bit."""
    pts = synth.urban(33, 90000, extent=60.0)
    k = 16
    with capi.Pcp(0) as full:
        full.set_cloud(pts)
        full.build_grid(0.0, k)
        want = full.knn_pca(k)
    edges, h, cap = [0.0, 20.0, 40.0, 60.0], 8.0, 32768
    gid = np.arange(pts.shape[0], dtype=np.int32)
    masks = [((pts[:, 0] >= edges[r]) | (r == 0)) & ((pts[:, 0] < edges[r + 1]) | (r == 2)) for r in range(3)]
    assert sum(int(m.sum()) for m in masks) == pts.shape[0]
    ctxs = [capi.Pcp(0) for _ in range(3)]
    exs = [sharding.SlabExchangePeer(ctxs[r], int(masks[r].sum()), cap, 0, r, 3, 0, edges[r], edges[r + 1], h) for r in range(3)]
    metas = [ex.meta() for ex in exs]
    for r, ex in enumerate(exs):
        ex.connect(metas)
        ex.set_own(torch.from_numpy(pts[masks[r]]).cuda(), torch.from_numpy(gid[masks[r]]).cuda())
    torch.cuda.synchronize()
    outs = [torch.empty(ex.n_own * 68, dtype=torch.uint8, device="cuda") for ex in exs]
    for epoch in range(3):  # both buffer parities, flags reused
        for ex in exs:      # enqueue only: a rank's wait completes when its neighbours' pushes have run
            ex.step()
        got = np.zeros(pts.shape[0], capi.FEATURE_DTYPE)
        for r, (ctx, ex) in enumerate(zip(ctxs, exs)):
            ctx.build_grid(0.0, k)
            outs[r].zero_()
            torch.cuda.synchronize()  # torch's stream and the library's stream are not ordered
            ctx.knn_pca(k, capi.CENTER_QUERY, out_device=outs[r].data_ptr())
            ctx.synchronize()
            got[gid[masks[r]]] = np.frombuffer(outs[r].cpu().numpy().tobytes(), dtype=capi.FEATURE_DTYPE)
        assert got.tobytes() == want.tobytes(), epoch
    counts = [ex.check() for ex in exs]
    assert counts[1][0] > 0 and counts[1][1] > 0
    # the middle rank's low halo region holds rank 0's high strip: same multiset as the NCCL-path packer's
    rp, rg = exs[1].halo_region(0)
    m = rg != sharding.SlabExchangePeer.INT_MAX
    ids = np.sort(rg[m].cpu().numpy())
    want_ids = np.sort(gid[masks[0]][pts[masks[0], 0] >= edges[1] - h])
    assert np.array_equal(ids, want_ids)
    assert torch.isnan(rp[~m][:, 0]).all()
    for ex in exs:
        ex.close()
    for c in ctxs:
        c.close()


def test_halo_wait_times_out_instead_of_hanging():
    with capi.Pcp(0) as p:
        ex = sharding.SlabExchangePeer(p, 1000, 256, 0, 0, 2, 0, 0.0, 10.0, 1.0, timeout_ms=50)
        other = sharding.SlabExchangePeer(p, 1000, 256, 0, 1, 2, 0, 10.0, 20.0, 1.0, timeout_ms=50)
        metas = [ex.meta(), other.meta()]
        ex.connect(metas)
        pts = torch.rand(1000, 4, device="cuda") * 10.0
        ex.set_own(pts, torch.arange(1000, dtype=torch.int32, device="cuda"))
        ex.step()  # rank 1 never pushes: rank 0's wait must give up after 50 ms
        with pytest.raises(RuntimeError, match="timed out"):
            ex.check()
        ex.close()
        other.close()


def test_shard_id_ranking_paths_agree(monkeypatch):
    """A shard's tie order is the order of the global ids.  Small ids serve as the order key directly (no ranking);
    PCP_SHARD_SORT=rank ranks them: from a sort of the halo entries alone when the own ids ascend (a slab cut from a
    cloud in index order), by the full sort of all ids otherwise or with PCP_SHARD_SORT=full.  Every path must give the
    same records, halo padding (ids INT_MAX, non-finite points) included, and neighbour lists naming the same ids."""
    pts = synth.urban(35, 40000, extent=40.0)
    k, split, h = 12, 20.0, 4.0
    own = pts[:, 0] < split
    halo = (~own) & (pts[:, 0] < split + h)
    gid = np.arange(pts.shape[0], dtype=np.int32) * 3 + 7  # not contiguous
    rs = np.random.default_rng(5)
    hsel = rs.permutation(np.flatnonzero(halo))             # halo entries arrive in any order
    pad = 700
    cloud = np.concatenate([pts[own], pts[hsel], np.full((pad, 4), np.nan, np.float32)])
    ids = np.concatenate([gid[own], gid[hsel], np.full(pad, 0x7FFFFFFF, np.int32)])
    n_own = int(own.sum())

    def run(cloud_, ids_, mode):
        if mode:
            monkeypatch.setenv("PCP_SHARD_SORT", mode)
        else:
            monkeypatch.delenv("PCP_SHARD_SORT", raising=False)
        with capi.Pcp(0) as p:  # the knob is read when the context is created
            d_c, d_i = torch.from_numpy(cloud_).cuda(), torch.from_numpy(ids_).cuda()
            torch.cuda.synchronize()  # torch's stream and the library's stream are not ordered
            p.set_cloud_device(d_c.data_ptr(), cloud_.shape[0], d_i.data_ptr(), n_own)
            p.build_grid(0.0, k)
            idx, d2 = p.knn_neighbors(k)
            rec = p.knn_pca(k)
            p.build_grid(0.45, 0)
            return rec, idx, d2, p.radius_pca(0.45, 13, capi.CENTER_QUERY, False)

    def same(a, b, perm=None):
        a = a if perm is None else tuple(x[perm] for x in a)
        assert a[0].tobytes() == b[0].tobytes() and a[3].tobytes() == b[3].tobytes()
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[2].view(np.uint32), b[2].view(np.uint32))

    direct = run(cloud, ids, None)
    same(direct, run(cloud, ids, "rank"))                   # own ids ascend: halo-only sort
    same(direct, run(cloud, ids, "full"))
    perm = rs.permutation(n_own)                            # own ids no longer ascend
    cloud_p, ids_p = cloud.copy(), ids.copy()
    cloud_p[:n_own], ids_p[:n_own] = cloud[:n_own][perm], ids[:n_own][perm]
    same(direct, run(cloud_p, ids_p, None), perm)
    same(direct, run(cloud_p, ids_p, "rank"), perm)         # falls back to the full sort
    # ids too large for the dense table are ranked (fast path here): same geometry, so the same records up to the id fields
    big = ids.copy()
    big[big != 0x7FFFFFFF] += 1_500_000_000
    far = run(cloud, big, None)
    assert np.array_equal(far[1], direct[1] + 1_500_000_000) and np.array_equal(far[2].view(np.uint32), direct[2].view(np.uint32))
    for name in capi.FEATURE_DTYPE.names:
        if name != "ptID":
            assert far[0][name].tobytes() == direct[0][name].tobytes(), name
    assert np.array_equal(far[0]["ptID"], direct[0]["ptID"] + 1_500_000_000)
    # the lists name global ids of real points (never the padding)
    assert np.isin(direct[1], gid).all()
