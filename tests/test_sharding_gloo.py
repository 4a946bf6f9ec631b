import pytest
"""CPU, world_size 2 over gloo: slab sharding + halo exchange (the N>1 host logic of bench.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pointcloudprocessing_b200 import sharding, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, pts_np, k, h, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle_py as O

    pts = torch.from_numpy(pts_np)
    gid_all = torch.arange(pts.shape[0], dtype=torch.int32)
    lo_g, hi_g = float(pts[:, 0].min()), float(pts[:, 0].max()) + 1e-3
    # every rank starts from an arbitrary half of the cloud and derives the same splitters
    mine = pts[rank::world]
    edges = sharding.equal_count_splitters(mine[:, 0], lo_g, hi_g, world, bins=512)
    lo, hi = edges[rank], edges[rank + 1]
    own = (pts[:, 0] >= lo) & (pts[:, 0] < hi)
    p_all, g_all, n_own = sharding.exchange_halo(pts[own].contiguous(), gid_all[own].contiguous(), 0, lo, hi, h, rank, world)
    assert n_own == int(own.sum())
    # the shard holds exactly the points of [lo-h, hi+h) that the neighbours own
    want = (pts[:, 0] >= lo - h) & (pts[:, 0] < hi + h) if 0 < rank < world - 1 else (
        (pts[:, 0] < hi + h) if rank == 0 else (pts[:, 0] >= lo - h))
    assert sorted(g_all.tolist()) == sorted(gid_all[want].tolist())
    # local k-NN on the shard (oracle as the stand-in engine), mapped to global ids
    local = p_all.numpy()
    idx, d2 = O.KdTree(local).knn_all(k)
    # ties must break by GLOBAL id: emulate with the rank of the gid, as libpcp does
    order = np.argsort(g_all.numpy(), kind="stable")
    local_sorted = np.ascontiguousarray(local[order])
    idx_s, d2_s = O.KdTree(local_sorted).knn_all(k)
    inv = np.empty_like(order)
    inv[order] = np.arange(order.size)
    gids_sorted = g_all.numpy()[order]
    own_rows = inv[:n_own]
    nn_gid = gids_sorted[idx_s[own_rows]]
    kth = torch.from_numpy(d2_s[own_rows, k - 1].copy())
    bad = sharding.knn_halo_violations(p_all[:n_own], kth, 0, lo, hi, h, rank, world)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), gid=g_all.numpy()[:n_own], nn=nn_gid, bad=bad, edges=np.array(edges))
    dist.barrier()
    dist.destroy_process_group()


def test_slab_halo_exchange_world2(tmp_path, oracle):
    pts = synth.urban(31, 6000, extent=25.0)
    k, h = 10, 3.0
    port = _free_port()
    mp.spawn(_worker, args=(2, port, pts, k, h, str(tmp_path)), nprocs=2, join=True)
    ref, _ = oracle.KdTree(pts).knn_all(k)
    seen = []
    for r in range(2):
        z = np.load(tmp_path / f"r{r}.npz")
        assert int(z["bad"]) == 0
        np.testing.assert_array_equal(z["nn"], ref[z["gid"]])  # identical to the unsharded run
        seen += z["gid"].tolist()
    assert sorted(seen) == list(range(pts.shape[0]))  # every point owned exactly once
    e = np.load(tmp_path / "r0.npz")["edges"]
    n0 = ((pts[:, 0] >= e[0]) & (pts[:, 0] < e[1])).sum()
    assert abs(n0 - 3000) < 150  # equal-count splitters


def test_halo_violation_detection():
    pts = torch.tensor([[5.0, 0, 0, 0], [9.5, 0, 0, 0]])
    kth = torch.tensor([1.0, 4.0])  # k-th distances 1 and 2
    # slab [0,10), h = 1, interior rank: covered region is [-1, 11)
    assert sharding.knn_halo_violations(pts, kth, 0, 0.0, 10.0, 1.0, 1, 3) == 1
    assert sharding.knn_halo_violations(pts, kth, 0, 0.0, 10.0, 2.0, 1, 3) == 0
    assert sharding.knn_halo_violations(pts, kth, 0, 0.0, 10.0, 0.1, 0, 1) == 0  # single rank: open faces


def test_frame_shard_partition():
    for n, w in ((10000, 8), (7, 3), (2, 4)):
        spans = [sharding.frame_shard(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [e - b for b, e in spans]
        assert max(sizes) - min(sizes) <= 1


def test_halo_reach_check():
    """A halo wider than a slab cannot be served by the two adjacent slabs alone: refused up front."""
    from pointcloudprocessing_b200 import sharding

    sharding.check_halo_reach([0.0, 10.0, 20.0, 30.0], 6.0)
    sharding.check_halo_reach([float("-inf"), 10.0, 20.0, float("inf")], 6.0)
    sharding.check_halo_reach([0.0, 5.0], 100.0)  # one rank: nothing to exchange
    with pytest.raises(ValueError):
        sharding.check_halo_reach([0.0, 10.0, 14.0, 30.0], 6.0)
