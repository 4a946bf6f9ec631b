"""Sharded step with the peer-store halo exchange (torchrun, N >= 2): CUDA-event times of its three parts
(exchange = push + wait, grid build incl. the global-id sort, search) next to the unsharded step on the same points."""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import capi, sharding, synth

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, k, halo, L = 1_000_000, 20, 6.0, 200.0
pts = synth.urban(3 + 1000 * rank, n, extent=L, order="random")
pts[:, 0] = np.clip(pts[:, 0], 0.0, L - 1e-3) + np.float32(rank * L)
p = capi.Pcp(local)
stream = torch.cuda.ExternalStream(p.stream, device=local)
d_pts = torch.from_numpy(pts).cuda(local)
d_gid = torch.arange(n, dtype=torch.int32, device=local) + rank * n
d_out = torch.empty(n * 68, dtype=torch.uint8, device=local)
d_kth = torch.empty(n, dtype=torch.float32, device=local)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=local)
cap = max(4096, int(n * min(0.5, 1.6 * halo / L)))
ex = sharding.SlabExchangePeer(p, n, cap, local, rank, world, 0, rank * L, (rank + 1) * L, halo)
ex.connect_group()
ex.set_own(d_pts, d_gid)
torch.cuda.synchronize()
dist.barrier()


def run(sharded, steps=20, sync_ranks=False):
    acc = np.zeros(4)
    for s in range(steps + 3):
        if sync_ranks:
            dist.barrier()
        with torch.cuda.stream(stream):
            flush.fill_(s & 0xFF)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record(stream)
        if sharded:
            with torch.cuda.stream(stream):
                ex.step()
        else:
            p.set_cloud_device(d_pts.data_ptr(), n)
        e[1].record(stream)
        p.build_grid(0.0, k)
        e[2].record(stream)
        p.knn_pca(k, capi.CENTER_QUERY, out_device=d_out.data_ptr(), kth_device=d_kth.data_ptr())
        e[3].record(stream)
        p.synchronize()
        if s >= 3:
            acc += [e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), e[2].elapsed_time(e[3]), e[0].elapsed_time(e[3])]
    return acc / steps


for name, sh, sy in (("unsharded", False, False), ("sharded", True, False), ("sharded, ranks aligned by a barrier per step", True, True)):
    a = run(sh, sync_ranks=sy)
    dist.barrier()
    print(f"rank {rank} {name}: exchange {a[0]:.3f}  grid {a[1]:.3f}  search {a[2]:.3f}  step {a[3]:.3f} ms", flush=True)
    dist.barrier()
dist.destroy_process_group()
os._exit(0)
