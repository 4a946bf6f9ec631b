"""Where the sharded step's time goes (run under torchrun, N >= 2): full step vs. no exchange vs. no pack."""
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
pts = synth.urban(100 + rank, n, extent=L)
pts[:, 0] = np.clip(pts[:, 0], 0.0, L - 1e-3) + np.float32(rank * L)
p = capi.Pcp(local)
stream = torch.cuda.ExternalStream(p.stream, device=local)
cap = max(4096, int(n * 1.6 * halo / L))
ex = sharding.SlabExchange(p, n, cap, local, rank, world, 0, rank * L, (rank + 1) * L, halo)
ex.own_pts.copy_(torch.from_numpy(pts).cuda(local))
ex.own_gid.copy_(torch.arange(n, dtype=torch.int32, device=local) + rank * n)
d_out = torch.empty(n * 68, dtype=torch.uint8, device=local)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=local)
torch.cuda.synchronize()


def hand_over():
    p.set_cloud_device(ex.buf_pts.data_ptr(), ex.buf_pts.shape[0], ex.buf_gid.data_ptr(), n, keep=(ex.buf_pts, ex.buf_gid))


def run(mode, steps=20):
    tot = 0.0
    stage = {}
    for s in range(steps + 3):
        with torch.cuda.stream(stream):
            flush.fill_(s & 0xFF)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        with torch.cuda.stream(stream):
            if mode == "full":
                ex.step()
            elif mode == "pack_only":
                ex.pack()
                hand_over()
            else:
                hand_over()
        p.build_grid(0.0, k)
        p.knn_pca(k, capi.CENTER_QUERY, out_device=d_out.data_ptr())
        b.record(stream)
        t = p.timings()
        if s >= 3:
            tot += a.elapsed_time(b)
            for key, v in t.items():
                stage[key] = stage.get(key, 0.0) + v
    dist.barrier()
    return tot / steps, {key: round(v / steps, 3) for key, v in stage.items()}


for mode in ("full", "pack_only", "static", "full"):
    ms, st = run(mode)
    print(f"rank {rank} {mode:10s} {ms:.3f} ms/step  stages {st}", flush=True)
p.set_cloud_device(ex.own_pts.data_ptr(), n)
tot = 0.0
for s in range(13):
    with torch.cuda.stream(stream):
        flush.fill_(s & 0xFF)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    p.set_cloud_device(ex.own_pts.data_ptr(), n)
    p.build_grid(0.0, k)
    p.knn_pca(k, capi.CENTER_QUERY, out_device=d_out.data_ptr())
    b.record(stream)
    p.synchronize()
    if s >= 3:
        tot += a.elapsed_time(b)
print(f"rank {rank} own-only   {tot / 10:.3f} ms/step", flush=True)
dist.destroy_process_group()
