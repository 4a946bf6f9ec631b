"""bench.py — kNN+PCA feature points/sec (k=20) on B200, BASELINE.json's metric.

A step = one pass of the hot path over one cloud: grid build (bbox, keys, radix sort, cell table) +
the fused k-NN search / moment / eigen kernel, 68-byte pcaFeature records out.  `value` is measured
with the cloud resident in HBM; `e2e` goes through the C-ABI with pinned HOST buffers (H2D of the
cloud and D2H of the records inside the timed region).  torch is used for device buffers, events on
the library's stream and the multi-rank barrier only.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--points P] [--k K]
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALGO_BYTES_PER_POINT = 84  # 16 B float4 read + 68 B pcaFeature written


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(n: int):
    """DRAM bytes per launch of the search stage (dram__bytes_read.sum + dram__bytes_write.sum of its two
    kernels) from the committed `ncu --set full` capture of this workload; null for any other size."""
    path = os.path.join(ROOT, "profiles", "r01f_knn_tile_and_worklist_ncu_summary.txt")
    if n != 1_000_000 or not os.path.exists(path):
        return None
    total = 0.0
    with open(path) as f:
        for line in f:
            if line.startswith("dram__bytes_read.sum [Mbyte]") or line.startswith("dram__bytes_write.sum [Mbyte]"):
                total += float(line.split("=")[1]) * 1e6
    return total or None


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def make_cloud(points: int, rank: int, per_rank_seed: bool = False):
    from pointcloudprocessing_b200 import synth

    # configs[2] shape: 1 M-point urban scan over 200 m x 200 m, random point order (worst case).
    # Weak scaling (N > 1): every rank owns a translated copy of the SAME tile, so the work per GPU is identical by
    # construction and the per-N values measure the sharding, not the luck of a seed (tiles of different seeds
    # differ by up to 17 % in search time on one GPU: 1.43 vs 1.71 ms, tools/shard_step_breakdown.py);
    # --rank-clouds seeded gives every rank its own tile as before.
    seed = 3 + 1000 * rank if per_rank_seed else 3
    return synth.urban(seed, points, extent=200.0 * (points / 1e6) ** 0.5, order="random")


def cpu_tile(pts: np.ndarray, target: int) -> np.ndarray:
    """Spatial crop (same density as the workload) of about `target` points."""
    if pts.shape[0] <= target:
        return pts
    frac = target / pts.shape[0]
    x0, x1 = pts[:, 0].min(), pts[:, 0].max()
    return np.ascontiguousarray(pts[pts[:, 0] < x0 + (x1 - x0) * frac])


def run_cpu(pts: np.ndarray, k: int, threads: int, target: int):
    from oracle import oracle_py as O

    tile = cpu_tile(pts, target)
    t0 = time.perf_counter()
    O.knn_pca_cloud(tile, k, 0, nthreads=threads)  # kd-tree build + per-point search + PCA, like the reference loop
    dt = time.perf_counter() - t0
    return tile.shape[0] / dt, tile.shape[0], dt


def reference_arm(args, rank, world):
    """The reference's CPU path (restated: oracle port; PCL/FLANN/Eigen cannot be built here) on the
    host cores, all threads, bounded sample per step."""
    if rank != 0:
        return
    from oracle import oracle_py as O

    pts = make_cloud(args.points, 0)
    threads = O.num_threads()
    target = args.cpu_sample
    for _ in range(min(args.warmup, 1)):
        run_cpu(pts, args.k, threads, target // 4)
    times, npts = [], 0
    for _ in range(args.steps):
        _, m, dt = run_cpu(pts, args.k, threads, target)
        times.append(dt)
        npts += m
    total = sum(times)
    value = npts / total
    line = {
        "impl": "reference", "metric": "kNN+PCA feature points/sec (k=%d)" % args.k, "value": value, "unit": "points/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"urban_{args.points}pts_knn_k{args.k}_pca (configs[2] shape, metric's k)"},
        "cpu_baseline": {"value": value, "unit": "points/s", "cores": threads, "kind": "port",
                         "sample": f"spatial tile of ~{target} points of the same cloud per step, OpenMP over points"},
        "e2e": {"value": value, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit_line(line)


def emit_line(line):
    """The bench contract's single stdout line (see the __main__ block: fd 1 itself points at stderr)."""
    fd = globals().get("_JSON_FD")
    if fd is None:
        emit_line(line)
    else:
        os.write(fd, (json.dumps(line) + "\n").encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--points", type=int, default=1_000_000, help="points per GPU")
    ap.add_argument("--k", type=int, default=20)
    ap.add_argument("--cell", type=float, default=0.0)
    ap.add_argument("--cpu-sample", type=int, default=1_000_000)
    ap.add_argument("--halo", type=float, default=6.0, help="slab halo width in metres (N > 1)")
    ap.add_argument("--halo-impl", default="peer", choices=["peer", "nccl"],
                    help="N > 1: halo strips by peer stores + flags (pcp_halo_push / pcp_halo_wait) or by NCCL send/recv")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--rank-clouds", default="tile", choices=["tile", "seeded"],
                    help="N > 1: every rank's slab is a translated copy of the same 1 M-point tile (identical work per GPU), or a tile of its own seed")
    ap.add_argument("--e2e-contexts", type=int, default=3, help="C-ABI contexts in flight for the end-to-end figure (N = 1)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    from pointcloudprocessing_b200 import capi

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from pointcloudprocessing_b200 import sharding

    n, k = args.points, args.k
    pts = make_cloud(n, rank, args.rank_clouds == "seeded")
    L = 200.0 * (n / 1e6) ** 0.5
    halo = args.halo
    if world > 1:
        # weak scaling: rank r owns the slab x in [r*L, (r+1)*L) of a world*L x L scan; neighbours
        # exchange the strips within `halo` of the shared faces every step (peer stores, or NCCL send/recv)
        pts[:, 0] = np.clip(pts[:, 0], 0.0, L - 1e-3) + np.float32(rank * L)
    lo, hi = rank * L, (rank + 1) * L
    p = capi.Pcp(local)  # raises if libpcp.so / the GPU is missing: no fallback
    stream = torch.cuda.ExternalStream(p.stream, device=local)
    d_pts = torch.from_numpy(pts).cuda(local)
    d_gid = (torch.arange(n, dtype=torch.int32, device=local) + rank * n) if world > 1 else None
    d_kth = torch.empty(n, dtype=torch.float32, device=local)
    d_out = torch.empty(n * 68, dtype=torch.uint8, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=local)  # > 126 MB L2
    h_in = torch.from_numpy(pts).pin_memory()
    h_out = torch.empty(n * 68, dtype=torch.uint8).pin_memory()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    keep = {}
    ex = None
    if world > 1:
        # fixed-capacity halo buffers: own points first, then the two received strips (sharding.SlabExchange)
        cap = max(4096, int(n * min(0.5, 1.6 * halo / L)))
        if args.halo_impl == "peer":
            # strips are written straight into the neighbours' halo regions (CUDA IPC mappings, NVLink stores)
            ex = sharding.SlabExchangePeer(p, n, cap, local, rank, world, 0, lo, hi, halo)
            ex.connect_group()
            ex.set_own(d_pts, d_gid)
        else:
            ex = sharding.SlabExchange(p, n, cap, local, rank, world, 0, lo, hi, halo)
            ex.own_pts.copy_(d_pts)
            ex.own_gid.copy_(d_gid)
        torch.cuda.synchronize()
        dist.barrier()

    def shard_step(src_pts):
        # halo classification kernel + grouped NCCL send/recv + sharded hand-over, all enqueued on the
        # library's stream: the exchange needs no host synchronisation (sizes are fixed)
        with torch.cuda.stream(stream):
            if src_pts is not None:
                ex.own_pts.copy_(src_pts, non_blocking=True)
            ex.step()
        p.build_grid(args.cell, k)
        p.knn_pca(k, capi.CENTER_QUERY, out_device=d_out.data_ptr(), kth_device=d_kth.data_ptr())

    def step_device():
        if world > 1:
            return shard_step(None)
        p.set_cloud_device(d_pts.data_ptr(), n)
        p.build_grid(args.cell, k)
        p.knn_pca(k, capi.CENTER_QUERY, out_device=d_out.data_ptr())


    def step_e2e():
        if world > 1:
            shard_step(h_in)  # the step's H2D copy goes straight into the own region of the halo buffer
            with torch.cuda.stream(stream):
                h_out.copy_(d_out, non_blocking=True)
            p.synchronize()
            return
        p.set_cloud_host_ptr(h_in.data_ptr(), n, 16)
        p.build_grid(args.cell, k)
        p.knn_pca(k, capi.CENTER_QUERY, out_host_ptr=h_out.data_ptr())

    for _ in range(max(args.warmup, 3)):
        step_device()
    p.synchronize()
    halo_unproven = 0
    if world > 1:
        # prove the halo was wide enough for every own query (k-th distance inside the covered region)
        ex.check()  # both strips fit the fixed-capacity buffers
        with torch.cuda.stream(stream):
            halo_unproven = sharding.knn_halo_violations(d_pts, d_kth, 0, lo, hi, halo, rank, world)
        if halo_unproven:
            raise SystemExit(f"rank {rank}: halo {halo} m too narrow for {halo_unproven} queries; raise --halo")

    # ---- device-resident timing: CUDA events on the library's stream, L2 flushed between steps
    sampler = ClockSampler(local)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kern_ms = 0.0
    barrier()
    sampler.start()
    l0 = p.launch_count()
    for s in range(args.steps):
        with torch.cuda.stream(stream):
            flush.fill_(s & 0xFF)
        ev[s][0].record(stream)
        step_device()
        ev[s][1].record(stream)
        kern_ms += p.timings()["search_pca"]  # synchronises the stream; per-stage events inside the library
    launches = p.launch_count() - l0
    barrier()
    clocks = sampler.stop()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)

    # ---- end to end through the C-ABI with pinned host buffers
    # (a) sequential: one context, every call synchronous, like the reference's class methods
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for s in range(args.steps):
        step_e2e()
    p.synchronize()
    e2e_seq_s = time.perf_counter() - t0
    barrier()
    # (b) pipelined: several contexts in pcp_set_host_async mode, the next steps' upload + grid + search
    # overlap a step's 68 MB result copy; every step still uploads its cloud and downloads its records
    e2e_s = e2e_seq_s
    e2e_mode = "sequential, 1 context"
    if world == 1:
        extra = [capi.Pcp(local) for _ in range(args.e2e_contexts - 1)]
        ctxs = [p] + extra
        nctx = len(ctxs)
        h_outs = [h_out] + [torch.empty(n * 68, dtype=torch.uint8).pin_memory() for _ in extra]
        for c in ctxs:
            c.set_host_async(True)

        def step_async(s):
            c = ctxs[s % nctx]
            c.synchronize()  # its previous cloud (step s-2) is complete; buffers may be reused
            c.set_cloud_host_ptr(h_in.data_ptr(), n, 16)
            c.build_grid(args.cell, k)
            c.knn_pca(k, capi.CENTER_QUERY, out_host_ptr=h_outs[s % nctx].data_ptr())

        for s in range(2 * nctx):
            step_async(s)
        for c in ctxs:
            c.synchronize()
        barrier()
        t0 = time.perf_counter()
        for s in range(args.steps):
            step_async(s)
        for c in ctxs:
            c.synchronize()
        e2e_s = time.perf_counter() - t0
        e2e_mode = f"{nctx} contexts in flight (pcp_set_host_async), each step = H2D cloud + grid + search + D2H records"
        for c in ctxs:
            c.set_host_async(False)
        for c in extra:
            c.close()
        barrier()

    t = torch.tensor([dev_ms, e2e_s * 1e3, kern_ms], dtype=torch.float64, device=local)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kern_ms = t.tolist()
    total_pts = n * world * args.steps
    value = total_pts / (dev_ms * 1e-3)
    peak, peak_src = peaks()
    kern_avg_s = kern_ms * 1e-3 / args.steps
    achieved = ALGO_BYTES_PER_POINT * n / kern_avg_s / 1e9
    line = {
        "metric": "kNN+PCA feature points/sec (k=%d)" % k, "value": value, "unit": "points/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"urban_{n}pts_knn_k{k}_pca (configs[2] shape, metric's k)", "points_per_gpu": n, "k": k,
                   "order": "random", "l2": "flushed between steps (256 MiB fill, untimed)",
                   "step": "bbox+keys+radix sort+cell table+fused search/moment/eigen, 68 B records",
                   "sharding": "single cloud" if world == 1 else
                   f"x-slabs of one {world}x{L:.0f} m x {L:.0f} m scan (" +
                   ("each slab a translated copy of the N = 1 tile: identical work per GPU" if args.rank_clouds == "tile" else "one tile per rank, seeds 3 + 1000 rank") +
                   f"), halo {halo} m exchanged every step (" +
                   ("pcp_halo_push: the classification kernel stores the strips straight into the neighbours' halo regions over NVLink and raises their flags; pcp_halo_wait on the receiver; no NCCL, no host sync"
                    if args.halo_impl == "peer" else
                    "pcp_halo_pack kernel + grouped NCCL send/recv into fixed-capacity buffers, no host sync in the exchange") +
                   "), global-index tie-break"},
        "clocks": clocks,
        "e2e": {"value": total_pts / (e2e_ms * 1e-3), "unit": "points/s", "h2d_bytes_per_step": 16 * n, "d2h_bytes_per_step": 68 * n,
                "mode": e2e_mode, "sequential_value": total_pts / e2e_seq_s},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(n),
                     "kernel": "knn_tile_kernel + search_pca_worklist_kernel<0> (one stage: the tile kernel serves ~90 % of the queries, the work-list kernel the rest)", "bytes_per_point": ALGO_BYTES_PER_POINT, "peak_source": peak_src,
                     "kernel_ms": kern_avg_s * 1e3,
                     "note": "k-NN selection is issue-bound (≈1.5-3 kFLOP + top-k per point), far below the HBM roof by construction"},
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle_py as O

        threads = O.num_threads()
        v1, m1, _ = run_cpu(pts, k, 1, args.cpu_sample // 4)
        best = None
        for _ in range(3):
            r_ = run_cpu(pts, k, threads, args.cpu_sample)
            best = r_ if best is None or r_[0] > best[0] else best
        vall, mall, dt = best
        line["cpu_baseline"] = {"value": vall, "unit": "points/s", "cores": threads, "kind": "port", "value_1core": v1,
                                "sample": f"spatial tile of {mall} points (all cores, {dt:.1f} s) / {m1} points (1 core) of the same cloud"}
    if rank == 0:
        emit_line(line)
    # orderly teardown: tensors that lived on the library's stream go first, then the stream itself
    p.synchronize()
    keep.clear()
    if world > 1:
        dist.barrier()  # nobody unmaps while a neighbour may still store into the mapping
        if hasattr(ex, "close"):
            ex.close()
    ex = None
    del d_pts, d_out, d_kth, flush, d_gid
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    p.close()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    # stdout carries the ONE JSON line and nothing else: libraries that write to file descriptor 1 (NCCL prints
    # its version banner there when NCCL_DEBUG is set on the box) are sent to stderr for the whole run
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    main()
