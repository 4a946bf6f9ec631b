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


def host_threads() -> int:
    """Host cores this process may use.  torchrun exports OMP_NUM_THREADS=1; the CPU arm passes this count to the
    oracle explicitly (num_threads clause), so the environment variable does not shrink it."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def source_id() -> str:
    """Hash of the CUDA sources the search stage (grid build + the two k-NN kernels) is compiled from: names the ncu
    capture the traffic figure belongs to.  Sources of other entry points (text parsing, clustering, ...) do not enter."""
    import hashlib

    h = hashlib.sha1()
    csrc = os.path.join(ROOT, "pointcloudprocessing_b200", "csrc")
    stage = ("Makefile", "eigen3.cuh", "grid.cu", "knn_finish.cuh", "knn_tile.cu", "knn_tile6.cu", "pcp_internal.cuh",
             "search_common.cuh", "search_pca.cu", "sortnet32.cuh", "tile_common.cuh")
    import re

    for name in sorted(os.listdir(csrc)):
        if name in stage:
            with open(os.path.join(csrc, name), "r", encoding="utf-8", errors="replace") as f:
                text = f.read()
            # comments and layout do not change the kernels: hash the code alone
            text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
            text = re.sub(r"//[^\n]*", " ", text)
            text = " ".join(text.split())
            h.update(name.encode() + b"\0" + text.encode())
    return h.hexdigest()[:12]


def workload_config(n: int, k: int, world: int, rank_clouds: str = "tile", halo: float = 6.0, halo_impl: str = "peer") -> dict:
    """`config` of the JSON line; both arms (ours / --impl reference) emit the same keys and values."""
    L = 200.0 * (n / 1e6) ** 0.5
    return {"workload": f"urban_{n}pts_knn_k{k}_pca (configs[2] shape, metric's k)", "points_per_gpu": n, "k": k,
            "order": "random", "l2": "flushed between steps (256 MiB fill, untimed)",
            "step": "bbox+keys+radix sort+cell table+fused search/moment/eigen, 68 B records",
            "sharding": "single cloud" if world == 1 else
            f"x-slabs of one {world}x{L:.0f} m x {L:.0f} m scan (" +
            ("each slab a translated copy of the N = 1 tile: identical work per GPU" if rank_clouds == "tile" else "one tile per rank, seeds 3 + 1000 rank") +
            f"), halo {halo} m exchanged every step (" +
            ("pcp_halo_push: the classification kernel stores the strips straight into the neighbours' halo regions over NVLink and raises their flags; pcp_halo_wait on the receiver; no NCCL, no host sync"
             if halo_impl == "peer" else
             "pcp_halo_pack kernel + grouped NCCL send/recv into fixed-capacity buffers, no host sync in the exchange") +
            "), global-index tie-break"}


def ncu_capture(n: int):
    """(DRAM bytes, warp instructions) per launch of the search stage from the `ncu --set full` capture of THIS
    build — profiles/r02_search_ncu_<source_id()>.txt, written by tools/ncu_summary.py from a capture of this very
    command — or (None, None): a capture of other sources says nothing about these kernels."""
    path = os.path.join(ROOT, "profiles", f"r02_search_ncu_{source_id()}.txt")
    if n != 1_000_000 or not os.path.exists(path):
        return None, None
    dram, inst = 0.0, 0.0
    with open(path) as f:
        for line in f:
            if line.startswith("dram__bytes_read.sum [Mbyte]") or line.startswith("dram__bytes_write.sum [Mbyte]"):
                dram += float(line.split("=")[1]) * 1e6
            elif line.startswith("smsp__inst_executed.sum [inst]"):
                inst += float(line.split("=")[1])
    return (dram or None), (inst or None)


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
    threads = host_threads()  # not O.num_threads(): torchrun sets OMP_NUM_THREADS=1
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
        "config": workload_config(args.points, args.k, world, args.rank_clouds, args.halo, args.halo_impl),
        "cpu_baseline": {"value": value, "unit": "points/s", "cores": threads, "kind": "port",
                         "sample": f"spatial tile of ~{target} points of the same cloud per step, OpenMP over points"},
        "e2e": {"value": value, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit_line(line)


def bind_to_gpu_numa_node(index: int):
    """N > 1: run this rank (and first-touch its pinned buffers) on the NUMA node its GPU hangs off, so that eight ranks'
    host<->device copies do not all cross the socket interconnect.  Returns a note for the JSON line; never raises."""
    try:
        import pynvml

        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        if len(bus.split(":")[0]) == 8:  # NVML prints an 8-digit domain, sysfs a 4-digit one
            bus = bus[4:]
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            node = int(f.read().strip())
        if node < 0:
            return "no NUMA information for the GPU"
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"NUMA node {node} has no allowed CPUs"
        os.sched_setaffinity(0, cpus)
        return f"bound to NUMA node {node} ({len(cpus)} CPUs) of GPU {bus}"
    except Exception as e:  # noqa: BLE001 — binding is an optimisation, never a reason to fail
        return f"not bound ({type(e).__name__})"


def rank_cloud(n: int, rank: int, world: int, seeded: bool) -> np.ndarray:
    """The slab rank `rank` owns: the N = 1 cloud, shifted to x in [rank L, (rank + 1) L) when world > 1."""
    pts = make_cloud(n, rank, seeded)
    if world > 1:
        L = 200.0 * (n / 1e6) ** 0.5
        pts[:, 0] = np.clip(pts[:, 0], 0.0, L - 1e-3) + np.float32(rank * L)
    return pts


def check_parity(args, p, stream, pts, d_out, n, k, rank, world, local, step_device):
    """Compares the records of one more (untimed) step with the checker; raises SystemExit on any difference.
    N = 1: neighbour indices and d2 of EVERY query bit-exact against the CPU oracle (pcp_knn_neighbors), the 68-byte
    records against the oracle's within tests/parity.py's tolerances.  N > 1: this rank's records bitwise equal to the
    records of the same points in an unsharded run of the union cloud on this GPU (global-index tie-break)."""
    import torch

    from pointcloudprocessing_b200 import capi

    step_device()
    p.synchronize()
    got = d_out.cpu().numpy().view(capi.FEATURE_DTYPE)
    if world > 1:
        union = np.concatenate([rank_cloud(n, r, world, args.rank_clouds == "seeded") for r in range(world)])
        d_union = torch.from_numpy(union).cuda(local)
        d_uout = torch.empty(union.shape[0] * 68, dtype=torch.uint8, device=local)
        with capi.Pcp(local) as q:
            q.set_cloud_device(d_union.data_ptr(), union.shape[0])
            q.build_grid(args.cell, k)
            q.knn_pca(k, capi.CENTER_QUERY, out_device=d_uout.data_ptr())
            q.synchronize()
        mine = d_uout[rank * n * 68:(rank + 1) * n * 68]
        if not torch.equal(mine, d_out):
            diff = int((mine.view(-1, 68) != d_out.view(-1, 68)).any(1).sum().item())
            raise SystemExit(f"rank {rank}: {diff} sharded records differ from the unsharded run of the union cloud")
        return n, f"every rank's {n} records bitwise equal to an unsharded run of the {world * n}-point union cloud on its own GPU"
    from oracle import oracle_py as O

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from parity import check_features

    threads = host_threads()
    idx, d2 = p.knn_neighbors(k)
    ridx, rd2 = O.KdTree(pts).knn_all(k, threads)
    if not np.array_equal(idx, ridx):
        raise SystemExit(f"parity: neighbour indices differ from the oracle on {int((idx != ridx).any(1).sum())} of {n} queries")
    if not np.array_equal(d2.view(np.uint32), rd2.view(np.uint32)):
        raise SystemExit("parity: neighbour distances differ from the oracle (bit compare)")
    ref = O.knn_pca_cloud(pts, k, 0, nthreads=threads)
    try:
        check_features(got, ref, None, "bench records")
    except AssertionError as e:
        raise SystemExit(f"parity: {e}")
    return n, (f"all {n} queries: k-NN indices and d2 bit-exact vs the CPU oracle (pcp_knn_neighbors), the timed step's records within "
               "tests/parity.py's tolerances (ptID/ptNum exact, eigenvalues 1e-4, eigenvectors up to sign 1e-4)")


def emit_line(line):
    """The bench contract's single stdout line (see the __main__ block: fd 1 itself points at stderr)."""
    fd = globals().get("_JSON_FD")
    data = (json.dumps(line) + "\n").encode()
    if fd is None:  # imported as a module: plain stdout
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(fd, data)


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
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle / unsharded-run comparison of the timed records")
    ap.add_argument("--rank-clouds", default="tile", choices=["tile", "seeded"],
                    help="N > 1: every rank's slab is a translated copy of the same 1 M-point tile (identical work per GPU), or a tile of its own seed")
    ap.add_argument("--e2e-contexts", type=int, default=3, help="C-ABI contexts in flight for the end-to-end figure (N = 1)")
    ap.add_argument("--e2e-prefetch", type=int, default=0, help="1: enqueue the next step's upload before building the current step's grid (N = 1 pipelined leg; measured slower: 412 vs 583 M points/s with 3 contexts)")
    ap.add_argument("--e2e-threads", type=int, default=0,
                    help="0: one host thread issues every context's calls (default: measured faster, 506 vs 454 M points/s); 1: one host thread per context")
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
    numa_note = bind_to_gpu_numa_node(local) if world > 1 else "N = 1: not bound (the CPU baseline uses every core)"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from pointcloudprocessing_b200 import sharding

    n, k = args.points, args.k
    # weak scaling: rank r owns the slab x in [r*L, (r+1)*L) of a world*L x L scan; neighbours
    # exchange the strips within `halo` of the shared faces every step (peer stores, or NCCL send/recv)
    pts = rank_cloud(n, rank, world, args.rank_clouds == "seeded")
    L = 200.0 * (n / 1e6) ** 0.5
    halo = args.halo
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
        sharding.check_halo_reach([r_ * L for r_ in range(world + 1)], halo)  # halos come from the adjacent slabs only
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

    def shard_step(src_pts, out=None):
        # halo classification kernel + grouped NCCL send/recv + sharded hand-over, all enqueued on the
        # library's stream: the exchange needs no host synchronisation (sizes are fixed)
        with torch.cuda.stream(stream):
            if src_pts is not None:
                ex.own_pts.copy_(src_pts, non_blocking=True)
            ex.step()
        p.build_grid(args.cell, k)
        p.knn_pca(k, capi.CENTER_QUERY, out_device=(d_out if out is None else out).data_ptr(), kth_device=d_kth.data_ptr())

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

        def upload(s):
            c = ctxs[s % nctx]
            c.synchronize()  # its previous cloud (step s - nctx) is complete; buffers may be reused
            c.set_cloud_host_ptr(h_in.data_ptr(), n, 16)  # asynchronous mode: the bounding-box kernels follow the copy

        def step_async(s, last=True):
            # --e2e-prefetch 1 (measured and rejected, kept for A/B): step s + 1's upload is enqueued (on its own context) BEFORE step s's grid is built, so
            # that the one host wait of a build (the bounding box) finds its cloud long uploaded instead of holding the
            # host — and with it the launches of the other contexts — for the duration of the copy
            c = ctxs[s % nctx]
            if not args.e2e_prefetch:
                upload(s)
            elif not last:
                upload(s + 1)
            c.build_grid(args.cell, k)
            c.knn_pca(k, capi.CENTER_QUERY, out_host_ptr=h_outs[s % nctx].data_ptr())

        # One host thread per context (ctypes releases the GIL inside the library): a context's grid build waits on the
        # host for its bounding box, and with a single thread that wait also held back the launches of the other contexts.
        # Thread t runs the steps s = t, t + nctx, ... of the args.steps timed steps; the clock stops when all are done.
        def run_steps(first, count):
            if args.e2e_threads:
                go = threading.Barrier(nctx + 1)

                def worker(t):
                    go.wait()
                    for s in range(first + t, first + count, nctx):
                        if args.e2e_prefetch:
                            upload(s)
                        step_async(s, last=True)
                    ctxs[t].synchronize()

                ths = [threading.Thread(target=worker, args=(t,)) for t in range(nctx)]
                for th in ths:
                    th.start()
                go.wait()
                t_start = time.perf_counter()
                for th in ths:
                    th.join()
                return time.perf_counter() - t_start
            t_start = time.perf_counter()
            if args.e2e_prefetch:
                upload(first)
            for s in range(first, first + count):
                step_async(s, last=(s == first + count - 1))
            for c in ctxs:
                c.synchronize()
            return time.perf_counter() - t_start

        run_steps(0, 2 * nctx)
        barrier()
        e2e_s = run_steps(2 * nctx, args.steps)
        e2e_mode = (f"{nctx} contexts in flight (pcp_set_host_async), " + ("one host thread each, " if args.e2e_threads else "one host thread, ") +
                    ("step s + 1's upload enqueued before step s's grid build, " if args.e2e_prefetch and not args.e2e_threads else "") +
                    "each step = H2D cloud + grid + search + D2H records")
        for c in ctxs:
            c.set_host_async(False)
        for c in extra:
            c.close()
        barrier()

    # ---- parity of the records that were just timed (outside every timed region)
    parity_checked, parity_note = 0, "skipped (--no-parity)"
    if not args.no_parity:
        parity_checked, parity_note = check_parity(args, p, stream, pts, d_out, n, k, rank, world, local, step_device)
        barrier()

    if world > 1:
        # (b') N > 1, one context per rank: the copies run on two streams of their own, double-buffered, so that the
        # upload of step s + 1 and the download of step s - 1 overlap the exchange / grid / search of step s; every step
        # still moves its 16 n bytes up from pinned memory and its 68 n bytes down
        cs_in, cs_out = torch.cuda.Stream(device=local), torch.cuda.Stream(device=local)
        d_in2 = [torch.empty_like(d_pts) for _ in range(2)]
        d_out2 = [torch.empty_like(d_out) for _ in range(2)]
        h_out2 = [h_out, torch.empty(n * 68, dtype=torch.uint8).pin_memory()]
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_cmp = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]

        def step_pipe(s):
            b = s & 1
            with torch.cuda.stream(cs_in):
                if s >= 2:
                    cs_in.wait_event(ev_cmp[b])  # step s - 2 has consumed d_in2[b]
                d_in2[b].copy_(h_in, non_blocking=True)
                ev_in[b].record(cs_in)
            stream.wait_event(ev_in[b])
            if s >= 2:
                stream.wait_event(ev_out[b])  # the records of step s - 2 have left d_out2[b]
            shard_step(d_in2[b], d_out2[b])
            ev_cmp[b].record(stream)
            with torch.cuda.stream(cs_out):
                cs_out.wait_event(ev_cmp[b])
                h_out2[b].copy_(d_out2[b], non_blocking=True)
                ev_out[b].record(cs_out)

        for s in range(4):
            step_pipe(s)
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for s in range(args.steps):
            step_pipe(s + 4)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        e2e_mode = ("one context per rank; upload and download on their own streams, double-buffered (upload of step s + 1 and download of "
                    "step s - 1 overlap the halo exchange + grid + search of step s); each step = H2D cloud + exchange + grid + search + D2H records")
        barrier()
        del d_in2, d_out2, h_out2, ev_in, ev_cmp, ev_out, cs_in, cs_out

    t = torch.tensor([dev_ms, e2e_s * 1e3, kern_ms, float(parity_checked)], dtype=torch.float64, device=local)
    if world > 1:
        # parity counts add up over the ranks; the timings take the slowest rank
        cnt = t[3:].clone()
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        parity_checked = int(cnt.item())
    t = t[:3]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kern_ms = t.tolist()
    total_pts = n * world * args.steps
    value = total_pts / (dev_ms * 1e-3)
    peak, peak_src = peaks()
    kern_avg_s = kern_ms * 1e-3 / args.steps
    achieved = ALGO_BYTES_PER_POINT * n / kern_avg_s / 1e9
    traffic, warp_inst = ncu_capture(n)
    sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
    issue_frac = (warp_inst / (kern_avg_s * 148 * 4 * sm_hz)) if warp_inst else None
    line = {
        "metric": "kNN+PCA feature points/sec (k=%d)" % k, "value": value, "unit": "points/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(n, k, world, args.rank_clouds, halo, args.halo_impl),
        "clocks": clocks,
        "e2e": {"value": total_pts / (e2e_ms * 1e-3), "unit": "points/s", "h2d_bytes_per_step": 16 * n, "d2h_bytes_per_step": 68 * n,
                "mode": e2e_mode, "sequential_value": total_pts / e2e_seq_s, "host_binding_rank0": numa_note},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": "knn_tile6_kernel + search_pca_worklist_kernel<0> (one stage: the tile kernel serves ~90 % of the queries, the work-list kernel the rest)",
                     "bytes_per_point": ALGO_BYTES_PER_POINT, "peak_source": peak_src, "kernel_ms": kern_avg_s * 1e3,
                     "traffic_source": (f"profiles/r02_search_ncu_{source_id()}.txt (ncu --set full of this build)" if traffic else
                                        f"none: no capture of source id {source_id()} committed"),
                     "issue_slot_frac": issue_frac,
                     "note": "k-NN selection is issue-bound, not HBM-bound: issue_slot_frac = warp instructions of the stage (ncu, same build) / "
                             "(kernel_ms x 148 SMs x 4 schedulers x SM clock) is the roofline that moves with the work done here"},
        "parity_checked": parity_checked, "parity": parity_note,
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import oracle_py as O

        threads = host_threads()
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
    # Orderly teardown, so that the process can simply return (no os._exit): EVERY torch object that was used on the
    # library's stream goes before the stream does — device tensors, the pinned host tensors (torch's host allocator
    # records an event on each stream a block was used on when the block is freed), the events and the stream wrapper.
    p.synchronize()
    keep.clear()
    if world > 1:
        dist.barrier()  # nobody unmaps while a neighbour may still store into the mapping
        if hasattr(ex, "close"):
            ex.close()
    ex = None
    del d_pts, d_out, d_kth, flush, d_gid, h_in, h_out, ev, t
    if world == 1:
        del h_outs
    import gc

    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    del stream
    gc.collect()
    p.close()
    sys.stdout.flush()


if __name__ == "__main__":
    # stdout carries the ONE JSON line and nothing else: libraries that write to file descriptor 1 (NCCL prints
    # its version banner there when NCCL_DEBUG is set on the box) are sent to stderr for the whole run
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    main()
