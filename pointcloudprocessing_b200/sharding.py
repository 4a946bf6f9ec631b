"""One process per GPU: spatial-slab sharding with a halo exchange, and frame sharding.

The per-point work of the hot path is independent given read access to every neighbour within the
search radius (or the k-th-neighbour distance), so a big cloud shards into slabs along its longest
axis with ONE exchange step — each rank sends its neighbours the points within `h` of the shared
face — and frame batches shard with no exchange at all.  Global indices travel with the
halo points, so tie-breaking, ptID and therefore every output equal the unsharded run.

torch.distributed is the plumbing (NCCL over NVLink on the GPUs, gloo in the CPU tests); the halo
payload is ~1 % of a shard, so latency, not bandwidth, is what matters.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def frame_shard(n_frames: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block of frames [begin, end) owned by `rank` (config 5: no collective needed)."""
    base, rem = divmod(n_frames, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def equal_count_splitters(coord: torch.Tensor, lo: float, hi: float, world: int, bins: int = 4096,
                          group=None) -> List[float]:
    """Slab boundaries along one axis holding equal point counts: histogram of the local points,
    all-reduce (a few KB), then the quantiles.  Returns world+1 ascending edges."""
    hist = torch.histc(coord.float(), bins=bins, min=lo, max=hi)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(hist, group=group)
    cdf = torch.cumsum(hist, 0)
    total = float(cdf[-1])
    edges = [lo]
    width = (hi - lo) / bins
    for r in range(1, world):
        b = int(torch.searchsorted(cdf, torch.tensor(total * r / world, device=cdf.device, dtype=cdf.dtype)))
        edges.append(lo + width * min(b + 1, bins))
    edges.append(hi)
    return edges


def select_halo(pts: torch.Tensor, axis: int, lo: float, hi: float, h: float):
    """Masks of own points lying within h of the low / high face of the slab [lo, hi)."""
    c = pts[:, axis]
    return c < lo + h, c >= hi - h


def exchange_halo(pts: torch.Tensor, gid: torch.Tensor, axis: int, lo: float, hi: float, h: float,
                  rank: int, world: int, group=None):
    """Sends the strips within `h` of each shared face to the slab neighbours and receives theirs.

    pts: (n, 4) float32 own points; gid: (n,) int32 global indices.  Returns (pts_all, gid_all, n_own)
    with the own points first and the received halo after them — the layout
    pcp_set_cloud_device_sharded expects."""
    n_own = pts.shape[0]
    if world == 1:
        return pts, gid, n_own
    to_left, to_right = select_halo(pts, axis, lo, hi, h)
    send = {}
    if rank > 0:
        send[rank - 1] = (pts[to_left].contiguous(), gid[to_left].contiguous())
    if rank < world - 1:
        send[rank + 1] = (pts[to_right].contiguous(), gid[to_right].contiguous())
    # counts: one small all-gather (each rank: how many points go left / right)
    cnt = torch.zeros(2, dtype=torch.int64, device=pts.device)
    if rank > 0:
        cnt[0] = send[rank - 1][0].shape[0]
    if rank < world - 1:
        cnt[1] = send[rank + 1][0].shape[0]
    allcnt = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(allcnt, cnt, group=group)
    ops, recv = [], {}
    for nb in (rank - 1, rank + 1):
        if nb < 0 or nb >= world:
            continue
        n_in = int(allcnt[nb][1 if nb < rank else 0])  # the left neighbour sends its "right" strip
        rp = torch.empty((n_in, 4), dtype=pts.dtype, device=pts.device)
        rg = torch.empty((n_in,), dtype=gid.dtype, device=pts.device)
        recv[nb] = (rp, rg)
        sp, sg = send[nb]
        ops += [dist.P2POp(dist.isend, sp, nb, group), dist.P2POp(dist.isend, sg, nb, group),
                dist.P2POp(dist.irecv, rp, nb, group), dist.P2POp(dist.irecv, rg, nb, group)]
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    parts_p = [pts] + [recv[nb][0] for nb in sorted(recv)]
    parts_g = [gid] + [recv[nb][1] for nb in sorted(recv)]
    return torch.cat(parts_p).contiguous(), torch.cat(parts_g).contiguous(), n_own


def knn_halo_violations(own_pts: torch.Tensor, kth_d2: torch.Tensor, axis: int, lo: float, hi: float, h: float,
                        rank: int, world: int) -> int:
    """k-NN needs a data-dependent halo: a query is proven correct when its k-th neighbour distance
    does not reach past the covered region [lo − h, hi + h) (faces at the ends of the cloud are
    open).  Returns how many own queries are NOT proven; the caller widens h and repeats if > 0."""
    c = own_pts[:, axis]
    inf = torch.full_like(c, float("inf"))
    d_lo = (c - (lo - h)) if rank > 0 else inf
    d_hi = ((hi + h) - c) if rank < world - 1 else inf
    reach = torch.minimum(d_lo, d_hi).clamp_min(0)
    return int((kth_d2 > reach * reach).sum())


def check_halo_reach(edges: List[float], h: float) -> None:
    """Halos come from the two ADJACENT slabs only: a halo of width h is complete only if no slab is narrower than h
    (else points two slabs away fall inside [lo - h, hi + h) and are never received, while knn_halo_violations and
    radius_halo_width still count that range as covered).  Raises ValueError; call it once after the splitters are
    known (edges = the world + 1 slab boundaries; infinite outer faces are fine)."""
    widths = [b - a for a, b in zip(edges[:-1], edges[1:])]
    inner = widths[1:-1] if len(widths) > 2 else []
    # the end slabs only have to supply their inner neighbour; an inner slab lies between two halos
    narrow = [w for w in (inner + widths[:1] + widths[-1:]) if w < h]
    if len(widths) > 1 and narrow:
        raise ValueError(f"halo width {h} exceeds the width of a slab ({min(narrow):.6g}): fewer ranks, or an exchange with further neighbours, is needed")


def radius_halo_width(radius: float) -> float:
    """Radius search: a halo of r (plus rounding slack) is exact."""
    return float(radius) * (1.0 + 1e-5) + 1e-6


class SlabExchange:
    """Fixed-capacity halo exchange of one slab rank, with NO host synchronisation per step.

    Persistent device buffers hold `own ‖ halo-from-low-neighbour ‖ halo-from-high-neighbour`
    (n_own + 2·cap points; unused entries are non-finite padding that the grid ignores).  A step is:
    `pcp_halo_pack` (one classification kernel into the two send buffers) → one grouped NCCL
    send/recv per neighbour straight into the halo regions → `pcp_set_cloud_device_sharded`.  Sizes
    never depend on data, so nothing has to reach the host before the grid build is enqueued; the true
    strip sizes stay in `counts` (device) and `check()` verifies them once, outside the hot loop."""

    INT_MAX = 0x7FFFFFFF

    def __init__(self, pcp, n_own: int, cap: int, device, rank: int, world: int, axis: int, lo: float, hi: float, h: float,
                 group=None):
        self.pcp, self.n_own, self.cap, self.rank, self.world = pcp, n_own, cap, rank, world
        self.axis, self.lo, self.hi, self.h, self.group = axis, float(lo), float(hi), float(h), group
        n_all = n_own + 2 * cap
        self.buf_pts = torch.full((n_all, 4), float("nan"), dtype=torch.float32, device=device)
        self.buf_pts[:, 3] = 0
        self.buf_gid = torch.full((n_all,), self.INT_MAX, dtype=torch.int32, device=device)
        self.send_pts = [torch.empty((cap, 4), dtype=torch.float32, device=device) for _ in range(2)]
        self.send_gid = [torch.empty((cap,), dtype=torch.int32, device=device) for _ in range(2)]
        self.counts = torch.zeros(2, dtype=torch.int32, device=device)

    @property
    def own_pts(self) -> torch.Tensor:
        return self.buf_pts[: self.n_own]

    @property
    def own_gid(self) -> torch.Tensor:
        return self.buf_gid[: self.n_own]

    def halo_region(self, side: int):
        a = self.n_own + side * self.cap
        return self.buf_pts[a: a + self.cap], self.buf_gid[a: a + self.cap]

    def pack(self):
        self.pcp.halo_pack(self.own_pts.data_ptr(), self.own_gid.data_ptr(), self.n_own, self.axis, self.lo, self.hi, self.h,
                           self.cap, self.send_pts[0].data_ptr(), self.send_gid[0].data_ptr(), self.send_pts[1].data_ptr(),
                           self.send_gid[1].data_ptr(), self.counts.data_ptr())

    def step(self):
        """Enqueue pack + exchange + hand-over on the CURRENT torch stream (use the library's stream)."""
        self.pack()
        ops = []
        for side, nb in ((0, self.rank - 1), (1, self.rank + 1)):
            if nb < 0 or nb >= self.world:
                continue
            rp, rg = self.halo_region(side)
            ops += [dist.P2POp(dist.isend, self.send_pts[side], nb, self.group), dist.P2POp(dist.isend, self.send_gid[side], nb, self.group),
                    dist.P2POp(dist.irecv, rp, nb, self.group), dist.P2POp(dist.irecv, rg, nb, self.group)]
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        self.pcp.set_cloud_device(self.buf_pts.data_ptr(), self.buf_pts.shape[0], self.buf_gid.data_ptr(), self.n_own,
                                  keep=(self.buf_pts, self.buf_gid))

    def check(self):
        """Host check (synchronises): did both strips fit their buffers?"""
        c = self.counts.tolist()
        if max(c) > self.cap:
            raise RuntimeError(f"rank {self.rank}: halo strips of {c} points exceed the capacity {self.cap}")
        return c


class _DevView:
    """Zero-copy torch view of raw device memory (the __cuda_array_interface__ protocol)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class SlabExchangePeer:
    """Halo exchange by peer stores: the classification kernel of `pcp_halo_push` writes a rank's two
    strips STRAIGHT INTO its slab neighbours' halo regions (their memory mapped over CUDA IPC, NVLink
    peer writes) and raises a flag there; the receiver's stream only waits for its two flags
    (`pcp_halo_wait`).  No message, no NCCL kernel, no host synchronisation per step.

    Every rank owns ONE peer-mappable allocation holding two `own ‖ halo ‖ halo` buffers (even / odd
    epochs, so a neighbour's strip for step t+1 never lands in memory step t still reads), their
    global-id arrays, four flags, an error word and the two strip counters.  All ranks use the same
    `cap`.  Set-up: `meta = alloc()`, exchange the metas of all ranks by any transport (here
    `all_gather_object`), `connect(metas)`; then `step()` per cloud as with SlabExchange."""

    INT_MAX = 0x7FFFFFFF

    def __init__(self, pcp, n_own: int, cap: int, device, rank: int, world: int, axis: int, lo: float, hi: float, h: float,
                 timeout_ms: int = 2000):
        import os
        self.pcp, self.n_own, self.cap, self.rank, self.world = pcp, n_own, cap, rank, world
        self.axis, self.lo, self.hi, self.h, self.timeout_ms = axis, float(lo), float(hi), float(h), timeout_ms
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self.n_all = n_own + 2 * cap
        lay = self.layout(n_own, cap)
        self.base, self.handle = pcp.peer_alloc(lay["bytes"])
        self.lay, self.pid, self.epoch = lay, os.getpid(), 0
        self.peers = {}      # neighbour rank -> (base pointer, layout, opened over IPC?)
        raw = torch.as_tensor(_DevView(self.base, (lay["bytes"] // 4,), "<u4"), device=self.device)
        raw.zero_()
        for b in (0, 1):
            self.buf_pts(b).fill_(float("nan"))
            self.buf_pts(b)[:, 3] = 0
            self.buf_gid(b).fill_(self.INT_MAX)
        # strips towards a face without a neighbour go nowhere useful: local scratch
        self.scratch_pts = torch.empty((cap, 4), dtype=torch.float32, device=self.device)
        self.scratch_gid = torch.empty((cap,), dtype=torch.int32, device=self.device)
        torch.cuda.synchronize(self.device)

    @staticmethod
    def layout(n_own: int, cap: int) -> dict:
        n_all = n_own + 2 * cap
        pts = [0, n_all * 16]
        gid = [2 * n_all * 16, 2 * n_all * 16 + n_all * 4]
        flags = (2 * n_all * 16 + 2 * n_all * 4 + 255) // 256 * 256
        return {"n_own": n_own, "cap": cap, "n_all": n_all, "pts": pts, "gid": gid, "flags": flags, "err": flags + 16,
                "counts": flags + 32, "bytes": flags + 64}

    # ---- views / addresses of the own allocation
    def buf_pts(self, b: int) -> torch.Tensor:
        return torch.as_tensor(_DevView(self.base + self.lay["pts"][b], (self.n_all, 4), "<f4"), device=self.device)

    def buf_gid(self, b: int) -> torch.Tensor:
        return torch.as_tensor(_DevView(self.base + self.lay["gid"][b], (self.n_all,), "<i4"), device=self.device)

    @property
    def own_pts(self) -> torch.Tensor:
        """Own region of the buffer the NEXT step() uses."""
        return self.buf_pts(self.epoch & 1)[: self.n_own]

    def set_own(self, pts: torch.Tensor, gid: torch.Tensor):
        for b in (0, 1):
            self.buf_pts(b)[: self.n_own].copy_(pts)
            self.buf_gid(b)[: self.n_own].copy_(gid)

    @staticmethod
    def _halo_addr(base: int, lay: dict, b: int, side: int):
        o = lay["n_own"] + side * lay["cap"]
        return base + lay["pts"][b] + o * 16, base + lay["gid"][b] + o * 4, base + lay["flags"] + (b * 2 + side) * 4

    # ---- set-up
    def meta(self) -> dict:
        return {"rank": self.rank, "pid": self.pid, "ptr": self.base, "handle": self.handle, "n_own": self.n_own, "cap": self.cap}

    def connect(self, metas):
        for m in metas:
            if m["cap"] != self.cap:
                raise ValueError("SlabExchangePeer: every rank must use the same halo capacity")
        for nb in (self.rank - 1, self.rank + 1):
            if nb < 0 or nb >= self.world:
                continue
            m = next(x for x in metas if x["rank"] == nb)
            same = m["pid"] == self.pid  # two ranks emulated in one process: the pointer is valid as it is
            base = m["ptr"] if same else self.pcp.peer_open(m["handle"])
            self.peers[nb] = (base, self.layout(m["n_own"], m["cap"]), not same)

    def connect_group(self, group=None):
        metas = [None] * self.world
        dist.all_gather_object(metas, self.meta(), group=group)
        self.connect(metas)
        dist.barrier(group=group)

    # ---- per cloud
    def step(self):
        """Enqueue push + wait + hand-over on the library's stream."""
        b = self.epoch & 1
        self.epoch += 1
        e = self.epoch
        tgt = []
        for nb, side_there in ((self.rank - 1, 1), (self.rank + 1, 0)):
            if nb in self.peers:
                base, lay, _ = self.peers[nb]
                tgt.append(self._halo_addr(base, lay, b, side_there))
            else:
                tgt.append((self.scratch_pts.data_ptr(), self.scratch_gid.data_ptr(), None))
        own_p = self.base + self.lay["pts"][b]
        own_g = self.base + self.lay["gid"][b]
        self.pcp.halo_push(own_p, own_g, self.n_own, self.axis, self.lo, self.hi, self.h, self.cap, tgt[0][0], tgt[0][1], tgt[1][0],
                           tgt[1][1], self.base + self.lay["counts"], tgt[0][2], tgt[1][2], e)
        fa = self._halo_addr(self.base, self.lay, b, 0)[2] if (self.rank - 1) in self.peers else None
        fb = self._halo_addr(self.base, self.lay, b, 1)[2] if (self.rank + 1) in self.peers else None
        self.pcp.halo_wait(fa, fb, e, self.timeout_ms, self.base + self.lay["err"])
        self.pcp.set_cloud_device(own_p, self.n_all, own_g, self.n_own)

    def halo_region(self, side: int, b: int = None):
        b = ((self.epoch - 1) & 1) if b is None else b
        a = self.n_own + side * self.cap
        return self.buf_pts(b)[a: a + self.cap], self.buf_gid(b)[a: a + self.cap]

    def check(self):
        """Host check (synchronises): both strips fitted, no wait timed out."""
        self.pcp.synchronize()
        raw = torch.as_tensor(_DevView(self.base + self.lay["err"], (8,), "<u4"), device=self.device).tolist()
        if raw[0]:
            raise RuntimeError(f"rank {self.rank}: a halo wait timed out (a neighbour never delivered its strip)")
        c = [raw[4], raw[5]]
        if max(c) > self.cap:
            raise RuntimeError(f"rank {self.rank}: halo strips of {c} points exceed the capacity {self.cap}")
        return c

    def close(self):
        for nb, (base, _, opened) in self.peers.items():
            if opened:
                self.pcp.peer_close(base)
        self.peers = {}
        if self.base:
            self.pcp.peer_free(self.base)
            self.base = 0
