"""Synthetic clouds of the shapes BASELINE.json names.  numpy PCG64 streams keyed
by seed, so every machine generates the same bytes.  All outputs are (n, 4) float32 xyzi."""
from __future__ import annotations

import numpy as np

# VLP-16 beam table = 16db.xml vertCorrection_ of lasers 0..15 (16db.xml:813-1249)
VLP16_VERT_DEG = np.array([-15, 1, -13, 3, -11, 5, -9, 7, -7, 9, -5, 11, -3, 13, -1, 15], np.float64)


def urban(seed: int, n: int, extent: float = 200.0, order: str = "random") -> np.ndarray:
    """50 % ground (σz 0.02), 25 % façades on a 20 m lattice (0-15 m high, σ 0.02), 15 % vegetation
    blobs (σ 1.5 m), 10 % uniform clutter 0-10 m.  order: "random" permutation, "scan" (by azimuth) or
    anything else = as generated (class by class)."""
    rng = np.random.default_rng(seed)
    n_g = n // 2
    n_f = n // 4
    n_v = (n * 15) // 100
    n_c = n - n_g - n_f - n_v
    parts = []
    g = np.empty((n_g, 3))
    g[:, 0:2] = rng.uniform(0, extent, (n_g, 2))
    g[:, 2] = rng.normal(0, 0.02, n_g)
    parts.append(g)
    f = np.empty((n_f, 3))
    nl = max(int(extent // 20), 1)
    along = rng.uniform(0, extent, n_f)
    plane = (rng.integers(0, nl, n_f) + 0.5) * 20.0 + rng.normal(0, 0.02, n_f)
    axis = rng.integers(0, 2, n_f).astype(bool)
    f[:, 0] = np.where(axis, along, plane)
    f[:, 1] = np.where(axis, plane, along)
    f[:, 2] = rng.uniform(0, 15, n_f)
    parts.append(f)
    nb = max(int(extent * extent / 800), 1)
    centres = np.column_stack([rng.uniform(0, extent, nb), rng.uniform(0, extent, nb), rng.uniform(2, 6, nb)])
    v = centres[rng.integers(0, nb, n_v)] + rng.normal(0, 1.5, (n_v, 3))
    parts.append(v)
    c = np.column_stack([rng.uniform(0, extent, n_c), rng.uniform(0, extent, n_c), rng.uniform(0, 10, n_c)])
    parts.append(c)
    xyz = np.concatenate(parts).astype(np.float32)
    inten = rng.uniform(0, 255, n).astype(np.float32)
    out = np.column_stack([xyz, inten]).astype(np.float32)
    if order == "random":
        out = out[rng.permutation(n)]
    elif order == "scan":
        ang = np.arctan2(out[:, 1] - extent / 2, out[:, 0] - extent / 2)
        out = out[np.argsort(ang, kind="stable")]
    return np.ascontiguousarray(out)


def playground(seed: int = 1, n: int = 500_000) -> np.ndarray:
    """Stand-in for the missing data/playground_subsampled.txt: 60×40 m sloped ground, boxes, poles,
    canopy blobs; scan-like order (sorted by azimuth from a corner)."""
    rng = np.random.default_rng(seed)
    n_g = int(n * 0.6)
    n_b = int(n * 0.2)
    n_p = int(n * 0.04)
    n_c = n - n_g - n_b - n_p
    g = np.empty((n_g, 3))
    g[:, 0] = rng.uniform(0, 60, n_g)
    g[:, 1] = rng.uniform(0, 40, n_g)
    g[:, 2] = 0.01 * g[:, 0] + rng.normal(0, 0.01, n_g)
    bc = np.column_stack([rng.uniform(5, 55, 6), rng.uniform(5, 35, 6)])
    bs = rng.uniform(2, 4, (6, 3))
    bi = rng.integers(0, 6, n_b)
    face = rng.integers(0, 3, n_b)
    u = rng.uniform(-0.5, 0.5, (n_b, 3))
    sign = rng.integers(0, 2, n_b) * 1.0 - 0.5
    u[np.arange(n_b), face] = sign
    b = np.empty((n_b, 3))
    b[:, 0] = bc[bi, 0] + u[:, 0] * bs[bi, 0]
    b[:, 1] = bc[bi, 1] + u[:, 1] * bs[bi, 1]
    b[:, 2] = (u[:, 2] + 0.5) * bs[bi, 2]
    pc = np.column_stack([rng.uniform(2, 58, 8), rng.uniform(2, 38, 8)])
    pi_ = rng.integers(0, 8, n_p)
    th = rng.uniform(0, 2 * np.pi, n_p)
    p = np.column_stack([pc[pi_, 0] + 0.05 * np.cos(th), pc[pi_, 1] + 0.05 * np.sin(th), rng.uniform(0, 3, n_p)])
    cc = np.column_stack([rng.uniform(5, 55, 5), rng.uniform(5, 35, 5), rng.uniform(3, 6, 5)])
    c = cc[rng.integers(0, 5, n_c)] + rng.normal(0, 1.2, (n_c, 3))
    xyz = np.concatenate([g, b, p, c]).astype(np.float32)
    out = np.column_stack([xyz, rng.uniform(0, 255, n).astype(np.float32)]).astype(np.float32)
    ang = np.arctan2(out[:, 1] + 1.0, out[:, 0] + 1.0)
    return np.ascontiguousarray(out[np.argsort(ang, kind="stable")])


def vlp16_frame(seed: int = 2, n_az: int = 1800, pose_jitter: float = 0.0) -> np.ndarray:
    """One VLP-16 revolution ray-cast into a courtyard (ground z = −1.8 m, 40×30 m walls 6 m high,
    a few boxes): 16 beams (16db.xml table) × n_az azimuths in firing order, range quantised to 2 mm
    (distLSB_ 0.2 cm), x = d·cos v·sin a, y = d·cos v·cos a, z = d·sin v (the PacketDecoder closed
    form).  Misses / > 100 m are dropped."""
    rng = np.random.default_rng(seed)
    ox, oy = rng.normal(0, pose_jitter, 2) if pose_jitter > 0 else (0.0, 0.0)
    az = np.repeat(np.arange(n_az) * (360.0 / n_az), 16)
    vert = np.tile(VLP16_VERT_DEG, n_az)
    a = np.deg2rad(az)
    v = np.deg2rad(vert)
    dx, dy, dz = np.cos(v) * np.sin(a), np.cos(v) * np.cos(a), np.sin(v)
    t = np.full(a.shape, np.inf)
    with np.errstate(divide="ignore", invalid="ignore"):
        tg = np.where(dz < 0, -1.8 / dz, np.inf)  # ground
        t = np.minimum(t, tg)
        for wall, d, o in ((20.0 - ox, dx, 0), (-20.0 - ox, dx, 0), (15.0 - oy, dy, 1), (-15.0 - oy, dy, 1)):
            tw = np.where(d * wall > 0, wall / d, np.inf)
            zhit = tw * dz
            tw = np.where((zhit > -1.8) & (zhit < 4.2), tw, np.inf)
            t = np.minimum(t, tw)
        nbox = 5
        bc = np.column_stack([rng.uniform(-15, 15, nbox), rng.uniform(-10, 10, nbox)])
        bh = rng.uniform(0.5, 1.5, nbox)
        for (cx, cy), h in zip(bc, bh):
            lo = np.array([cx - h - ox, cy - h - oy, -1.8])
            hi = np.array([cx + h - ox, cy + h - oy, -1.8 + 2 * h])
            d3 = np.stack([dx, dy, dz], 1)
            t1 = lo / d3
            t2 = hi / d3
            tn = np.nanmax(np.minimum(t1, t2), 1)
            tf = np.nanmin(np.maximum(t1, t2), 1)
            tb = np.where((tn <= tf) & (tn > 0.3), tn, np.inf)
            t = np.minimum(t, tb)
    t = t + rng.normal(0, 0.01, t.shape)
    raw = np.round(t / 0.002)
    ok = np.isfinite(t) & (t < 100.0) & (raw > 0)
    d = (raw * 0.002)[ok]
    xyz = np.column_stack([d * dx[ok], d * dy[ok], d * dz[ok]]).astype(np.float32)
    inten = rng.integers(0, 256, xyz.shape[0]).astype(np.float32)
    return np.ascontiguousarray(np.column_stack([xyz, inten]).astype(np.float32))


def vlp16_batch(n_frames: int, seed0: int = 1000):
    """Frames back to back + offsets (config 5 shape)."""
    frames = [vlp16_frame(seed0 + f, pose_jitter=0.5) for f in range(n_frames)]
    offsets = np.zeros(n_frames + 1, np.int64)
    np.cumsum([f.shape[0] for f in frames], out=offsets[1:])
    return np.ascontiguousarray(np.concatenate(frames)), offsets
