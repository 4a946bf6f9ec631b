"""Synthetic Velodyne VLP-16 data packets (1206 B: 12 x {0xEEFF, azimuth 0.01 deg, 32 x {u16 distance, u8 intensity}} +
u32 timestamp + 2 B; include/PacketDecoder.h:21-49) for the decode oracle and its golden fixtures."""
import numpy as np


def make_packets(seed: int, n_packets: int, az0: int = 35000, az_step: int = 40, zero_frac: float = 0.1) -> np.ndarray:
    """-> (n_packets, 1206) uint8.  Azimuth advances az_step (0.01 deg) per block and wraps at 36000, so a
    revolution is 36000 / az_step blocks; a share of the returns has distance 0 (no return)."""
    rng = np.random.default_rng(seed)
    pk = np.zeros((n_packets, 1206), np.uint8)
    az = az0
    for p in range(n_packets):
        for b in range(12):
            o = b * 100
            pk[p, o:o + 2] = (0xFF, 0xEE)
            pk[p, o + 2:o + 4] = np.frombuffer(np.uint16(az % 36000).tobytes(), np.uint8)
            dist = rng.integers(200, 50000, 32).astype(np.uint16)
            dist[rng.random(32) < zero_frac] = 0
            inten = rng.integers(0, 256, 32).astype(np.uint8)
            ret = np.zeros((32, 3), np.uint8)
            ret[:, 0:2] = np.frombuffer(dist.tobytes(), np.uint8).reshape(32, 2)
            ret[:, 2] = inten
            pk[p, o + 4:o + 100] = ret.reshape(-1)
            az += az_step
        pk[p, 1200:1204] = np.frombuffer(np.uint32(1000 + 553 * p).tobytes(), np.uint8)
        pk[p, 1204:1206] = (0x37, 0x22)
    return pk


def read_frames(path: str):
    """Frames dumped by ref_decoder: list of dicts of arrays."""
    raw = open(path, "rb").read()
    nf = int(np.frombuffer(raw, np.uint64, 1, 0)[0])
    off, out = 8, []
    for _ in range(nf):
        n = int(np.frombuffer(raw, np.uint64, 1, off)[0]); off += 8
        fr = {}
        for name, dt in (("x", np.float64), ("y", np.float64), ("z", np.float64), ("distance", np.float64), ("intensity", np.uint8),
                         ("laser_id", np.uint8), ("azimuth", np.uint16), ("ms", np.uint32)):
            fr[name] = np.frombuffer(raw, dt, n, off).copy(); off += n * np.dtype(dt).itemsize
        out.append(fr)
    return out
