"""Writes tests/golden/decoder_*.npz: synthetic Velodyne packets and what the REFERENCE's own prebuilt decoder
(lib/libPacketDecoder.so through oracle/_ref/ref_decoder, built by `make -C oracle ref`) makes of them with
16db.xml.  Run in the build container (needs /root/reference); the fixtures travel, the binary does not.

  python oracle/ref_decoder/make_golden.py
"""
import os, subprocess, sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
from make_packets import make_packets, read_frames  # noqa: E402

REF = os.environ.get("PCP_REFERENCE", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref", "ref_decoder")
OUT = os.path.join(ROOT, "tests", "golden")
TMP = os.path.join(ROOT, "build")


def run(name: str, pk: np.ndarray):
    os.makedirs(TMP, exist_ok=True)
    pk.tofile(os.path.join(TMP, "pk.bin"))
    subprocess.run([BIN, os.path.join(REF, "16db.xml"), os.path.join(TMP, "pk.bin"), os.path.join(TMP, "fr.bin")], check=True)
    fr = read_frames(os.path.join(TMP, "fr.bin"))
    offs = np.concatenate([[0], np.cumsum([f["x"].shape[0] for f in fr])]).astype(np.int64)
    cat = {k: np.concatenate([f[k] for f in fr]) if fr else np.zeros(0) for k in ("x", "y", "z", "distance", "intensity", "laser_id", "azimuth", "ms")}
    np.savez_compressed(os.path.join(OUT, f"decoder_{name}.npz"), packets=pk, frame_offsets=offs, **cat)
    print(name, "packets", pk.shape[0], "complete frames", len(fr), "points", int(offs[-1]))


if __name__ == "__main__":
    # regular sweep, 30 blocks per revolution, first frame partial
    run("sweep", make_packets(11, 14, az0=35000, az_step=1200, zero_frac=0.1))
    # irregular steps, repeated azimuths (no split on equality), many missing returns, starts at azimuth 0
    pk = make_packets(12, 10, az0=0, az_step=1500, zero_frac=0.6)
    blocks = pk[:, :1200].reshape(-1, 100)
    az = (blocks[:, 2].astype(np.int64) | (blocks[:, 3].astype(np.int64) << 8))
    az[5] = az[4]                 # equal azimuth: same frame
    az[17] = 35999                # just below the wrap
    az[40] = 36000 if az[39] < 36000 else az[40]
    blocks[:, 2] = az & 0xFF
    blocks[:, 3] = az >> 8
    pk[:, :1200] = blocks.reshape(-1, 1200)
    run("ragged", pk)
    # upper-bank blocks (0xDDFF: laser ids 32..63) mixed in
    pk = make_packets(13, 8, az0=100, az_step=2000, zero_frac=0.2)
    blocks = pk[:, :1200].reshape(-1, 100)
    blocks[1::3, 0:2] = (0xFF, 0xDD)
    pk[:, :1200] = blocks.reshape(-1, 1200)
    run("upper_bank", pk)
