"""Packet decode throughput: 128 revolutions of VLP-16 packets -> points on the device vs the CPU oracle port."""
import os, sys, time
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, os.path.join("oracle", "ref_decoder"))
from make_packets import make_packets
from pointcloudprocessing_b200 import capi
from oracle import oracle_py as oracle
import torch

n_pk = 75 * 128
pk = make_packets(1, n_pk, az0=0, az_step=40, zero_frac=0.1)
calib = capi.ring_calib_default()
t0 = time.perf_counter()
want = oracle.decode_packets(pk, np.ctypeslib.as_array(calib.vert_deg))
cpu_s = time.perf_counter() - t0
d_pk = torch.from_numpy(pk).cuda()
with capi.Pcp(0) as p:
    for it in range(3):
        got = p.decode_packets(None, want_arrays=False, packets_device=d_pk.data_ptr(), n_packets=n_pk)
    t0 = time.perf_counter()
    reps = 10
    for it in range(reps):
        got = p.decode_packets(None, want_arrays=False, packets_device=d_pk.data_ptr(), n_packets=n_pk)
    dev_s = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for it in range(reps):
        got = p.decode_packets(pk, want_arrays=False)
    e2e_s = (time.perf_counter() - t0) / reps
n = got["n_points"]
print(f"packets {n_pk} ({pk.nbytes / 1e6:.1f} MB) -> {n} points, {got['n_complete']} complete frames")
print(f"device-resident packets: {dev_s * 1e3:.3f} ms = {n / dev_s / 1e9:.2f} G points/s ({pk.nbytes / dev_s / 1e9:.1f} GB/s of packets)")
print(f"host packets (pageable H2D inside): {e2e_s * 1e3:.3f} ms = {n / e2e_s / 1e9:.2f} G points/s")
print(f"CPU oracle port, 1 thread: {cpu_s * 1e3:.1f} ms = {n / cpu_s / 1e6:.1f} M points/s")
