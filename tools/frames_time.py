"""Throughput of the frame pipeline (configs[4] shape: VLP-16 frames, ring projection + k-NN PCA k = 10)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import frames, synth
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 400
distinct = [synth.vlp16_frame(300 + f, pose_jitter=0.5) for f in range(16)]
batch = [distinct[f % 16] for f in range(nf)]
pts = sum(b.shape[0] for b in batch)
for n_ctx in (1, 2, 4, 8):
    with frames.FramePipeline(0, n_ctx=n_ctx) as pipe:
        pipe.run(batch[:32], k=10, keep=False)
        t0 = time.perf_counter()
        pipe.run(batch, k=10, keep=False)
        dt = time.perf_counter() - t0
    print(f"n_ctx={n_ctx} frames={nf} points={pts} {nf/dt:.0f} frames/s {pts/dt/1e6:.1f} M points/s", flush=True)
for n_ctx, bsz in ((1, 32), (2, 32), (2, 64), (3, 64), (2, 128), (3, 128)):
    if nf < bsz * n_ctx:
        continue
    for cell in (0.0, 0.5):
        with frames.FrameBatcher(0, n_ctx=n_ctx, batch=bsz) as bat:
            bat.run(batch[: bsz * n_ctx], k=10, cell=cell, keep=False)
            t0 = time.perf_counter()
            bat.run(batch, k=10, cell=cell, keep=False)
            dt = time.perf_counter() - t0
        print(f"batched n_ctx={n_ctx} batch={bsz} cell={cell} frames={nf} {nf/dt:.0f} frames/s {pts/dt/1e6:.1f} M points/s", flush=True)
