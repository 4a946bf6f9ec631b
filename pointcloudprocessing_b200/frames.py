"""Frame batches (BASELINE configs[1] and configs[4]): per VLP-16 frame, ring projection
(`RingProjector::projectPointCloud`, src/ringProjection.cpp:24-81) + k-NN PCA features
(the pcaAnalysis path with the lidarOdometry-sized k = 10), many frames in flight.

Frames are independent units: no halo, no collective.  Two drivers:
`FrameBatcher` (the fast one) lays `batch` frames end to end and runs them as ONE segmented grid
(`pcp_set_frames`: cell key = frame ‖ cell, searches never leave their frame) — a handful of launches per
batch instead of a dozen per frame; `FramePipeline` runs one frame per call.  One process drives `n_ctx` contexts
of the C-ABI, one host thread each, with page-locked buffers, so the frames of different contexts
overlap on the GPU (upload, grid build, search, result copy) and on the host (launch latency); across GPUs the frames are
split into contiguous blocks (`sharding.frame_shard`).  No torch here: numpy + ctypes over libpcp.so.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import List, Optional, Sequence

import numpy as np

from . import capi


class FramePipeline:
    def __init__(self, device: int = 0, n_ctx: int = 4, max_points: int = 32768, calib: Optional[capi.RingCalib] = None):
        self.calib = calib or capi.ring_calib_default()
        self.cells = self.calib.n_scan * self.calib.horizon_scan
        self.ctx = [capi.Pcp(device) for _ in range(n_ctx)]
        self.max_points = max_points
        self.slots = []
        for c in self.ctx:
            c.set_host_async(True)
            self.slots.append({
                "pts": capi.PinnedArray((max_points, 4), np.float32),
                "feat": capi.PinnedArray((max_points,), capi.FEATURE_DTYPE),
                "range": capi.PinnedArray((self.cells,), np.float32),
                "index": capi.PinnedArray((self.cells,), np.int32),
                "long": capi.PinnedArray((1,), np.int32),
                "frame": -1, "n": 0,
            })

    def close(self):
        for c in self.ctx:
            c.close()
        self.ctx, self.slots = [], []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _collect(self, i: int, out: list, copy: bool):
        s = self.slots[i]
        if s["frame"] < 0:
            return
        self.ctx[i].synchronize()
        n = s["n"]
        if copy:
            out[s["frame"]] = {
                "features": s["feat"].array[:n].copy(),
                "range": s["range"].array.reshape(self.calib.n_scan, self.calib.horizon_scan).copy(),
                "index": s["index"].array.reshape(self.calib.n_scan, self.calib.horizon_scan).copy(),
                "long_range": int(s["long"].array[0]),
            }
        s["frame"] = -1

    def run(self, frames: Sequence[np.ndarray], k: int = 10, cell: float = 0.0, ring: bool = True, keep: bool = True) -> List[dict]:
        """frames[f] = (n_f, 4) float32 x,y,z,intensity in firing order.  Returns one dict per frame
        (features = n_f pcaFeature records, range / index = the 16x1800 images, long_range) — or, with
        keep=False, only runs the work (throughput measurements)."""
        out: List[Optional[dict]] = [None] * len(frames)
        nc = len(self.ctx)

        def worker(i: int):
            # context i serves frames i, i + nc, ...; each context has its own host thread (the C-ABI
            # calls release the GIL), so the per-frame launch latency and the grid build's wait for the
            # bounding box overlap across contexts
            c, s = self.ctx[i], self.slots[i]
            for f in range(i, len(frames), nc):
                fr = frames[f]
                n = fr.shape[0]
                if n > self.max_points:
                    raise ValueError(f"frame {f} has {n} points; pipeline was sized for {self.max_points}")
                s["pts"].array[:n] = fr
                s["frame"], s["n"] = f, n
                c.set_cloud_host_ptr(s["pts"].ptr, n, 16)
                if ring:
                    c.ring_project_host_ptr(self.calib, s["range"].ptr, s["index"].ptr, s["long"].ptr)
                c.build_grid(cell, k)
                c.knn_pca(k, capi.CENTER_QUERY, out_host_ptr=s["feat"].ptr)
                self._collect(i, out, keep)

        if nc == 1:
            worker(0)
        else:
            threads = [threading.Thread(target=self._guard, args=(worker, i)) for i in range(nc)]
            self._errors = []
            for t in threads:
                t.start()
            for t in threads:
                t.join()
            if self._errors:
                raise self._errors[0]
        return out

    def _guard(self, fn, i):
        try:
            fn(i)
        except BaseException as e:  # re-raised by run() in the calling thread
            self._errors.append(e)


class FrameBatcher:
    """Many frames per launch ("many frames per launch (segmented grid)").

    `batch` frames are copied back to back into one page-locked buffer and handed over with
    `pcp_set_cloud_host` + `pcp_set_frames`; ring projection (`pcp_ring_project` with the frame offsets),
    grid build and k-NN PCA then run once per batch.  `n_ctx` contexts, one host thread each, keep one batch
    uploading / computing while another one's records travel back.  Per frame the results are bit-identical
    to `FramePipeline` (tests/test_gpu_frames.py)."""

    def __init__(self, device: int = 0, n_ctx: int = 2, batch: int = 64, max_points: int = 32768,
                 calib: Optional[capi.RingCalib] = None, segment: bool = False, odom: bool = False):
        self.calib = calib or capi.ring_calib_default()
        self.segment = segment  # also run groundRemoval + cloudSegmentation (RingProjector::fullSegmentation)
        self.odom = odom        # also lidarOdometryClass::getRangeAndptIdImage + calculateSmoothness per frame
        self.odom_prm = capi.odom_params_default()
        self.odom_cells = self.odom_prm.n_rows * self.odom_prm.n_cols
        self.seg_prm = capi.ring_seg_params_default()
        self.cells = self.calib.n_scan * self.calib.horizon_scan
        self.batch, self.max_points = batch, max_points
        self.ctx = [capi.Pcp(device) for _ in range(n_ctx)]
        self.slots = []
        cap = batch * max_points
        for c in self.ctx:
            c.set_host_async(True)
            self.slots.append({
                "pts": capi.PinnedArray((cap, 4), np.float32),
                "feat": capi.PinnedArray((cap,), capi.FEATURE_DTYPE),
                "range": capi.PinnedArray((batch * self.cells,), np.float32),
                "index": capi.PinnedArray((batch * self.cells,), np.int32),
                "long": capi.PinnedArray((batch,), np.int32),
            })
            if odom:
                self.slots[-1].update({
                    "odom_range": capi.PinnedArray((batch * self.odom_cells,), np.float32),
                    "odom_type": capi.PinnedArray((batch * self.odom_cells,), np.float32),
                })
            if segment:
                self.slots[-1].update({
                    "ground": capi.PinnedArray((batch * self.cells,), np.int8),
                    "label": capi.PinnedArray((batch * self.cells,), np.int32),
                    "colmax": capi.PinnedArray((batch * self.calib.horizon_scan,), np.float32),
                    "above": capi.PinnedArray((batch * self.cells,), np.int32),
                    "above_count": capi.PinnedArray((batch,), np.int32),
                    "n_labels": capi.PinnedArray((batch,), np.int32),
                })

    def close(self):
        for c in self.ctx:
            c.close()
        self.ctx, self.slots = [], []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _run_batch(self, i: int, frames, f0: int, f1: int, k: int, cell: float, ring: bool, out, keep: bool):
        c, s = self.ctx[i], self.slots[i]
        nb = f1 - f0
        offs = np.zeros(nb + 1, np.int64)
        for j in range(nb):
            n = frames[f0 + j].shape[0]
            if n > self.max_points:
                raise ValueError(f"frame {f0 + j} has {n} points; batcher was sized for {self.max_points}")
            offs[j + 1] = offs[j] + n
            s["pts"].array[offs[j]:offs[j + 1]] = frames[f0 + j]
        n_all = int(offs[nb])
        if n_all == 0:
            return
        c.set_cloud_host_ptr(s["pts"].ptr, n_all, 16)
        c.set_frames(offs)
        if ring and self.segment:
            c.ring_segment_host_ptr(self.calib, self.seg_prm, offs, s["range"].ptr, s["index"].ptr, s["long"].ptr, s["ground"].ptr,
                                    s["label"].ptr, s["colmax"].ptr, s["above"].ptr, s["above_count"].ptr, s["n_labels"].ptr)
        elif ring:
            c.ring_project_host_ptr(self.calib, s["range"].ptr, s["index"].ptr, s["long"].ptr, frame_offsets=offs)
        c.build_grid(cell, k)
        c.knn_pca(k, capi.CENTER_QUERY, out_host_ptr=s["feat"].ptr)
        if self.odom:
            # on the k-NN grid of the batch: capped radius count (3 m, 6 hits), images, smoothness from the device images
            c._ck(c.lib.pcp_odom_range_image(c.h, C.byref(self.odom_prm), offs.ctypes.data, nb, s["odom_range"].ptr, None, None, None,
                                             None, 0))
            c._ck(c.lib.pcp_odom_smoothness(c.h, C.byref(self.odom_prm), None, 1, nb, None, s["odom_type"].ptr, 0))
        c.synchronize()
        if not keep:
            return
        ns, hs = self.calib.n_scan, self.calib.horizon_scan
        for j in range(nb):
            feat = s["feat"].array[offs[j]:offs[j + 1]].copy()
            feat["ptID"] -= np.int32(offs[j])  # frame-local, like a run on the frame alone
            out[f0 + j] = {
                "features": feat,
                "range": s["range"].array[j * self.cells:(j + 1) * self.cells].reshape(ns, hs).copy(),
                "index": s["index"].array[j * self.cells:(j + 1) * self.cells].reshape(ns, hs).copy(),
                "long_range": int(s["long"].array[j]),
            }
            if self.odom:
                oc, shp = self.odom_cells, (self.odom_prm.n_rows, self.odom_prm.n_cols)
                out[f0 + j]["odom_range"] = s["odom_range"].array[j * oc:(j + 1) * oc].reshape(shp).copy()
                out[f0 + j]["odom_type"] = s["odom_type"].array[j * oc:(j + 1) * oc].reshape(shp).copy()
            if ring and self.segment:
                na = int(s["above_count"].array[j])
                out[f0 + j].update({
                    "ground": s["ground"].array[j * self.cells:(j + 1) * self.cells].reshape(ns, hs).copy(),
                    "label": s["label"].array[j * self.cells:(j + 1) * self.cells].reshape(ns, hs).copy(),
                    "above_idx": s["above"].array[j * self.cells:j * self.cells + na].copy(),
                    "n_labels": int(s["n_labels"].array[j]),
                })

    def run(self, frames: Sequence[np.ndarray], k: int = 10, cell: float = 0.0, ring: bool = True, keep: bool = True) -> List[dict]:
        """Same contract as FramePipeline.run (plus ground / label / above_idx / n_labels with segment=True)."""
        out: List[Optional[dict]] = [None] * len(frames)
        starts = list(range(0, len(frames), self.batch))
        nc = len(self.ctx)
        errors = []

        def worker(i: int):
            try:
                for b in starts[i::nc]:
                    self._run_batch(i, frames, b, min(b + self.batch, len(frames)), k, cell, ring, out, keep)
            except BaseException as e:  # re-raised below in the calling thread
                errors.append(e)

        if nc == 1:
            worker(0)
        else:
            threads = [threading.Thread(target=worker, args=(i,)) for i in range(nc)]
            for t in threads:
                t.start()
            for t in threads:
                t.join()
        if errors:
            raise errors[0]
        return out
