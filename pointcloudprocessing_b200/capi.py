"""ctypes binding of libpcp.so — the C-ABI declared in include/pcp.h.

This is the reference-side stub a Python caller would use; the C++ drop-in (`host/`) binds
the same symbols.  There is no fallback: a missing library raises at import of `load()`,
and every compute call raises `PcpError` when the CUDA path fails.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpcp.so")

# numpy mirror of pcaAnalysist::pcaFeature (include/pcaAnalysis.h:34-45), 68 bytes
FEATURE_DTYPE = np.dtype(
    [
        ("la", np.float32, (3,)),
        ("v1", np.float32, (3,)),
        ("v2", np.float32, (3,)),
        ("v3", np.float32, (3,)),
        ("ptID", np.int32),
        ("ptNum", np.int32),
        ("lineStruc", np.float32),
        ("planeStruc", np.float32),
        ("scatterStruc", np.float32),
    ]
)
assert FEATURE_DTYPE.itemsize == 68

CENTER_QUERY, CENTER_MEAN = 0, 1
STAGES = ("bbox", "keys", "sort", "table", "search_pca", "carry", "h2d", "d2h")
MAX_K = 64

# every symbol include/pcp.h declares (checked by tests/test_abi.py)
EXPORTS = (
    "pcp_create", "pcp_destroy", "pcp_last_error", "pcp_version", "pcp_stream", "pcp_synchronize",
    "pcp_last_timings", "pcp_launch_count", "pcp_set_cloud_host", "pcp_set_cloud_device",
    "pcp_set_cloud_device_sharded", "pcp_build_grid", "pcp_grid_info", "pcp_radius_pca", "pcp_knn_pca",
    "pcp_knn_pca_kth", "pcp_knn_neighbors", "pcp_radius_neighbors", "pcp_segment_pca", "pcp_ring_calib_default",
    "pcp_ring_calib_from_xml", "pcp_ring_project", "pcp_set_host_async", "pcp_alloc_pinned", "pcp_free_pinned", "pcp_halo_pack", "pcp_set_frames",
    "pcp_halo_push", "pcp_halo_wait", "pcp_peer_alloc", "pcp_peer_free", "pcp_peer_open", "pcp_peer_close",
    "pcp_ring_seg_params_default", "pcp_ring_segment", "pcp_decode_packets",
    "pcp_dbscan", "pcp_dbscan_d", "pcp_cluster_boxes", "pcp_mean_shift", "pcp_set_cloud_txt",
    "pcp_radius_count", "pcp_odom_params_default", "pcp_odom_range_image", "pcp_odom_smoothness",
    "pcp_bbox_pca", "pcp_carry_forward_static", "pcp_carry_forward_static_max", "pcp_set_cloud_xyz",
)


class PinnedArray:
    """numpy view over page-locked host memory from pcp_alloc_pinned (freed with the object)."""

    def __init__(self, shape, dtype):
        dtype = np.dtype(dtype)
        nbytes = int(np.prod(shape)) * dtype.itemsize
        self._lib = load()
        self._ptr = self._lib.pcp_alloc_pinned(max(nbytes, 1))
        if not self._ptr:
            raise MemoryError(f"pcp_alloc_pinned({nbytes}) failed")
        buf = (C.c_uint8 * max(nbytes, 1)).from_address(self._ptr)
        self.array = np.frombuffer(buf, dtype=np.uint8, count=nbytes).view(dtype).reshape(shape)
        self.ptr = self._ptr

    def __del__(self):
        if getattr(self, "_ptr", None):
            self.array = None
            self._lib.pcp_free_pinned(self._ptr)
            self._ptr = None


class RingCalib(C.Structure):
    _fields_ = [
        ("n_scan", C.c_int32),
        ("horizon_scan", C.c_int32),
        ("ang_res_x", C.c_float),
        ("ang_res_y", C.c_float),
        ("ang_bottom", C.c_float),
        ("vert_deg", C.c_float * 64),
        ("dist_lsb_cm", C.c_float),
        ("n_enabled", C.c_int32),
    ]


class RingSegParams(C.Structure):
    _fields_ = [
        ("ground_scan_ind", C.c_int32),
        ("sensor_mount_angle", C.c_float),
        ("segment_theta", C.c_float),
        ("segment_valid_point_num", C.c_int32),
        ("segment_valid_line_num", C.c_int32),
    ]


class OdomParams(C.Structure):
    """pcp_odom_params: the literals of src/lidarOdometry.cpp:21-187."""
    _fields_ = [
        ("n_rows", C.c_int32), ("n_cols", C.c_int32), ("vert_res_deg", C.c_double), ("hori_res_deg", C.c_double),
        ("radius", C.c_float), ("min_neighbors", C.c_int32), ("n_blocks", C.c_int32), ("half_window", C.c_int32),
        ("near_range", C.c_float), ("min_valid", C.c_int32), ("pick", C.c_int32), ("min_block_valid", C.c_int32),
    ]


class Decoded(C.Structure):
    _fields_ = [("x", C.c_void_p), ("y", C.c_void_p), ("z", C.c_void_p), ("distance", C.c_void_p), ("intensity", C.c_void_p),
                ("laser_id", C.c_void_p), ("azimuth", C.c_void_p), ("ms", C.c_void_p)]


class Obb(C.Structure):
    """pcp_obb (include/pcp.h): getMiniBoundingBox's quantities (src/pcaAnalysis.cpp:123-148)."""
    _fields_ = [("centroid", C.c_float * 3), ("eigval", C.c_float * 3), ("axes", (C.c_float * 3) * 3), ("min_pt", C.c_float * 3),
                ("max_pt", C.c_float * 3), ("mean_diag", C.c_float * 3), ("n_points", C.c_int64)]


class PcpError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"pcp status {status}: {msg}")
        self.status = status


_lib = None


def load() -> C.CDLL:
    """Loads libpcp.so (built in-tree by __graft_entry__.build()).  No fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("PCP_LIB", LIB_PATH)  # PCP_LIB: a debug build of the same library (stats counters)
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(path)
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    sig = {
        "pcp_create": (i32, [C.POINTER(vp), i32]),
        "pcp_destroy": (i32, [vp]),
        "pcp_last_error": (C.c_char_p, [vp]),
        "pcp_version": (C.c_char_p, []),
        "pcp_stream": (vp, [vp]),
        "pcp_synchronize": (i32, [vp]),
        "pcp_last_timings": (i32, [vp, vp, i32]),
        "pcp_launch_count": (i64, [vp]),
        "pcp_set_host_async": (i32, [vp, i32]),
        "pcp_alloc_pinned": (vp, [C.c_size_t]),
        "pcp_free_pinned": (None, [vp]),
        "pcp_halo_pack": (i32, [vp, vp, vp, i64, i32, f32, f32, f32, i64, vp, vp, vp, vp, vp]),
        "pcp_halo_push": (i32, [vp, vp, vp, i64, i32, f32, f32, f32, i64, vp, vp, vp, vp, vp, vp, vp, C.c_uint32]),
        "pcp_halo_wait": (i32, [vp, vp, vp, C.c_uint32, i32, vp]),
        "pcp_peer_alloc": (i32, [vp, C.c_size_t, C.POINTER(vp), vp]),
        "pcp_peer_free": (i32, [vp, vp]),
        "pcp_peer_open": (i32, [vp, vp, C.POINTER(vp)]),
        "pcp_peer_close": (i32, [vp, vp]),
        "pcp_set_cloud_host": (i32, [vp, vp, i64, i32]),
        "pcp_set_cloud_device": (i32, [vp, vp, i64]),
        "pcp_set_cloud_device_sharded": (i32, [vp, vp, vp, i64, i64]),
        "pcp_set_frames": (i32, [vp, vp, i64]),
        "pcp_build_grid": (i32, [vp, f32, i32]),
        "pcp_grid_info": (i32, [vp, vp, vp, vp]),
        "pcp_radius_pca": (i32, [vp, f32, i32, i32, i32, vp, i32]),
        "pcp_knn_pca": (i32, [vp, i32, i32, vp, vp, i32]),
        "pcp_knn_pca_kth": (i32, [vp, i32, i32, vp, vp, vp, i32]),
        "pcp_knn_neighbors": (i32, [vp, i32, vp, vp, i32]),
        "pcp_radius_neighbors": (i32, [vp, f32, vp, vp, vp, vp, i32]),
        "pcp_segment_pca": (i32, [vp, vp, vp, vp, i64, vp]),
        "pcp_bbox_pca": (i32, [vp, C.POINTER(Obb)]),
        "pcp_carry_forward_static": (i32, [vp, vp, vp, i64, i32]),
        "pcp_carry_forward_static_max": (i64, []),
        "pcp_ring_calib_default": (i32, [C.POINTER(RingCalib)]),
        "pcp_ring_calib_from_xml": (i32, [C.c_char_p, C.POINTER(RingCalib)]),
        "pcp_ring_project": (i32, [vp, C.POINTER(RingCalib), vp, i64, vp, vp, vp, i32]),
        "pcp_ring_seg_params_default": (i32, [C.POINTER(RingSegParams)]),
        "pcp_decode_packets": (i32, [vp, C.POINTER(RingCalib), vp, i64, i32, vp, i32, vp, i64, vp, vp]),
        "pcp_radius_count": (i32, [vp, f32, i32, vp, i32]),
        "pcp_dbscan": (i32, [vp, f32, i32, vp, vp, i32]),
        "pcp_dbscan_d": (i32, [vp, C.c_double, i32, vp, vp, i32]),
        "pcp_set_cloud_txt": (i32, [vp, vp, i64, i32, vp, vp, i64, i32]),
        "pcp_set_cloud_xyz": (i32, [vp, vp, i64, i32, vp, vp, i64, i32]),
        "pcp_cluster_boxes": (i32, [vp, vp, i32, i32, vp, vp, vp, vp]),
        "pcp_mean_shift": (i32, [vp, f32, i32, i32, vp, i32, vp, vp, i32]),
        "pcp_odom_params_default": (i32, [C.POINTER(OdomParams)]),
        "pcp_odom_range_image": (i32, [vp, C.POINTER(OdomParams), vp, i64, vp, vp, vp, vp, vp, i32]),
        "pcp_odom_smoothness": (i32, [vp, C.POINTER(OdomParams), vp, i32, i64, vp, vp, i32]),
        "pcp_ring_segment": (i32, [vp, C.POINTER(RingCalib), C.POINTER(RingSegParams), vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _ptr(a) -> Optional[int]:
    """Address of a numpy array, a raw int device pointer, or None."""
    if a is None:
        return None
    if isinstance(a, (int, np.integer)):
        return int(a)
    return a.ctypes.data


def ring_seg_params_default() -> "RingSegParams":
    prm = RingSegParams()
    st = load().pcp_ring_seg_params_default(C.byref(prm))
    if st != 0:
        raise PcpError(st, "pcp_ring_seg_params_default")
    return prm


def ring_calib_default() -> RingCalib:
    c = RingCalib()
    st = load().pcp_ring_calib_default(C.byref(c))
    if st:
        raise PcpError(st, "pcp_ring_calib_default")
    return c


def odom_params_default() -> OdomParams:
    prm = OdomParams()
    st = load().pcp_odom_params_default(C.byref(prm))
    if st:
        raise PcpError(st, "pcp_odom_params_default")
    return prm


def ring_calib_from_xml(path: str) -> RingCalib:
    c = RingCalib()
    st = load().pcp_ring_calib_from_xml(path.encode(), C.byref(c))
    if st:
        raise PcpError(st, f"pcp_ring_calib_from_xml({path})")
    return c


class Pcp:
    """One pcp_ctx (one CUDA stream on one GPU)."""

    def __init__(self, device: int = 0):
        self.lib = load()
        h = C.c_void_p()
        st = self.lib.pcp_create(C.byref(h), device)
        if st:
            raise PcpError(st, "pcp_create failed (no sm_100 CUDA device?)")
        self.h = h
        self.n = 0
        self.n_query = 0
        self._keep = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.pcp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _ck(self, st: int):
        if st:
            raise PcpError(st, self.lib.pcp_last_error(self.h).decode())

    # ---- plumbing
    @property
    def stream(self) -> int:
        return int(self.lib.pcp_stream(self.h) or 0)

    def synchronize(self):
        self._ck(self.lib.pcp_synchronize(self.h))

    def halo_pack(self, d_pts: int, d_gid: Optional[int], n: int, axis: int, lo_face: float, hi_face: float, h: float, cap: int,
                  d_lo_pts: int, d_lo_gid: int, d_hi_pts: int, d_hi_gid: int, d_counts: int):
        """Device pointers throughout; see pcp_halo_pack."""
        self._ck(self.lib.pcp_halo_pack(self.h, d_pts, d_gid, n, axis, lo_face, hi_face, h, cap, d_lo_pts, d_lo_gid, d_hi_pts,
                                        d_hi_gid, d_counts))

    def halo_push(self, d_pts: int, d_gid: Optional[int], n: int, axis: int, lo_face: float, hi_face: float, h: float, cap: int,
                  p_lo_pts: int, p_lo_gid: int, p_hi_pts: int, p_hi_gid: int, d_counts: int, p_lo_flag: Optional[int],
                  p_hi_flag: Optional[int], epoch: int):
        self._ck(self.lib.pcp_halo_push(self.h, d_pts, d_gid, n, axis, lo_face, hi_face, h, cap, p_lo_pts, p_lo_gid, p_hi_pts, p_hi_gid,
                                        d_counts, p_lo_flag, p_hi_flag, epoch))

    def halo_wait(self, d_flag_a: Optional[int], d_flag_b: Optional[int], epoch: int, timeout_ms: int, d_err: int):
        self._ck(self.lib.pcp_halo_wait(self.h, d_flag_a, d_flag_b, epoch, timeout_ms, d_err))

    def peer_alloc(self, nbytes: int):
        """-> (device pointer, 64-byte IPC handle)"""
        ptr = C.c_void_p()
        handle = C.create_string_buffer(64)
        self._ck(self.lib.pcp_peer_alloc(self.h, nbytes, C.byref(ptr), handle))
        return ptr.value, handle.raw

    def peer_free(self, ptr: int):
        self._ck(self.lib.pcp_peer_free(self.h, ptr))

    def peer_open(self, handle: bytes) -> int:
        ptr = C.c_void_p()
        self._ck(self.lib.pcp_peer_open(self.h, C.create_string_buffer(handle, 64), C.byref(ptr)))
        return ptr.value

    def peer_close(self, ptr: int):
        self._ck(self.lib.pcp_peer_close(self.h, ptr))

    def set_host_async(self, enable: bool):
        """Host-pointer calls only enqueue (pinned buffers, caller synchronises): see pcp_set_host_async."""
        self._ck(self.lib.pcp_set_host_async(self.h, 1 if enable else 0))

    def timings(self) -> dict:
        ms = np.zeros(len(STAGES), np.float32)
        self._ck(self.lib.pcp_last_timings(self.h, ms.ctypes.data, len(STAGES)))
        return dict(zip(STAGES, ms.tolist()))

    def launch_count(self) -> int:
        return int(self.lib.pcp_launch_count(self.h))

    # ---- cloud + grid
    def set_cloud(self, pts: np.ndarray):
        """pts: (n, 4) float32 packed xyzi, or (n, 8) float32 = pcl::PointXYZI's 32-byte layout."""
        pts = np.ascontiguousarray(pts, np.float32)
        assert pts.ndim == 2 and pts.shape[1] in (3, 4, 8)
        self._ck(self.lib.pcp_set_cloud_host(self.h, pts.ctypes.data, pts.shape[0], pts.shape[1] * 4))
        self.n = self.n_query = pts.shape[0]

    def set_cloud_host_ptr(self, ptr: int, n: int, stride: int):
        self._ck(self.lib.pcp_set_cloud_host(self.h, ptr, n, stride))
        self.n = self.n_query = n

    def set_cloud_device(self, d_ptr: int, n: int, d_gid: Optional[int] = None, n_query: Optional[int] = None, keep=None):
        nq = n if n_query is None else n_query
        self._ck(self.lib.pcp_set_cloud_device_sharded(self.h, d_ptr, d_gid, n, nq))
        self.n, self.n_query, self._keep = n, nq, keep

    def set_frames(self, frame_offsets: np.ndarray):
        """The current cloud is len(frame_offsets) - 1 independent clouds laid end to end (pcp_set_frames)."""
        fo = np.ascontiguousarray(frame_offsets, np.int64)
        self._ck(self.lib.pcp_set_frames(self.h, fo.ctypes.data, fo.shape[0] - 1))

    def build_grid(self, cell: float = 0.0, k_hint: int = 0):
        self._ck(self.lib.pcp_build_grid(self.h, float(cell), int(k_hint)))

    def grid_info(self):
        dims = np.zeros(3, np.int32)
        cell = C.c_float()
        nc = C.c_int64()
        self._ck(self.lib.pcp_grid_info(self.h, dims.ctypes.data, C.addressof(cell), C.addressof(nc)))
        return dims.tolist(), cell.value, nc.value

    # ---- compute (host outputs unless device pointers are given)
    def radius_pca(self, radius: float, min_pts: int = 13, center_mode: int = CENTER_QUERY, carry_forward: bool = True,
                   out_device: Optional[int] = None):
        if out_device is not None:
            self._ck(self.lib.pcp_radius_pca(self.h, radius, min_pts, center_mode, int(carry_forward), out_device, 1))
            return None
        out = np.zeros(self.n_query, FEATURE_DTYPE)
        self._ck(self.lib.pcp_radius_pca(self.h, radius, min_pts, center_mode, int(carry_forward), out.ctypes.data, 0))
        return out

    def knn_pca(self, k: int, center_mode: int = CENTER_QUERY, want_normals: bool = False, out_device: Optional[int] = None,
                normals_device: Optional[int] = None, out_host_ptr: Optional[int] = None, kth_device: Optional[int] = None):
        if out_device is not None or normals_device is not None:
            self._ck(self.lib.pcp_knn_pca_kth(self.h, k, center_mode, out_device, normals_device, kth_device, 1))
            return None
        if out_host_ptr is not None:
            self._ck(self.lib.pcp_knn_pca(self.h, k, center_mode, out_host_ptr, None, 0))
            return None
        out = np.zeros(self.n_query, FEATURE_DTYPE)
        nrm = np.zeros((self.n_query, 4), np.float32) if want_normals else None
        self._ck(self.lib.pcp_knn_pca(self.h, k, center_mode, out.ctypes.data, _ptr(nrm), 0))
        return (out, nrm) if want_normals else out

    def knn_pca_kth(self, k: int, center_mode: int = CENTER_QUERY):
        """Records + the squared distance of every query's k-th neighbour (inf when the cloud holds fewer than k)."""
        out = np.zeros(self.n_query, FEATURE_DTYPE)
        kth = np.zeros(self.n_query, np.float32)
        self._ck(self.lib.pcp_knn_pca_kth(self.h, k, center_mode, out.ctypes.data, None, kth.ctypes.data, 0))
        return out, kth

    def knn_neighbors(self, k: int):
        idx = np.zeros((self.n_query, k), np.int32)
        d2 = np.zeros((self.n_query, k), np.float32)
        self._ck(self.lib.pcp_knn_neighbors(self.h, k, idx.ctypes.data, d2.ctypes.data, 0))
        return idx, d2

    def radius_neighbors(self, radius: float):
        counts = np.zeros(self.n_query, np.int32)
        self._ck(self.lib.pcp_radius_neighbors(self.h, radius, counts.ctypes.data, None, None, None, 0))
        offsets = np.zeros(self.n_query + 1, np.int64)
        np.cumsum(counts, out=offsets[1:])
        total = int(offsets[-1])
        idx = np.zeros(max(total, 1), np.int32)
        d2 = np.zeros(max(total, 1), np.float32)
        self._ck(self.lib.pcp_radius_neighbors(self.h, radius, None, offsets.ctypes.data, idx.ctypes.data, d2.ctypes.data, 0))
        return offsets, idx[:total], d2[:total]

    def segment_pca(self, centres: np.ndarray, offsets: np.ndarray, idx: np.ndarray):
        centres = np.ascontiguousarray(centres, np.float32)
        offsets = np.ascontiguousarray(offsets, np.int64)
        idx = np.ascontiguousarray(idx, np.int32)
        nseg = offsets.shape[0] - 1
        out = np.zeros(nseg, FEATURE_DTYPE)
        self._ck(self.lib.pcp_segment_pca(self.h, centres.ctypes.data, offsets.ctypes.data, idx.ctypes.data, nseg, out.ctypes.data))
        return out

    def carry_forward_static(self, d_in: int, d_out: int, n: int, min_pts: int = 13):
        """Stale-record carry on DEVICE records with the library's static scratch (no allocation): pcp_carry_forward_static."""
        self._ck(self.lib.pcp_carry_forward_static(self.h, d_in, d_out, n, min_pts))

    def bbox_pca(self) -> dict:
        """getMiniBoundingBox (src/pcaAnalysis.cpp:123-148) of the current cloud: centroid, eigenvalues, rotation rows, extents."""
        o = Obb()
        self._ck(self.lib.pcp_bbox_pca(self.h, C.byref(o)))
        return {"centroid": np.array(o.centroid, np.float32), "eigval": np.array(o.eigval, np.float32),
                "axes": np.array([list(r) for r in o.axes], np.float32), "min_pt": np.array(o.min_pt, np.float32),
                "max_pt": np.array(o.max_pt, np.float32), "mean_diag": np.array(o.mean_diag, np.float32), "n_points": int(o.n_points)}

    def ring_project_host_ptr(self, calib: RingCalib, range_ptr: int, index_ptr: int, long_ptr: int,
                              frame_offsets: Optional[np.ndarray] = None):
        """Outputs to raw (page-locked) host pointers; one frame = the whole current cloud unless
        frame_offsets (int64, n_frames + 1) is given."""
        if frame_offsets is None:
            self._ck(self.lib.pcp_ring_project(self.h, C.byref(calib), None, 1, range_ptr, index_ptr, long_ptr, 0))
        else:
            fo = np.ascontiguousarray(frame_offsets, np.int64)
            self._ck(self.lib.pcp_ring_project(self.h, C.byref(calib), fo.ctypes.data, fo.shape[0] - 1, range_ptr, index_ptr, long_ptr, 0))

    def decode_packets(self, packets: np.ndarray, calib: Optional[RingCalib] = None, want_arrays: bool = True,
                       packets_device: Optional[int] = None, n_packets: Optional[int] = None) -> dict:
        """Velodyne packets ((n, 1206) uint8 on the host, or a device pointer) -> the decoded points become the
        current cloud (with its frames); returns frame_offsets / n_complete / n_points and, with want_arrays,
        the per-return arrays of PacketDecoder::HDLFrame."""
        calib = calib or ring_calib_default()
        if packets_device is None:
            pk = np.ascontiguousarray(packets, np.uint8).reshape(-1, 1206)
            npk, ptr, on_dev = pk.shape[0], pk.ctypes.data, 0
        else:
            npk, ptr, on_dev = int(n_packets), packets_device, 1
        cap = npk * 384
        max_frames = npk * 12 + 2
        offs = np.zeros(max_frames, np.int64)
        nfc, npts = C.c_int64(), C.c_int64()
        arrays, dec = {}, None
        if want_arrays:
            arrays = {"x": np.zeros(cap), "y": np.zeros(cap), "z": np.zeros(cap), "distance": np.zeros(cap),
                      "intensity": np.zeros(cap, np.uint8), "laser_id": np.zeros(cap, np.uint8), "azimuth": np.zeros(cap, np.uint16),
                      "ms": np.zeros(cap, np.uint32)}
            dec = Decoded(**{k: v.ctypes.data for k, v in arrays.items()})
        self._ck(self.lib.pcp_decode_packets(self.h, C.byref(calib), ptr, npk, on_dev, C.byref(dec) if dec else None, 0, offs.ctypes.data,
                                             max_frames, C.byref(nfc), C.byref(npts)))
        n, nc = npts.value, nfc.value
        self.n = self.n_query = n
        res = {k: v[:n] for k, v in arrays.items()}
        res.update({"frame_offsets": offs[: nc + 2].copy(), "n_complete": nc, "n_points": n})
        return res

    def ring_segment(self, calib: Optional[RingCalib] = None, prm: Optional["RingSegParams"] = None,
                     frame_offsets: Optional[np.ndarray] = None) -> dict:
        """RingProjector::fullSegmentation on the current cloud (one frame, or frame_offsets.shape[0] - 1 frames)."""
        calib = calib or ring_calib_default()
        prm = prm or ring_seg_params_default()
        nf = 1 if frame_offsets is None else frame_offsets.shape[0] - 1
        fo = None if frame_offsets is None else np.ascontiguousarray(frame_offsets, np.int64)
        ns, hs = calib.n_scan, calib.horizon_scan
        out = {
            "range": np.zeros((nf, ns, hs), np.float32), "index": np.zeros((nf, ns, hs), np.int32), "long_range": np.zeros(nf, np.int32),
            "ground": np.zeros((nf, ns, hs), np.int8), "label": np.zeros((nf, ns, hs), np.int32),
            "col_max_ground": np.zeros((nf, hs), np.float32), "above_idx": np.zeros((nf, ns * hs), np.int32),
            "above_count": np.zeros(nf, np.int32), "n_labels": np.zeros(nf, np.int32),
        }
        self._ck(self.lib.pcp_ring_segment(self.h, C.byref(calib), C.byref(prm), _ptr(fo), nf, *[out[key].ctypes.data for key in (
            "range", "index", "long_range", "ground", "label", "col_max_ground", "above_idx", "above_count", "n_labels")], 0))
        return out

    def ring_segment_host_ptr(self, calib: RingCalib, prm: "RingSegParams", frame_offsets: np.ndarray, range_ptr: int, index_ptr: int,
                              long_ptr: int, ground_ptr: int, label_ptr: int, colmax_ptr: int, above_ptr: int, above_count_ptr: int,
                              n_labels_ptr: int):
        """pcp_ring_segment with raw (page-locked) host output pointers."""
        fo = np.ascontiguousarray(frame_offsets, np.int64)
        self._ck(self.lib.pcp_ring_segment(self.h, C.byref(calib), C.byref(prm), fo.ctypes.data, fo.shape[0] - 1, range_ptr, index_ptr,
                                           long_ptr, ground_ptr, label_ptr, colmax_ptr, above_ptr, above_count_ptr, n_labels_ptr, 0))

    def ring_project(self, calib: Optional[RingCalib] = None, frame_offsets: Optional[np.ndarray] = None,
                     range_device: Optional[int] = None, index_device: Optional[int] = None, long_device: Optional[int] = None):
        calib = calib or ring_calib_default()
        nf = 1 if frame_offsets is None else frame_offsets.shape[0] - 1
        fo = None if frame_offsets is None else np.ascontiguousarray(frame_offsets, np.int64)
        cells = calib.n_scan * calib.horizon_scan
        if range_device is not None:
            self._ck(self.lib.pcp_ring_project(self.h, C.byref(calib), _ptr(fo), nf, range_device, index_device, long_device, 1))
            return None
        rng = np.zeros((nf, calib.n_scan, calib.horizon_scan), np.float32)
        pti = np.zeros((nf, calib.n_scan, calib.horizon_scan), np.int32)
        lr = np.zeros(nf, np.int32)
        assert rng.size == cells * nf
        self._ck(self.lib.pcp_ring_project(self.h, C.byref(calib), _ptr(fo), nf, rng.ctypes.data, pti.ctypes.data, lr.ctypes.data, 0))
        return rng, pti, lr

    def radius_count(self, radius: float, cap: int = 0) -> np.ndarray:
        """Per-point neighbour count within radius (query included), min(count, cap) when cap > 0."""
        out = np.zeros(self.n_query, np.int32)
        self._ck(self.lib.pcp_radius_count(self.h, radius, cap, out.ctypes.data, 0))
        return out

    def odom_range_image(self, prm: Optional[OdomParams] = None, frame_offsets: Optional[np.ndarray] = None) -> dict:
        """lidarOdometryClass::getRangeAndptIdImage per frame: range / ptid / pt_index images, stats, angle min/max."""
        prm = prm or odom_params_default()
        nf = 1 if frame_offsets is None else frame_offsets.shape[0] - 1
        fo = None if frame_offsets is None else np.ascontiguousarray(frame_offsets, np.int64)
        shape = (nf, prm.n_rows, prm.n_cols)
        out = {"range": np.zeros(shape, np.float32), "ptid": np.zeros(shape, np.float32), "pt_index": np.zeros(shape, np.int32),
               "stats": np.zeros((nf, 2), np.int32), "minmax": np.zeros((nf, 4), np.float64)}
        self._ck(self.lib.pcp_odom_range_image(self.h, C.byref(prm), _ptr(fo), nf, out["range"].ctypes.data, out["ptid"].ctypes.data,
                                               out["pt_index"].ctypes.data, out["stats"].ctypes.data, out["minmax"].ctypes.data, 0))
        return out

    def odom_smoothness(self, range_img: Optional[np.ndarray] = None, prm: Optional[OdomParams] = None, n_frames: Optional[int] = None):
        """lidarOdometryClass::calculateSmoothness: (smoothness images, point-type images); range_img None = the device
        images of the last odom_range_image call."""
        prm = prm or odom_params_default()
        if range_img is not None:
            range_img = np.ascontiguousarray(range_img, np.float32).reshape(-1, prm.n_rows, prm.n_cols)
            n_frames = range_img.shape[0]
        shape = (int(n_frames), prm.n_rows, prm.n_cols)
        sm = np.zeros(shape, np.float32)
        ty = np.zeros(shape, np.float32)
        self._ck(self.lib.pcp_odom_smoothness(self.h, C.byref(prm), _ptr(range_img), 0, shape[0], sm.ctypes.data, ty.ctypes.data, 0))
        return sm, ty

    def dbscan(self, radius: float, min_pts: int):
        """DynaTools::clusterDBSCAN's state vector on the current (flattened by the caller) cloud: (labels, n_clusters)."""
        labels = np.zeros(self.n, np.int32)
        nc = C.c_int32()
        self._ck(self.lib.pcp_dbscan(self.h, radius, min_pts, labels.ctypes.data, C.addressof(nc), 0))
        return labels, int(nc.value)

    def dbscan_d(self, radius: float, min_pts: int):
        """The same with the radius as the reference's `double r` (pcp_dbscan_d)."""
        labels = np.zeros(self.n, np.int32)
        nc = C.c_int32()
        self._ck(self.lib.pcp_dbscan_d(self.h, C.c_double(radius), min_pts, labels.ctypes.data, C.addressof(nc), 0))
        return labels, int(nc.value)

    def cluster_boxes(self, labels: np.ndarray, n_clusters: int):
        """Per cluster of `labels` over the current cloud: (centroid, min, max, count)."""
        labels = np.ascontiguousarray(labels, np.int32)
        cen = np.zeros((n_clusters, 3), np.float32)
        lo = np.zeros((n_clusters, 3), np.float32)
        hi = np.zeros((n_clusters, 3), np.float32)
        cnt = np.zeros(n_clusters, np.int32)
        self._ck(self.lib.pcp_cluster_boxes(self.h, labels.ctypes.data, 0, n_clusters, cen.ctypes.data, lo.ctypes.data, hi.ctypes.data,
                                            cnt.ctypes.data))
        return cen, lo, hi, cnt

    def mean_shift(self, radius: float = 2.0, min_pts: int = 15, iterations: int = 1, start: Optional[np.ndarray] = None):
        """meanShiftClustering sweeps over the current cloud (grid = the original positions): (positions n x 4, counts)."""
        st = None if start is None else np.ascontiguousarray(start, np.float32)
        out = np.zeros((self.n, 4), np.float32)
        cnt = np.zeros(self.n, np.int32)
        self._ck(self.lib.pcp_mean_shift(self.h, radius, min_pts, iterations, _ptr(st), 0, out.ctypes.data, cnt.ctypes.data, 0))
        return out, cnt

    def set_cloud_txt(self, text: bytes) -> np.ndarray:
        """preprocess::txt2pc on the device: the parsed points become the current cloud; returns them (n, 4)."""
        buf = np.frombuffer(text, np.uint8)
        cap = text.count(b"\n") + 1
        out = np.zeros((cap, 4), np.float32)
        n = C.c_int64()
        self._ck(self.lib.pcp_set_cloud_txt(self.h, buf.ctypes.data if len(text) else None, len(text), 0, C.addressof(n), out.ctypes.data, cap, 0))
        self.n = self.n_query = int(n.value)
        return out[: int(n.value)].copy()

    def set_cloud_xyz(self, text: bytes) -> np.ndarray:
        """preprocess::xyz2pc on the device (pcp_set_cloud_xyz): the parsed points become the current cloud; returns them (n, 4)."""
        buf = np.frombuffer(text, np.uint8)
        cap = text.count(b"\n") + 2
        out = np.zeros((cap, 4), np.float32)
        n = C.c_int64()
        self._ck(self.lib.pcp_set_cloud_xyz(self.h, buf.ctypes.data if len(text) else None, len(text), 0, C.addressof(n), out.ctypes.data, cap, 0))
        self.n = self.n_query = int(n.value)
        return out[: int(n.value)].copy()
