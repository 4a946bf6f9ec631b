"""ctypes wrapper of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (see pcp_oracle.h).

Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs, never by the product package.  PARITY UNPINNED (no reference fixtures exist).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")

FEATURE_DTYPE = np.dtype(
    [("la", np.float32, (3,)), ("v1", np.float32, (3,)), ("v2", np.float32, (3,)), ("v3", np.float32, (3,)),
     ("ptID", np.int32), ("ptNum", np.int32), ("lineStruc", np.float32), ("planeStruc", np.float32),
     ("scatterStruc", np.float32)]
)
assert FEATURE_DTYPE.itemsize == 68

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    sig = {
        "orc_radius2": (f32, [f32]),
        "orc_radius_brute": (i64, [vp, i64, i64, f32, vp, vp, i64]),
        "orc_knn_brute": (i64, [vp, i64, i64, i32, vp, vp]),
        "orc_kdtree_build": (vp, [vp, i64, i32]),
        "orc_kdtree_free": (None, [vp]),
        "orc_kdtree_radius": (i64, [vp, i64, f32, vp, vp, i64]),
        "orc_kdtree_knn": (i64, [vp, i64, i32, vp, vp]),
        "orc_kdtree_knn_all": (None, [vp, i32, vp, vp, i32]),
        "orc_kdtree_radius_counts": (None, [vp, f32, vp, i32]),
        "orc_kdtree_radius_all": (None, [vp, f32, vp, vp, vp, i32]),
        "orc_covariance_about": (None, [vp, vp, i64, vp, vp]),
        "orc_eigen_sym3_qr": (i32, [vp, vp, vp]),
        "orc_pcl_eigen33": (None, [vp, vp, vp]),
        "orc_structure_features": (None, [f32, f32, f32, vp]),
        "orc_radius_pca_cloud": (None, [vp, i64, f32, i32, i32, vp, vp, i32]),
        "orc_knn_pca_cloud": (None, [vp, i64, i32, i32, vp, vp, i32]),
        "orc_segment_pca": (None, [vp, vp, vp, vp, i64, vp]),
        "orc_bbox_pca": (None, [vp, i64, vp]),
        "orc_ring_project": (i32, [vp, i64, i32, i32, f32, f32, f32, vp, vp, vp]),
        "orc_decode_packets": (i64, [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i64, vp]),
        "orc_ring_segment": (i32, [vp, i32, i32, f32, f32, i32, f32, f32, i32, i32, vp, vp, vp, vp, vp, vp, vp]),
        "orc_odom_range_image": (None, [vp, i64, vp, i32, i32, i32, C.c_double, C.c_double, vp, vp, vp, vp, vp]),
        "orc_odom_smoothness": (None, [vp, i32, i32, i32, i32, f32, i32, i32, i32, vp, vp]),
        "orc_atanf_mismatches": (i64, [C.c_uint32, C.c_uint32]),
        "orc_dbscan": (i32, [vp, i64, f32, i32, vp]),
        "orc_dbscan_d": (i32, [vp, i64, C.c_double, i32, vp]),
        "orc_cluster_boxes": (None, [vp, i64, vp, i32, vp, vp, vp, vp]),
        "orc_mean_shift_step": (None, [vp, vp, i64, f32, i32, vp, vp, vp]),
        "orc_txt2pc": (i64, [C.c_char_p, i64, vp, i64]),
        "orc_xyz2pc": (i64, [C.c_char_p, i64, vp, i64]),
        "orc_num_threads": (i32, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _f4(pts):
    pts = np.ascontiguousarray(pts, np.float32)
    assert pts.ndim == 2 and pts.shape[1] == 4
    return pts


def num_threads() -> int:
    return int(load().orc_num_threads())


def radius2(r: float) -> float:
    return float(load().orc_radius2(r))


def knn_brute(pts, q: int, k: int):
    pts = _f4(pts)
    idx = np.zeros(k, np.int32)
    d2 = np.zeros(k, np.float32)
    m = load().orc_knn_brute(pts.ctypes.data, pts.shape[0], q, k, idx.ctypes.data, d2.ctypes.data)
    return idx[:m], d2[:m]


def radius_brute(pts, q: int, radius: float):
    pts = _f4(pts)
    r2 = radius2(radius)
    m = load().orc_radius_brute(pts.ctypes.data, pts.shape[0], q, r2, None, None, 0)
    idx = np.zeros(max(m, 1), np.int32)
    d2 = np.zeros(max(m, 1), np.float32)
    load().orc_radius_brute(pts.ctypes.data, pts.shape[0], q, r2, idx.ctypes.data, d2.ctypes.data, m)
    return idx[:m], d2[:m]


class KdTree:
    def __init__(self, pts, leaf_max: int = 15):
        self.pts = _f4(pts)
        self.n = self.pts.shape[0]
        self.h = load().orc_kdtree_build(self.pts.ctypes.data, self.n, leaf_max)

    def __del__(self):
        if getattr(self, "h", None):
            load().orc_kdtree_free(self.h)
            self.h = None

    def knn_all(self, k: int, nthreads: int = 0):
        idx = np.zeros((self.n, k), np.int32)
        d2 = np.zeros((self.n, k), np.float32)
        load().orc_kdtree_knn_all(self.h, k, idx.ctypes.data, d2.ctypes.data, nthreads or num_threads())
        return idx, d2

    def radius_counts(self, radius: float, nthreads: int = 0):
        counts = np.zeros(self.n, np.int32)
        load().orc_kdtree_radius_counts(self.h, radius2(radius), counts.ctypes.data, nthreads or num_threads())
        return counts

    def radius_all(self, radius: float, nthreads: int = 0):
        r2 = radius2(radius)
        nt = nthreads or num_threads()
        counts = np.zeros(self.n, np.int32)
        load().orc_kdtree_radius_counts(self.h, r2, counts.ctypes.data, nt)
        offsets = np.zeros(self.n + 1, np.int64)
        np.cumsum(counts, out=offsets[1:])
        total = int(offsets[-1])
        idx = np.zeros(max(total, 1), np.int32)
        d2 = np.zeros(max(total, 1), np.float32)
        load().orc_kdtree_radius_all(self.h, r2, offsets.ctypes.data, idx.ctypes.data, d2.ctypes.data, nt)
        return offsets, idx[:total], d2[:total]


def covariance_about(pts, idx, centre):
    pts = _f4(pts)
    idx = np.ascontiguousarray(idx, np.int32)
    centre = np.ascontiguousarray(centre, np.float32)
    Cm = np.zeros(9, np.float32)
    load().orc_covariance_about(pts.ctypes.data, idx.ctypes.data, idx.shape[0], centre.ctypes.data, Cm.ctypes.data)
    return Cm.reshape(3, 3)


def eigen_sym3_qr(Cm):
    Cm = np.ascontiguousarray(Cm, np.float32).reshape(9)
    ev = np.zeros(3, np.float32)
    Q = np.zeros(9, np.float32)
    it = load().orc_eigen_sym3_qr(Cm.ctypes.data, ev.ctypes.data, Q.ctypes.data)
    return ev, Q.reshape(3, 3), it


def pcl_eigen33(Cm):
    Cm = np.ascontiguousarray(Cm, np.float32).reshape(9)
    ev = np.zeros(1, np.float32)
    v = np.zeros(3, np.float32)
    load().orc_pcl_eigen33(Cm.ctypes.data, ev.ctypes.data, v.ctypes.data)
    return float(ev[0]), v


def structure_features(la1, la2, la3):
    o = np.zeros(3, np.float32)
    load().orc_structure_features(la1, la2, la3, o.ctypes.data)
    return o


def bbox_pca(pts):
    """getMiniBoundingBox restated (src/pcaAnalysis.cpp:123-148): dict of centroid, eigval, axes (3x3 rows), min_pt,
    max_pt, mean_diag."""
    pts = _f4(pts)
    out = np.zeros(24, np.float32)
    load().orc_bbox_pca(pts.ctypes.data, pts.shape[0], out.ctypes.data)
    return {"centroid": out[0:3], "eigval": out[3:6], "axes": out[6:15].reshape(3, 3), "min_pt": out[15:18], "max_pt": out[18:21],
            "mean_diag": out[21:24]}


def radius_pca_cloud(pts, radius: float, min_pts: int = 13, carry_forward: bool = True, nthreads: int = 1):
    pts = _f4(pts)
    out = np.zeros(pts.shape[0], FEATURE_DTYPE)
    valid = np.zeros(pts.shape[0], np.uint8)
    load().orc_radius_pca_cloud(pts.ctypes.data, pts.shape[0], radius, min_pts, int(carry_forward), out.ctypes.data,
                                valid.ctypes.data, nthreads)
    return out, valid


def knn_pca_cloud(pts, k: int, center_mode: int = 0, want_normals: bool = False, nthreads: int = 0, want_features: bool = True):
    pts = _f4(pts)
    out = np.zeros(pts.shape[0], FEATURE_DTYPE) if want_features else None
    nrm = np.zeros((pts.shape[0], 4), np.float32) if want_normals else None
    load().orc_knn_pca_cloud(pts.ctypes.data, pts.shape[0], k, center_mode, out.ctypes.data if want_features else None,
                             nrm.ctypes.data if want_normals else None, nthreads or num_threads())
    return (out, nrm) if want_normals else out


def segment_pca(pts, centres, offsets, idx):
    pts = _f4(pts)
    centres = np.ascontiguousarray(centres, np.float32)
    offsets = np.ascontiguousarray(offsets, np.int64)
    idx = np.ascontiguousarray(idx, np.int32)
    nseg = offsets.shape[0] - 1
    out = np.zeros(nseg, FEATURE_DTYPE)
    load().orc_segment_pca(pts.ctypes.data, centres.ctypes.data, offsets.ctypes.data, idx.ctypes.data, nseg, out.ctypes.data)
    return out


def ring_project(pts, n_scan=16, horizon=1800, ang_res_x=0.2, ang_res_y=2.0, ang_bottom=15.0 + 0.1):
    pts = _f4(pts)
    rng = np.zeros((n_scan, horizon), np.float32)
    pti = np.zeros((n_scan, horizon), np.int32)
    tag = np.zeros((n_scan, horizon), np.float32)
    lr = load().orc_ring_project(pts.ctypes.data, pts.shape[0], n_scan, horizon, np.float32(ang_res_x), np.float32(ang_res_y),
                                 np.float32(ang_bottom), rng.ctypes.data, pti.ctypes.data, tag.ctypes.data)
    return rng, pti, tag, int(lr)


def ring_segment(pts, rng, pti, n_scan=16, horizon=1800, ang_res_x=0.2, ang_res_y=2.0, ground_scan_ind=7, sensor_mount_angle=0.0,
                 segment_theta=1.0472, valid_point_num=5, valid_line_num=3):
    """groundRemoval + cloudSegmentation (src/ringProjection.cpp:83-289) on one projected frame.
    Returns ground (int8 image), label (int32 image), col_max_ground, above_idx (original point indices
    of the cloud above ground, raster order), n_labels."""
    pts = _f4(pts)
    rng = np.ascontiguousarray(rng, np.float32)
    pti = np.ascontiguousarray(pti, np.int32)
    ground = np.zeros((n_scan, horizon), np.int8)
    label = np.zeros((n_scan, horizon), np.int32)
    colmax = np.zeros(horizon, np.float32)
    above = np.zeros(n_scan * horizon, np.int32)
    nl = np.zeros(1, np.int32)
    na = load().orc_ring_segment(pts.ctypes.data, n_scan, horizon, np.float32(ang_res_x), np.float32(ang_res_y), ground_scan_ind,
                                 np.float32(sensor_mount_angle), np.float32(segment_theta), valid_point_num, valid_line_num,
                                 rng.ctypes.data, pti.ctypes.data, ground.ctypes.data, label.ctypes.data, colmax.ctypes.data,
                                 above.ctypes.data, nl.ctypes.data)
    return ground, label, colmax, above[:na].copy(), int(nl[0])


def decode_packets(packets, vert_deg):
    """Velodyne packets (n, 1206) uint8 -> dict of per-return arrays + frame_offsets (complete frames first,
    then the frame in progress) + n_complete.  vert_deg: 64 vertical corrections in degrees (16db.xml)."""
    pk = np.ascontiguousarray(packets, np.uint8).reshape(-1, 1206)
    n_max = pk.shape[0] * 384
    vd = np.ascontiguousarray(vert_deg, np.float32)
    out = {"x": np.zeros(n_max), "y": np.zeros(n_max), "z": np.zeros(n_max), "distance": np.zeros(n_max),
           "intensity": np.zeros(n_max, np.uint8), "laser_id": np.zeros(n_max, np.uint8), "azimuth": np.zeros(n_max, np.uint16),
           "ms": np.zeros(n_max, np.uint32)}
    max_frames = pk.shape[0] * 12 + 2
    offs = np.zeros(max_frames, np.int64)
    nf = np.zeros(1, np.int64)
    n = load().orc_decode_packets(pk.ctypes.data, pk.shape[0], vd.ctypes.data, *[out[k].ctypes.data for k in (
        "x", "y", "z", "distance", "intensity", "laser_id", "azimuth", "ms")], offs.ctypes.data, max_frames, nf.ctypes.data)
    res = {k: v[:n].copy() for k, v in out.items()}
    res["frame_offsets"] = offs[: int(nf[0]) + 2].copy()
    res["n_complete"] = int(nf[0])
    return res


def odom_range_image(pts, nb_counts, min_neighbors=6, n_rows=16, n_cols=900, vert_res=2.0, hori_res=0.2):
    """getRangeAndptIdImage (src/lidarOdometry.cpp:21-113) on one frame; nb_counts = radius-search counts (r = 3)."""
    pts = _f4(pts)
    cnt = np.ascontiguousarray(nb_counts, np.int32)
    rng = np.zeros((n_rows, n_cols), np.float32)
    pid = np.zeros((n_rows, n_cols), np.float32)
    pti = np.zeros((n_rows, n_cols), np.int32)
    stats = np.zeros(2, np.int32)
    mm = np.zeros(4, np.float64)
    load().orc_odom_range_image(pts.ctypes.data, pts.shape[0], cnt.ctypes.data, min_neighbors, n_rows, n_cols,
                                C.c_double(vert_res), C.c_double(hori_res), rng.ctypes.data, pid.ctypes.data,
                                pti.ctypes.data, stats.ctypes.data, mm.ctypes.data)
    return rng, pid, pti, stats, mm


def odom_smoothness(rng, n_blocks=5, half_window=9, near_range=1.0, min_valid=9, pick=3, min_block_valid=9):
    """calculateSmoothness (src/lidarOdometry.cpp:121-187): (smoothness image, point-type image)."""
    rng = np.ascontiguousarray(rng, np.float32)
    n_rows, n_cols = rng.shape
    sm = np.zeros_like(rng)
    ty = np.zeros_like(rng)
    load().orc_odom_smoothness(rng.ctypes.data, n_rows, n_cols, n_blocks, half_window, C.c_float(near_range), min_valid,
                               pick, min_block_valid, sm.ctypes.data, ty.ctypes.data)
    return sm, ty


def dbscan(pts, radius: float, min_pts: int):
    """DynaTools::clusterDBSCAN on the given (already flattened) cloud: (state, n_clusters)."""
    pts = _f4(pts)
    state = np.zeros(pts.shape[0], np.int32)
    nc = load().orc_dbscan(pts.ctypes.data, pts.shape[0], radius, min_pts, state.ctypes.data)
    return state, int(nc)


def dbscan_d(pts, radius: float, min_pts: int):
    """The same with the radius taken as the reference's `double r` (r2 = (float)(r * r) in double)."""
    pts = _f4(pts)
    state = np.zeros(pts.shape[0], np.int32)
    nc = load().orc_dbscan_d(pts.ctypes.data, pts.shape[0], C.c_double(radius), min_pts, state.ctypes.data)
    return state, int(nc)


def cluster_boxes(pts, state, n_clusters: int):
    pts = _f4(pts)
    state = np.ascontiguousarray(state, np.int32)
    cen = np.zeros((n_clusters, 3), np.float32)
    lo = np.zeros((n_clusters, 3), np.float32)
    hi = np.zeros((n_clusters, 3), np.float32)
    cnt = np.zeros(n_clusters, np.int32)
    load().orc_cluster_boxes(pts.ctypes.data, pts.shape[0], state.ctypes.data, n_clusters, cen.ctypes.data, lo.ctypes.data,
                             hi.ctypes.data, cnt.ctypes.data)
    return cen, lo, hi, cnt


def mean_shift_step(orig, cur, radius: float = 2.0, min_pts: int = 15):
    """One sweep of meanShiftClustering: (next positions n x 4, shift n x 3, neighbour counts)."""
    orig, cur = _f4(orig), _f4(cur)
    nxt = np.zeros_like(cur)
    shift = np.zeros((cur.shape[0], 3), np.float32)
    cnt = np.zeros(cur.shape[0], np.int32)
    load().orc_mean_shift_step(orig.ctypes.data, cur.ctypes.data, cur.shape[0], radius, min_pts, nxt.ctypes.data,
                               shift.ctypes.data, cnt.ctypes.data)
    return nxt, shift, cnt


def txt2pc(text: bytes):
    """preprocess::txt2pc on an in-memory file: (n_lines, 4) float32."""
    cap = text.count(b"\n") + 1
    out = np.zeros((cap, 4), np.float32)
    n = load().orc_txt2pc(text, len(text), out.ctypes.data, cap)
    return out[:n].copy()


def xyz2pc(text: bytes):
    """preprocess::xyz2pc (src/smoothing.cpp:83-125) on an in-memory file: (n_points, 4) float32; None when a line has 255
    characters or more (the reference's loop never ends on such a file)."""
    cap = text.count(b"\n") + 2
    out = np.zeros((cap, 4), np.float32)
    n = load().orc_xyz2pc(text, len(text), out.ctypes.data, cap)
    return None if n < 0 else out[:n].copy()
