/*
 * pcp.h — C-ABI of the B200-native neighbourhood-PCA hot path (libpcp.so).
 *
 * This is the drop-in boundary for the reference's `pcaAnalysist` stage and for
 * `RingProjector::projectPointCloud`.  Plain C: pointers and sizes only, no C++ types,
 * no torch types, no exceptions across the boundary.  Every entry point returns a
 * `pcp_status` (0 = OK, negative = error); `pcp_last_error(ctx)` gives the message.
 *
 * There is NO CPU fallback behind this ABI: every compute call runs CUDA kernels compiled
 * for sm_100a and fails with PCP_ERR_CUDA / PCP_ERR_NO_DEVICE when no such GPU is present.
 *
 * One `pcp_ctx` per host thread per GPU.  A ctx owns one CUDA stream and all device
 * memory it allocates; the caller owns every pointer it passes in.  Calls taking host
 * pointers synchronise the ctx stream before returning; calls taking device pointers
 * (`*_device`, or `out_is_device != 0`) only enqueue work on the ctx stream.
 *
 * Each function cites the reference interface it replaces (paths relative to the
 * reference tree YangSiri/pointCloudProcessing).
 */
#ifndef PCP_H_
#define PCP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PCP_API __attribute__((visibility("default")))
#else
#define PCP_API
#endif

typedef struct pcp_ctx pcp_ctx;

typedef enum pcp_status {
  PCP_OK = 0,
  PCP_ERR_INVALID_ARG = -1,
  PCP_ERR_NO_DEVICE = -2,      /* no CUDA device / not sm_100 */
  PCP_ERR_CUDA = -3,           /* a CUDA runtime call or kernel failed */
  PCP_ERR_NO_CLOUD = -4,       /* compute call before pcp_set_cloud_* */
  PCP_ERR_NO_GRID = -5,        /* compute call before pcp_build_grid */
  PCP_ERR_GRID_TOO_LARGE = -6, /* dense cell table would exceed the configured budget */
  PCP_ERR_K_UNSUPPORTED = -7,  /* k outside [1, PCP_MAX_K] */
  PCP_ERR_OOM = -8,
  PCP_ERR_IO = -9,             /* calibration file could not be read / parsed */
  PCP_ERR_NEIGHBOR_OVERFLOW = -10 /* k-NN with 32 < k <= 64: more than 512 points at one distance (exact duplicates) inside the cover radius.
                                     (pcp_radius_neighbors no longer returns it: lists beyond the 512-entry on-chip buffer are filled
                                     through global memory; records and normals of such neighbourhoods come from an unordered
                                     double-precision accumulation) */
} pcp_status;

#define PCP_MAX_K 64 /* k <= 32: the tile / warp-per-query k-NN kernels; 33..64 (SOR k = 60, region growing k = 50 upstream): cover-radius form, same results */

/* Record layout-identical to `pcaAnalysist::pcaFeature` (include/pcaAnalysis.h:34-45):
 * lamda{la1,la2,la3} ascending; vec{v1,v2,v3} as three packed 3-float columns;
 * ptID; ptNum; lineStruc; planeStruc; scatterStruc.  68 bytes, align 4. */
typedef struct pcp_feature68 {
  float la1, la2, la3;
  float v1[3], v2[3], v3[3];
  int32_t ptID;
  int32_t ptNum;
  float lineStruc, planeStruc, scatterStruc;
} pcp_feature68;

/* (nx,ny,nz,curvature): layout of pcl::Normal's first four floats, the output of
 * `preprocess::normalestimate` (src/smoothing.cpp:236-248). */
typedef struct pcp_normal16 {
  float nx, ny, nz, curvature;
} pcp_normal16;

/* Centre of the second moment. */
enum {
  PCP_CENTER_QUERY = 0, /* about the query point itself — src/pcaAnalysis.cpp:53-55 */
  PCP_CENTER_MEAN = 1   /* about the neighbourhood mean — pcl::NormalEstimation twin */
};

/* Stage indices for pcp_last_timings (milliseconds, CUDA events on the ctx stream). */
enum {
  PCP_T_BBOX = 0,
  PCP_T_KEYS = 1,
  PCP_T_SORT = 2,
  PCP_T_TABLE = 3,
  PCP_T_SEARCH_PCA = 4, /* the fused search + moment + eigen kernel */
  PCP_T_CARRY = 5,
  PCP_T_H2D = 6,
  PCP_T_D2H = 7,
  PCP_T_COUNT = 8
};

/* ---- lifetime ------------------------------------------------------------------- */

/* Creates a context on CUDA device `device`.  Fails with PCP_ERR_NO_DEVICE when the
 * device is missing or is not compute capability 10.x. */
PCP_API int pcp_create(pcp_ctx** out, int device);
PCP_API int pcp_destroy(pcp_ctx* ctx);
PCP_API const char* pcp_last_error(const pcp_ctx* ctx);
PCP_API const char* pcp_version(void);
/* cudaStream_t of the context, as an opaque pointer (for event timing by the caller). */
PCP_API void* pcp_stream(pcp_ctx* ctx);
PCP_API int pcp_synchronize(pcp_ctx* ctx);
/* Pipelining across clouds (frame batches): with `enable` != 0, pcp_set_cloud_host and pcp_knn_pca[_kth]
 * with host outputs only ENQUEUE their copies and kernels on the ctx stream and return; the caller must
 * pass page-locked host memory, keep it untouched until pcp_synchronize, and may drive a second ctx
 * meanwhile so that one cloud's result copy overlaps the next cloud's upload and search (pcp_ring_project
 * with host outputs follows the same rule).  Off by default
 * (the reference's calls are synchronous, src/pcaAnalysis.cpp:11-42).  pcp_build_grid still waits for
 * the bounding box (it sizes the grid on the host), pcp_radius_pca for its overflow flag. */
PCP_API int pcp_set_host_async(pcp_ctx* ctx, int enable);
/* Page-locked host buffers for that mode (cudaHostAlloc / cudaFreeHost); NULL when the allocation fails. */
PCP_API void* pcp_alloc_pinned(size_t bytes);
PCP_API void pcp_free_pinned(void* p);
/* Copies the PCP_T_COUNT stage times of the last build/compute calls into ms[0..n). */
PCP_API int pcp_last_timings(pcp_ctx* ctx, float* ms, int n);
/* Number of kernel launches issued by this ctx since creation. */
PCP_API int64_t pcp_launch_count(const pcp_ctx* ctx);

/* ---- cloud input ------------------------------------------------------------------
 * Replaces `pcl::KdTreeFLANN<pcl::PointXYZI>::setInputCloud` at src/pcaAnalysis.cpp:14-15
 * (the cloud hand-over half of it; the index build is pcp_build_grid). */

/* Host AoS cloud: `stride_bytes` between points, x,y,z as three leading floats
 * (32 for pcl::PointXYZI, 16 for a packed float4).  Copies to the device. */
PCP_API int pcp_set_cloud_host(pcp_ctx* ctx, const void* pts, int64_t n, int stride_bytes);
/* Device-resident packed float4 cloud (x,y,z,intensity); not copied, must stay alive
 * until the next pcp_set_cloud_* or pcp_destroy. */
PCP_API int pcp_set_cloud_device(pcp_ctx* ctx, const void* d_xyzi, int64_t n);
/* Sharded form (multi-GPU tiles, frame batches): the first `n_query` points are this
 * shard's own points (they get outputs); points [n_query, n) are halo points that only
 * serve as neighbours.  `d_global_id` (device, n × int32, may be NULL = identity) is the
 * index used for tie-breaking and written to ptID, so results are identical to the
 * unsharded run. */
PCP_API int pcp_set_cloud_device_sharded(pcp_ctx* ctx, const void* d_xyzi, const int32_t* d_global_id,
                                         int64_t n, int64_t n_query);
/* Frame batch (BASELINE configs[4]: many VLP-16 frames per launch): declares the current cloud to be
 * `n_frames` independent clouds laid end to end, frame f = points [frame_offsets[f], frame_offsets[f+1])
 * (HOST array of n_frames + 1 entries, first 0, last n; n_frames <= 65535).  The next pcp_build_grid
 * builds ONE segmented grid (cell key = frame ‖ cell over the common bounding box) and every search
 * stays inside the query's frame, so pcp_knn_pca / pcp_radius_pca / pcp_knn_normals give, per frame,
 * exactly the records of a separate run on that frame — the reference's per-frame loop
 * (src/lidarOdometry.cpp:66-76 builds one tree and searches every point of ONE frame) in one launch.  ptID and neighbour
 * indices count from the start of the batch; the stale-record carry restarts at every
 * frame.  Reset by the next pcp_set_cloud_*; not available for sharded clouds. */
PCP_API int pcp_set_frames(pcp_ctx* ctx, const int64_t* frame_offsets, int64_t n_frames);

/* ---- slab sharding: halo classification ----------------------------------------------
 * Multi-GPU form of the loop at src/pcaAnalysis.cpp:18-40 (the reference is one process):
 * a rank owns the slab [lo_face, hi_face) along `axis` (0/1/2 = x/y/z) and gives each slab neighbour
 * its points within `h` of the shared face.  Appends every own point with coord < lo_face + h to the
 * "lo" buffer and every point with coord >= hi_face - h to the "hi" buffer (float4 + global index,
 * `cap` entries each, device memory), pads the unused tails with non-finite points (global index
 * INT32_MAX), which no search ever returns, and leaves the true counts in d_counts[0..1] (a count
 * above cap means the buffer was too small: points were dropped).  d_global_id may be NULL (identity).
 * Enqueues on the ctx stream; no host synchronisation. */
PCP_API int pcp_halo_pack(pcp_ctx* ctx, const void* d_xyzi, const int32_t* d_global_id, int64_t n, int axis, float lo_face,
                          float hi_face, float h, int64_t cap, void* d_lo_pts, int32_t* d_lo_gid, void* d_hi_pts,
                          int32_t* d_hi_gid, uint32_t* d_counts);

/* Peer form of the exchange (no message, no NCCL launch): p_lo_* / p_hi_* point INTO the slab neighbours'
 * halo regions (memory of another GPU mapped with pcp_peer_open, or plain device memory), so the
 * classification kernel's stores are the transfer (NVLink peer writes).  When the strips and their padding
 * are complete, the last block raises *p_lo_flag and *p_hi_flag (either may be NULL) to `epoch` with
 * system-scope release stores; d_counts as in pcp_halo_pack.  The receiver enqueues pcp_halo_wait before
 * it reads its halo regions: one thread polls its own flags (either may be NULL) until they reached
 * `epoch`, and gives up after timeout_ms, leaving 1 in *d_err (device word, zeroed by the caller) instead
 * of hanging the GPU.  Callers alternate between two sets of halo regions (epoch parity) so that a
 * neighbour's step t+1 strip never lands in memory step t is still reading. */
PCP_API int pcp_halo_push(pcp_ctx* ctx, const void* d_xyzi, const int32_t* d_global_id, int64_t n, int axis, float lo_face,
                          float hi_face, float h, int64_t cap, void* p_lo_pts, int32_t* p_lo_gid, void* p_hi_pts,
                          int32_t* p_hi_gid, uint32_t* d_counts, uint32_t* p_lo_flag, uint32_t* p_hi_flag, uint32_t epoch);
PCP_API int pcp_halo_wait(pcp_ctx* ctx, const uint32_t* d_flag_a, const uint32_t* d_flag_b, uint32_t epoch, int timeout_ms,
                          uint32_t* d_err);
/* Device memory that other processes of the node can map: cudaMalloc + a 64-byte CUDA IPC handle to send to
 * them (any transport); pcp_peer_open maps a received handle (peer access enabled lazily), pcp_peer_close
 * unmaps it, pcp_peer_free releases an own allocation.  open/close/free synchronise the ctx stream. */
PCP_API int pcp_peer_alloc(pcp_ctx* ctx, size_t bytes, void** d_ptr, void* handle64);
PCP_API int pcp_peer_free(pcp_ctx* ctx, void* d_ptr);
PCP_API int pcp_peer_open(pcp_ctx* ctx, const void* handle64, void** d_ptr);
PCP_API int pcp_peer_close(pcp_ctx* ctx, void* d_ptr);

/* ---- index build ------------------------------------------------------------------
 * Replaces the FLANN KDTreeSingleIndex build inside setInputCloud (src/pcaAnalysis.cpp:15):
 * bbox reduce → cell keys → radix sort → cell table (prefix scan).  `cell` > 0 is the
 * cell edge in metres; for radius search it must be >= the radius (the 27-cell stencil
 * is then exact).  `cell` <= 0 picks one from the point density for a k-NN of `k_hint`. */
PCP_API int pcp_build_grid(pcp_ctx* ctx, float cell, int k_hint);
/* Grid facts for diagnostics: dims[3], cell edge, number of cells. */
PCP_API int pcp_grid_info(pcp_ctx* ctx, int32_t dims[3], float* cell, int64_t* n_cells);

/* ---- fused search + PCA -----------------------------------------------------------
 * Replaces the hot loop of `pcaAnalysist::calculatePCAofPointCloud`
 * (src/pcaAnalysis.cpp:18-40): radiusSearch (:25) + calculatePCAofPoint(int) (:44-87) +
 * the three structure features (:29-31), one 68-byte record per point, out[i].ptID == i.
 *   radius        : as the reference's float argument; r2 = (float)((double)r*(double)r),
 *                   membership is strict d2 < r2, the query is its own neighbour.
 *   min_pts       : 13 in the reference (:50); records with ptNum < min_pts keep only
 *                   ptID/ptNum valid unless carry_forward is set.
 *   carry_forward : reproduce the stale-`tempFea` behaviour (:16,:26,:39): an invalid
 *                   record repeats λ/vectors/features of the last valid lower-index
 *                   record (zeros before the first valid one).
 *   out           : n_query records, host (out_is_device = 0) or device memory. */
PCP_API int pcp_radius_pca(pcp_ctx* ctx, float radius, int min_pts, int center_mode,
                           int carry_forward, pcp_feature68* out, int out_is_device);

/* k-nearest variant (the search primitive behind `pcl::NormalEstimation::setKSearch`,
 * src/smoothing.cpp:239-246): the k nearest including the query itself, ordered by
 * (d2, index).  Writes the pcaFeature record and, when `normals` is non-NULL, the
 * (normal, curvature) of the PCL twin: moment about the mean, curvature =
 * |λmin / trace|, normal flipped toward the viewpoint (0,0,0). */
PCP_API int pcp_knn_pca(pcp_ctx* ctx, int k, int center_mode, pcp_feature68* out,
                        pcp_normal16* normals, int out_is_device);
/* Same, plus kth_d2[i] = squared distance of point i's k-th neighbour (+inf when the cloud holds
 * fewer than k points): a tile-sharded caller uses it to prove its halo was wide enough (a query
 * is safe when kth_d2 <= squared distance to the nearest un-covered tile face). */
PCP_API int pcp_knn_pca_kth(pcp_ctx* ctx, int k, int center_mode, pcp_feature68* out,
                            pcp_normal16* normals, float* kth_d2, int out_is_device);

/* ---- neighbour lists (verification and the consumers: mean-shift, DBSCAN) ----------
 * Replace `kdtree.nearestKSearch` / `kdtree.radiusSearch` call sites
 * (src/pcaAnalysis.cpp:25, src/buildingFacadeExtraction.cpp:83, src/lidarOdometry.cpp:76,
 * include/dbscanCLuster.h:156,282,292). */

/* idx/d2: n_query × k, row i = neighbours of point i in (d2, index) order. */
PCP_API int pcp_knn_neighbors(pcp_ctx* ctx, int k, int32_t* idx, float* d2, int out_is_device);
/* Two-call protocol: first call with idx == NULL fills counts[n_query]; the caller builds
 * offsets (exclusive prefix sum, n_query+1 entries) and calls again with idx/d2 sized
 * offsets[n_query].  Rows are in (d2, index) order. */
PCP_API int pcp_radius_neighbors(pcp_ctx* ctx, float radius, int32_t* counts, const int64_t* offsets,
                                 int32_t* idx, float* d2, int out_is_device);

/* ---- stale-record carry without any allocation ----------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the stale-record
 * carry that pcp_radius_pca(carry_forward = 1) applies (a record below the min_pts gate repeats everything but
 * ptID / ptNum of the last record at or above it; leading ones keep their own words), as a call of its own whose
 * scratch lives in a static device array inside libpcp.so instead of cudaMalloc'ed buffers:
 *   - d_in / d_out: DEVICE pointers to n records (they may be the same pointer); a record is valid when ptNum >= min_pts;
 *   - n <= pcp_carry_forward_static_max() = 1 048 576 records, else PCP_ERR_INVALID_ARG (4 MiB + 4 KiB of device
 *     memory are held by every GPU that loads the library, until exit);
 *   - asynchronous on the context's stream; calls no cudaMalloc / cudaFree / allocating routine;
 *   - every call on a GPU shares the scratch array: calls that overlap on one GPU (two contexts) race — serialise them;
 *   - one cloud per call (frame batches: call once per frame). */
PCP_API int pcp_carry_forward_static(pcp_ctx* ctx, const pcp_feature68* d_in, pcp_feature68* d_out, int64_t n, int min_pts);
PCP_API int64_t pcp_carry_forward_static_max(void);

/* ---- whole-cloud PCA bounding box ---------------------------------------------------
 * Replaces `pcaAnalysist::getMiniBoundingBox(cloud)` (src/pcaAnalysis.cpp:123-148): centroid
 * (pcl::compute3DCentroid, :125-126), moment about it with divisor n (:128-129), eigenvectors
 * ascending with the third axis replaced by col0 x col1 (:131-133), extents of the cloud after the
 * rigid transform [R | -R c] (:135-146) and meanDiag (:147).  The reference discards all of it;
 * here it is returned.  Device reductions over the current cloud (no grid needed); `out` is a host
 * pointer; the call synchronises. */
typedef struct pcp_obb {
  float centroid[3];
  float eigval[3];     /* ascending */
  float axes[3][3];    /* rows: eigenvector 0, eigenvector 1, their cross product (rows of the rotation R) */
  float min_pt[3];     /* extents of R (p - centroid) */
  float max_pt[3];
  float mean_diag[3];  /* 0.5 (max_pt + min_pt) */
  int64_t n_points;    /* finite points that entered the sums */
} pcp_obb;
PCP_API int pcp_bbox_pca(pcp_ctx* ctx, pcp_obb* out);

/* ---- segment PCA ------------------------------------------------------------------
 * Replaces `pcaAnalysist::calculatePCAofPoint(cloud, centro, idx, fea)`
 * (src/pcaAnalysis.cpp:89-121; caller src/buildingFacadeExtraction.cpp:212-219):
 * CSR segments of point indices into the current cloud, one centre per segment, no
 * minimum-count gate, ptID left 0.  Host pointers. */
PCP_API int pcp_segment_pca(pcp_ctx* ctx, const float* centres_xyz, const int64_t* seg_offsets,
                            const int32_t* seg_idx, int64_t n_seg, pcp_feature68* out);

/* ---- ring projection ----------------------------------------------------------------
 * Replaces `RingProjector::projectPointCloud` (src/ringProjection.cpp:24-81). */
typedef struct pcp_ring_calib {
  int32_t n_scan;       /* 16   (include/ringProjection.h:16) */
  int32_t horizon_scan; /* 1800 (include/ringProjection.h:17) */
  float ang_res_x;      /* 0.2  */
  float ang_res_y;      /* 2.0  */
  float ang_bottom;     /* 15.0 + 0.1 */
  float vert_deg[64];   /* per-laser vertical correction from 16db.xml (for frame synthesis) */
  float dist_lsb_cm;    /* 16db.xml:5 */
  int32_t n_enabled;
} pcp_ring_calib;

/* Reference constants (include/ringProjection.h:16-20) + the VLP-16 beam table. */
PCP_API int pcp_ring_calib_default(pcp_ring_calib* calib);
/* Parses the Boost-serialization XML calibration (16db.xml) on the host and derives
 * n_scan / ang_res_y / ang_bottom from the enabled lasers' vertical corrections. */
PCP_API int pcp_ring_calib_from_xml(const char* path, pcp_ring_calib* calib);

/* Projects `n_frames` frames stored back to back in the current cloud
 * (frame f = points [frame_offsets[f], frame_offsets[f+1]); frame_offsets is always a HOST
 * array of n_frames+1 entries; NULL means a single frame covering the whole cloud).  Per frame, host or device outputs:
 *   range_img  : n_scan × horizon_scan floats, FLT_MAX where empty   (rangeMat)
 *   pt_index   : n_scan × horizon_scan int32, −1 where empty, else the frame-local index
 *                of the LAST point in input order that fell in the cell (fullInfoCloud.x)
 *   long_range : per frame, count of in-image points with range > 60 (may be NULL). */
PCP_API int pcp_ring_project(pcp_ctx* ctx, const pcp_ring_calib* calib, const int64_t* frame_offsets,
                             int64_t n_frames, float* range_img, int32_t* pt_index,
                             int32_t* long_range, int out_is_device);

/* Constants of the segmentation stage (include/ringProjection.h:21-26). */
typedef struct pcp_ring_seg_params {
  int32_t ground_scan_ind;         /* 7      rings tested for ground, bottom-up */
  float sensor_mount_angle;        /* 0.0    degrees */
  float segment_theta;             /* 1.0472 growth threshold, radians */
  int32_t segment_valid_point_num; /* 5 */
  int32_t segment_valid_line_num;  /* 3 */
} pcp_ring_seg_params;
PCP_API int pcp_ring_seg_params_default(pcp_ring_seg_params* prm);

/* RingProjector::fullSegmentation (include/ringProjection.h:174-179) for `n_frames` frames of the current
 * cloud: projectPointCloud (as pcp_ring_project), then groundRemoval (src/ringProjection.cpp:83-136) and
 * cloudSegmentation / labelComponents (:138-289), one thread block per frame.  Outputs, per frame, host or
 * device (range_img, pt_index, long_range as in pcp_ring_project; any of them and of the following may be
 * NULL except label):
 *   ground      : n_scan × horizon int8   groundMat: 1 ground, -1 no ring pair, 0 else
 *   label       : n_scan × horizon int32  labelMat: -1 ground or empty, 1.. valid segments numbered in the
 *                 raster order of their first pixel, 999999 rejected segments
 *   col_max_ground : horizon floats       maxGroundrangeOfcol (never initialised upstream: starts at 0)
 *   above_idx   : n_scan × horizon int32  frame-local indices of the points of cloudaboveGround in the
 *                 reference's order, padded with -1;  above_count : their number
 *   n_labels    : number of valid segments (labelCount - 1)
 * The reference's flood grows right, +255 columns and down only (its neighbour table is unsigned,
 * include/ringProjection.h:69); that behaviour is kept.  At most 16 rings. */
PCP_API int pcp_ring_segment(pcp_ctx* ctx, const pcp_ring_calib* calib, const pcp_ring_seg_params* prm,
                             const int64_t* frame_offsets, int64_t n_frames, float* range_img, int32_t* pt_index,
                             int32_t* long_range, int8_t* ground, int32_t* label, float* col_max_ground, int32_t* above_idx,
                             int32_t* above_count, int32_t* n_labels, int out_is_device);

/* ---- Velodyne packet decode -------------------------------------------------------------------
 * Replaces the host-side decode the reference does with its prebuilt lib/libPacketDecoder.so
 * (PacketDecoder::DecodePacket / ProcessHDLPacket / PushFiringData / SplitFrame, include/PacketDecoder.h:79-121,
 * driven at src/pcapReader.cpp:79-102): `packets` = n_packets x 1206 bytes (12 blocks of {u16 0xEEFF|0xDDFF,
 * u16 azimuth in 0.01 deg, 32 x {u16 distance in 2 mm, u8 intensity}}, u32 timestamp, 2 bytes), host or
 * device.  Per return with a non-zero distance, in arrival order, the optional arrays of `out` (capacity
 * n_packets * 384 each, host or device, any member may be NULL) receive what PacketDecoder::HDLFrame holds
 * (include/PacketDecoder.h:84-94), doubles bit-identical to the reference binary's.  frame_offsets (HOST,
 * max_frames entries): frame f = points [frame_offsets[f], frame_offsets[f+1]); frames 0 .. *n_frames_complete-1
 * are complete revolutions (what GetFrames() returns), frame *n_frames_complete is the one still being filled.
 * The decoded points (x, y, z, intensity narrowed to float) become the CURRENT CLOUD of the context with its
 * non-empty frames declared as by pcp_set_frames, so pcp_ring_project / pcp_ring_segment (pass the same
 * frame_offsets) / pcp_build_grid / pcp_knn_pca follow without the points ever visiting the host.
 * Synchronises (the sizes are needed on the host). */
typedef struct pcp_decoded {
  double *x, *y, *z, *distance;
  uint8_t *intensity, *laser_id;
  uint16_t* azimuth;
  uint32_t* ms;
} pcp_decoded;
PCP_API int pcp_decode_packets(pcp_ctx* ctx, const pcp_ring_calib* calib, const void* packets, int64_t n_packets,
                               int packets_on_device, const pcp_decoded* out, int out_is_device, int64_t* frame_offsets,
                               int64_t max_frames, int64_t* n_frames_complete, int64_t* n_points);

/* ---- lidarOdometry consumers ------------------------------------------------
 * Per-point neighbour COUNT within `radius` (the query included; FLANN's strict fp32 test
 * d2 < (float)(r*r)), stopping at `cap` hits when cap > 0: counts[i] = min(count, cap).  This is all the
 * reference takes from its radiusSearch(i, 3.0) in getRangeAndptIdImage (`size() < 6`,
 * src/lidarOdometry.cpp:76-78) and from DBSCAN's core test (include/dbscanCLuster.h:282-284).
 * Needs a grid (any cell size; cell = radius is the cheap choice). */
PCP_API int pcp_radius_count(pcp_ctx* ctx, float radius, int cap, int32_t* counts, int out_is_device);

typedef struct pcp_odom_params {
  int32_t n_rows;          /* 16   src/lidarOdometry.cpp:29 */
  int32_t n_cols;          /* 900  :29 */
  double vert_res_deg;     /* 2.0  :85 */
  double hori_res_deg;     /* 0.2  :86 */
  float radius;            /* 3.0  :76 */
  int32_t min_neighbors;   /* 6    :77 */
  int32_t n_blocks;        /* 5    :131 */
  int32_t half_window;     /* 9    :133,141 */
  float near_range;        /* 1    :136,143 */
  int32_t min_valid;       /* 9    :149 */
  int32_t pick;            /* 3    thresholds sorted[size-3] and sorted[3], :167,170 */
  int32_t min_block_valid; /* 9    :162 */
} pcp_odom_params;
PCP_API int pcp_odom_params_default(pcp_odom_params* prm);

/* lidarOdometryClass::getRangeAndptIdImage (src/lidarOdometry.cpp:21-113) for `n_frames` frames of the current
 * cloud (frame_offsets: HOST array as in pcp_ring_project, NULL = one frame; for frame batches declare the same
 * frames with pcp_set_frames BEFORE pcp_build_grid so the neighbour counts stay inside a frame).  Needs a grid.
 * Per frame, host or device, row-major n_rows x n_cols (the reference's MatrixXf is column-major; contents equal):
 *   range_img : float, 0 where empty (rangeImg)           ptid_img : float point id, 0 where empty (ptIdImg; may be NULL)
 *   pt_index  : int32 frame-local id, -1 where empty (may be NULL)
 *   stats     : per frame {overwritten returns ("losing points num"), occupied cells}   (may be NULL)
 *   angle_minmax : per frame {vert min, vert max, hori min, hori max} in degrees (doubles; may be NULL)
 * Where the reference is undefined (row >= n_rows / col >= n_cols writes outside its matrix, int(NaN)) the
 * point is skipped. */
PCP_API int pcp_odom_range_image(pcp_ctx* ctx, const pcp_odom_params* prm, const int64_t* frame_offsets, int64_t n_frames,
                                 float* range_img, float* ptid_img, int32_t* pt_index, int32_t* stats, double* angle_minmax,
                                 int out_is_device);

/* lidarOdometryClass::calculateSmoothness (src/lidarOdometry.cpp:121-187): range images (n_frames x n_rows x
 * n_cols floats; host or device per range_is_device; NULL = the device images the last pcp_odom_range_image of
 * this context produced) -> smooth_img (smoothnessImg; may be NULL) and type_img (ptTypeImg: 0, 1 = among the
 * three roughest of its block, 2 = below the fourth smallest smoothness of its block), host or device.
 * Reads past the last column (:141-143 for the last 8 columns) see range 0. */
PCP_API int pcp_odom_smoothness(pcp_ctx* ctx, const pcp_odom_params* prm, const float* range_img, int range_is_device,
                                int64_t n_frames, float* smooth_img, float* type_img, int out_is_device);

/* ---- DBSCAN and mean shift ---------------------------------------------------
 * DynaTools::clusterDBSCAN's labelling (include/dbscanCLuster.h:256-300) on the current cloud and grid.  The
 * reference clusters a copy of the cloud with z set to 0 (:264-267): hand that flattened cloud to
 * pcp_set_cloud_* (the distance is then the planar one).  labels[i] = the reference's state[i]: cluster id 1..
 * (numbered by their first core point in index order) or -2 for noise; a border point belongs to the
 * first-seeded cluster that reaches it, and stays noise when its own index precedes that cluster's seed, as in
 * the reference's sequential walk.  *n_clusters = clusterID (HOST int32).  Synchronises. */
PCP_API int pcp_dbscan(pcp_ctx* ctx, float radius, int min_pts, int32_t* labels, int32_t* n_clusters, int out_is_device);
/* The same with the radius as the reference declares it: `double r` (include/dbscanCLuster.h:256; the call site passes
 * the literal 0.2, src/dynaObjTest.cpp:133).  PCL compares squared distances with (float)(r * r) evaluated in double:
 * for r = 0.2 that is 0x3d23d70a, one ulp below what the float 0.2f squares to — points whose squared distance is exactly
 * that float are neighbours for pcp_dbscan(0.2f) and not for the reference.  pcp_dbscan(r) == pcp_dbscan_d((double)r). */
PCP_API int pcp_dbscan_d(pcp_ctx* ctx, double radius, int min_pts, int32_t* labels, int32_t* n_clusters, int out_is_device);

/* Per cluster 1..n_clusters of `labels` (host or device per labels_is_device), over the CURRENT cloud (set the
 * original 3-D cloud again after pcp_dbscan): centroid (pcl::compute3DCentroid), min and max corner
 * (pcl::getMinMax3D) and member count — what include/dbscanCLuster.h:312-325 derives per cluster; outputs are HOST
 * arrays (n_clusters x 3 floats each, counts may be NULL).  The centroid is the double-precision mean rounded to
 * float (the reference adds in fp32 in index order: equal within fp32 rounding of the sum). */
PCP_API int pcp_cluster_boxes(pcp_ctx* ctx, const int32_t* labels, int labels_is_device, int32_t n_clusters, float* centroid_xyz,
                              float* min_xyz, float* max_xyz, int32_t* counts);

/* buildingFacadeExtractor::meanShiftClustering (src/buildingFacadeExtraction.cpp:59-132): `iterations` sweeps
 * (at least one: the reference loop is a do-while) over the current cloud, whose grid (any cell; cell = radius is
 * the cheap choice) indexes the ORIGINAL positions for every sweep, as the reference's kd-tree does.  The reference
 * uses radius 2.0 and min_pts 15 (:79-80).  start_xyzi: the positions the first sweep starts from (n x 4 floats,
 * host or device per start_is_device; NULL = the cloud itself).  out_xyzi: n x 4 floats (x, y, z moved, intensity
 * kept), host or device; counts (optional): neighbours found by the last sweep.  Not for frame batches. */
PCP_API int pcp_mean_shift(pcp_ctx* ctx, float radius, int min_pts, int iterations, const void* start_xyzi, int start_is_device,
                           void* out_xyzi, int32_t* counts, int out_is_device);

/* ---- text clouds ---------------------------------------------------------------
 * Replaces preprocess::txt2pc (src/smoothing.cpp:39-81): `text` = the bytes of an "x y z intensity" file (host or
 * device per text_is_device), parsed on the device with the reference's stream semantics — one point per line
 * (empty lines included), a field the line does not supply keeps the previous line's value, a field that does
 * not parse becomes 0 and ends the line, values are the correctly rounded floats strtof gives (exact for up to
 * 19 significant digits).  The points become the CURRENT CLOUD of the context; *n_points (HOST) receives their
 * number and out_xyzi (optional, n x 4 floats, host or device, capacity in points) a copy.  Synchronises.
 * preprocess::xyz2pc (:83-125) has its own entry point below. */
PCP_API int pcp_set_cloud_txt(pcp_ctx* ctx, const void* text, int64_t n_bytes, int text_is_device, int64_t* n_points,
                              float* out_xyzi, int64_t out_capacity, int out_is_device);
/* Replaces preprocess::xyz2pc (src/smoothing.cpp:83-125): `while (!eof()) { getline(line, 256); sscanf(line, "%f %f %f %s",
 * &x, &y, &z, &intens); intensity = intens; push_back }` — three floats and the FIRST BYTE of the fourth token as a
 * signed char (the reference scans %s into one char: the overflow is the one undefined part and is not reproduced);
 * a field that does not parse ends the line and keeps the previous line's value, as every field after it; a value
 * beyond the float range becomes +-inf; the eof() loop pushes one more copy of the last point after a final newline
 * (and one zero point for an empty file).  A line of 256 characters or more is PCP_ERR_INVALID_ARG (the reference
 * never returns from it).  inf / nan / hexadecimal fields are not recognised.  Same arguments as pcp_set_cloud_txt. */
PCP_API int pcp_set_cloud_xyz(pcp_ctx* ctx, const void* text, int64_t n_bytes, int text_is_device, int64_t* n_points,
                              float* out_xyzi, int64_t out_capacity, int out_is_device);

#ifdef __cplusplus
}
#endif
#endif /* PCP_H_ */
