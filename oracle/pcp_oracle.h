/*
 * pcp_oracle.h — CPU restatement of the reference's neighbourhood-PCA path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker or the timed CPU baseline.  libpcp.so never links or calls it.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or fixtures for this path
 *  and PCL 1.8.1 / FLANN 1.8.x / Eigen 3.3 are not vendored in the
 * reference tree nor installed in this image, so the upstream arithmetic is restated from
 * the published algorithms and pinned by brute-force cross-checks and analytic known-answer
 * cases (tests/test_oracle_*.py), not by outputs of the reference itself.
 * Exception — the packet-decode row (orc_decode_packets) IS pinned: the reference's prebuilt
 * lib/libPacketDecoder.so runs in the build container, oracle/ref_decoder/ drives it, and
 * tests/golden/decoder_*.npz hold its outputs (tests/test_golden.py).
 *
 * The lidarOdometry images (odom.cpp) and the DBSCAN / mean-shift consumers (consumers.cpp) are restated from the
 * reference's OWN sources (src/lidarOdometry.cpp, include/dbscanCLuster.h, src/buildingFacadeExtraction.cpp) on top
 * of the same search; their libm dependency (atanf) is pinned to this image's glibc on all 2^32 inputs.
 *
 * Clouds are packed float4 (x,y,z,intensity), n×4 floats.
 */
#ifndef PCP_ORACLE_H_
#define PCP_ORACLE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* same 68-byte layout as pcaAnalysist::pcaFeature (include/pcaAnalysis.h:34-45) */
typedef struct orc_feature68 {
  float la1, la2, la3;
  float v1[3], v2[3], v3[3];
  int32_t ptID, ptNum;
  float lineStruc, planeStruc, scatterStruc;
} orc_feature68;

/* r2 exactly as PCL hands it to FLANN: (float)((double)r*(double)r)   */
float orc_radius2(float radius);

/* ---- search: brute force (defines membership/order) and a FLANN-like kd-tree ---- */
/* neighbours of query q (an index into the cloud) with d2 < r2, sorted by (d2, idx).
 * Returns the count; writes at most cap entries. */
int64_t orc_radius_brute(const float* xyzi, int64_t n, int64_t q, float r2, int32_t* idx, float* d2, int64_t cap);
/* first k of the (d2, idx) order; returns min(k, n) */
int64_t orc_knn_brute(const float* xyzi, int64_t n, int64_t q, int k, int32_t* idx, float* d2);

typedef struct orc_kdtree orc_kdtree;
orc_kdtree* orc_kdtree_build(const float* xyzi, int64_t n, int leaf_max);
void orc_kdtree_free(orc_kdtree* t);
int64_t orc_kdtree_radius(const orc_kdtree* t, int64_t q, float r2, int32_t* idx, float* d2, int64_t cap);
int64_t orc_kdtree_knn(const orc_kdtree* t, int64_t q, int k, int32_t* idx, float* d2);
/* all queries, kNN lists (n×k), nthreads OpenMP threads */
void orc_kdtree_knn_all(const orc_kdtree* t, int k, int32_t* idx, float* d2, int nthreads);
/* per-query counts for a radius search (for CSR sizing) */
void orc_kdtree_radius_counts(const orc_kdtree* t, float r2, int32_t* counts, int nthreads);
/* CSR fill: offsets has n+1 entries */
void orc_kdtree_radius_all(const orc_kdtree* t, float r2, const int64_t* offsets, int32_t* idx, float* d2, int nthreads);

/* ---- numeric kernels ---- */
/* pcl::computeCovarianceMatrixNormalized(cloud, indices, centroid, C) — SURVEY A.2.
 * C is row-major 3×3. */
void orc_covariance_about(const float* xyzi, const int32_t* idx, int64_t m, const float centre[3], float C[9]);
/* Eigen::SelfAdjointEigenSolver<Matrix3f>::compute (tridiagonal QR) — SURVEY A.3.
 * evals ascending; evecs row-major, column k = eigenvector k.  Returns iterations used. */
int orc_eigen_sym3_qr(const float C[9], float evals[3], float evecs[9]);
/* pcl::eigen33 smallest eigenpair — SURVEY A.4 */
void orc_pcl_eigen33(const float C[9], float* eval, float evec[3]);
/* features of src/pcaAnalysis.cpp:29-31 on ascending λ */
void orc_structure_features(float la1, float la2, float la3, float out3[3]);

/* ---- drivers: restated reference path ---- */
/* pcaAnalysist::calculatePCAofPointCloud (src/pcaAnalysis.cpp:11-42): kd-tree build,
 * per point radius search, PCA about the query, features, stale-tempFea carry (Q4, leading
 * invalid entries zero-filled).  out has n records.  nthreads == 1 is the faithful
 * single-thread run; > 1 parallelises the loop and applies the carry afterwards.
 * valid[i] (optional) = 1 when ptNum >= min_pts. */
void orc_radius_pca_cloud(const float* xyzi, int64_t n, float radius, int min_pts, int carry_forward,
                          orc_feature68* out, uint8_t* valid, int nthreads);
/* k-NN flavour: neighbours = first k of (d2, idx); PCA about the query (center_mode 0) or
 * the mean (1); optional PCL-twin normals (nx,ny,nz,curvature) per point (n×4) */
void orc_knn_pca_cloud(const float* xyzi, int64_t n, int k, int center_mode, orc_feature68* out,
                       float* normals4, int nthreads);
/* pcaAnalysist::calculatePCAofPoint(cloud, centro, idx, fea) (src/pcaAnalysis.cpp:89-121) */
/* pcaAnalysist::getMiniBoundingBox (src/pcaAnalysis.cpp:123-148): 24 floats out (see pca.cpp) */
void orc_bbox_pca(const float* xyzi, int64_t n, float* out);
void orc_segment_pca(const float* xyzi, const float* centres_xyz, const int64_t* seg_offsets,
                     const int32_t* seg_idx, int64_t n_seg, orc_feature68* out);

/* RingProjector::projectPointCloud (src/ringProjection.cpp:24-81) for one frame.
 * range_img: n_scan*horizon floats (FLT_MAX init); pt_index: int32 (−1 init), last writer
 * wins; full_intensity (optional): the row + col/10000 tag per cell (NaN init... here −1).
 * Returns the long-range (>60) count. */
int32_t orc_ring_project(const float* xyzi, int64_t n, int n_scan, int horizon, float ang_res_x,
                         float ang_res_y, float ang_bottom, float* range_img, int32_t* pt_index,
                         float* full_intensity);

/* RingProjector::groundRemoval + cloudSegmentation + labelComponents (src/ringProjection.cpp:83-289) for one
 * frame, given its projection.  ground: int8 image (1 ground, -1 no pair, 0 else); label: int32 image
 * (-1 ground / empty, 1.. valid segments in seed raster order, 999999 rejected); col_max_ground: per column;
 * above_idx: original indices of the "cloud above ground" in raster order, -1 padded to n_scan*horizon.
 * Returns the above-ground count; *n_labels = number of valid segments. */
int32_t orc_ring_segment(const float* xyzi, int n_scan, int horizon, float ang_res_x, float ang_res_y, int ground_scan_ind,
                         float sensor_mount_angle, float segment_theta, int valid_point_num, int valid_line_num,
                         const float* range_img, const int32_t* pt_index, int8_t* ground, int32_t* label,
                         float* col_max_ground, int32_t* above_idx, int32_t* n_labels);

/* lidarOdometryClass::getRangeAndptIdImage (src/lidarOdometry.cpp:21-113) for one frame, given every point's
 * radius-search count (r = 3.0 upstream, computed by the caller with orc_kdtree_radius_counts): 16 x 900 range image
 * and point-id image (both 0 where empty, row-major), pt_index (optional, -1 where empty), stats2 = {overwrites
 * ("losing points"), occupied cells}, minmax4 = {vert min, vert max, hori min, hori max} in degrees. */
void orc_odom_range_image(const float* xyzi, int64_t n, const int32_t* nb_counts, int min_neighbors, int n_rows, int n_cols,
                          double vert_res, double hori_res, float* range_img, float* ptid_img, int32_t* pt_index,
                          int32_t* stats2, double* minmax4);
/* lidarOdometryClass::calculateSmoothness (src/lidarOdometry.cpp:121-187): smoothness image and point-type image
 * (0 none, 1 edge, 2 flat), row-major. */
void orc_odom_smoothness(const float* range_img, int n_rows, int n_cols, int n_blocks, int half_window, float near_range,
                         int min_valid, int pick, int min_block_valid, float* smooth_img, float* type_img);

/* inputs (every stride-th float bit pattern from start) where the fdlibm atanf the CUDA path evaluates differs
 * from this C library's atanf */
int64_t orc_atanf_mismatches(uint32_t start, uint32_t stride);

/* DynaTools::clusterDBSCAN (include/dbscanCLuster.h:256-300) on the cloud it searches in (z zeroed by the caller, as
 * :264-267): state[i] = cluster id 1.. or -2 (noise).  Returns the number of clusters. */
int32_t orc_dbscan(const float* xyzi, int64_t n, float radius, int min_pts, int32_t* state);
int32_t orc_dbscan_d(const float* xyzi, int64_t n, double radius, int min_pts, int32_t* state);  /* radius as the reference declares it */
/* per cluster: compute3DCentroid, getMinMax3D (:312-325) and the member count */
void orc_cluster_boxes(const float* xyzi, int64_t n, const int32_t* state, int32_t n_clusters, float* centroid3, float* min3,
                       float* max3, int32_t* counts);
/* one sweep of buildingFacadeExtractor::meanShiftClustering (src/buildingFacadeExtraction.cpp:72-123): tree on the
 * original positions, queries and differences from the current ones.  shift_xyz (n x 3) and counts optional. */
void orc_mean_shift_step(const float* orig_xyzi, const float* cur_xyzi, int64_t n, float radius, int min_pts, float* next_xyzi,
                         float* shift_xyz, int32_t* counts);

/* preprocess::txt2pc (src/smoothing.cpp:39-81): text -> points (x y z intensity per line), the reference's own
 * stream statements.  Returns the number of lines = points; writes at most `capacity` of them. */
int64_t orc_txt2pc(const char* text, int64_t n_bytes, float* out_xyzi, int64_t capacity);
/* preprocess::xyz2pc (src/smoothing.cpp:83-125): getline(line, 256) + sscanf("%f %f %f %s"); -1 for a line of 256+ characters (255 still fit: the delimiter is tested before the count, [istream.unformatted]) */
int64_t orc_xyz2pc(const char* text, int64_t n_bytes, float* out_xyzi, int64_t capacity);

int orc_num_threads(void);

/* Velodyne packet decode (lib/libPacketDecoder.so restated, see decoder.cpp): n_packets x 1206 bytes ->
 * per-return arrays in arrival order (zero distances dropped).  frame_offsets[f] = first point of frame f;
 * *n_frames = number of COMPLETE frames (the reference's GetFrames()); frame_offsets[*n_frames + 1] = total,
 * the trailing frame being the one still in progress.  Returns the number of points. */
int64_t orc_decode_packets(const uint8_t* packets, int64_t n_packets, const float* vert_deg, double* x, double* y, double* z,
                           double* distance, uint8_t* intensity, uint8_t* laser_id, uint16_t* azimuth, uint32_t* ms,
                           int64_t* frame_offsets, int64_t max_frames, int64_t* n_frames);

#ifdef __cplusplus
}
#endif
#endif
