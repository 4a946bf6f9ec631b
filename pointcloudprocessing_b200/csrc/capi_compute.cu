// capi_compute.cu — compute entry points of the pcp_* C-ABI (include/pcp.h).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#include "pcp_internal.cuh"
#include "search_common.cuh"

int pcp_segment_pca_impl(pcp_ctx* ctx, const float* d_centres, const int64_t* d_offs, const int32_t* d_idx, int64_t n_seg,
                         pcp_feature68* d_out);
int pcp_ring_project_impl(pcp_ctx* ctx, const pcp_ring_calib* calib, const int64_t* h_offsets, int64_t n_frames, float* d_range,
                          int32_t* d_index, int32_t* d_long);
int pcp_ring_segment_impl(pcp_ctx* ctx, const pcp_ring_calib* calib, const pcp_ring_seg_params* prm, const int64_t* d_offsets,
                          int64_t n_frames, const float* d_range, const int32_t* d_index, int8_t* d_ground, int32_t* d_label,
                          float* d_colmax, int32_t* d_above, int32_t* d_above_count, int32_t* d_n_labels);

int pcp_radius_count_impl(pcp_ctx* ctx, float r2, int R, int cap, int32_t* d_counts);
int pcp_odom_range_impl(pcp_ctx* ctx, const pcp_odom_params* prm, const int64_t* d_off, int64_t n_frames, int64_t max_len,
                        const int32_t* d_counts, float* d_range, float* d_ptid, int32_t* d_index, int32_t* d_stats, double* d_minmax);
int pcp_odom_smooth_impl(pcp_ctx* ctx, const pcp_odom_params* prm, const float* d_range, int64_t n_frames, float* d_smooth,
                         float* d_type);

int pcp_dbscan_impl(pcp_ctx* ctx, float r2, int R, int min_pts, int32_t* d_labels, int32_t* d_n_clusters);
int pcp_cluster_boxes_impl(pcp_ctx* ctx, const int32_t* d_labels, int n_clusters, float* d_centroid, float* d_lo, float* d_hi,
                           int32_t* d_counts);
int pcp_mean_shift_impl(pcp_ctx* ctx, const float4* d_cur, float r2, int R, int min_pts, float4* d_next, int32_t* d_counts);

int pcp_parse_txt_impl(pcp_ctx* ctx, const unsigned char* d_text, int64_t n_bytes, int64_t* n_lines_out, int xyz, int* too_long_out);

namespace {

int require_grid(pcp_ctx* ctx, const char* who) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, who);
  if (!ctx->grid_ok) return pcp_fail(ctx, PCP_ERR_NO_GRID, who);
  cudaError_t e = cudaSetDevice(ctx->device);
  if (e != cudaSuccess) return pcp_fail(ctx, PCP_ERR_CUDA, "cudaSetDevice", e);
  return PCP_OK;
}

// device output: the caller's pointer when out_is_device, else a ctx-owned buffer
template <typename T>
int pick_out(pcp_ctx* ctx, DevBuf& own, T* user, int out_is_device, size_t count, T** dev) {
  if (!user) {
    *dev = nullptr;
    return PCP_OK;
  }
  if (out_is_device) {
    *dev = user;
    return PCP_OK;
  }
  int rc = pcp_reserve(ctx, own, count * sizeof(T));
  if (rc) return rc;
  *dev = (T*)own.p;
  return PCP_OK;
}

template <typename T>
int copy_back(pcp_ctx* ctx, T* user, const T* dev, int out_is_device, size_t count) {
  if (!user || out_is_device || count == 0) return PCP_OK;
  PCP_CUDA(cudaMemcpyAsync(user, dev, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  return PCP_OK;
}

// 32 < k <= PCP_MAX_K: the k-NN kernels keep one key per lane, so these run through the radius machinery of
// the warp-per-query kernel ("k by cover radius", search_pca.cu) — same results, same tie rule
int knn_wide(pcp_ctx* ctx, SearchParams& sp) {
  if (sp.k <= 32) return PCP_OK;
  sp.mode = 1;
  sp.min_pts = 1;
  int rc = pcp_reserve(ctx, ctx->flags, sizeof(int));
  if (rc) return rc;
  sp.overflow_flag = (int*)ctx->flags.p;
  PCP_CUDA(cudaMemsetAsync(ctx->flags.p, 0, sizeof(int), ctx->stream));
  return PCP_OK;
}
int knn_wide_check(pcp_ctx* ctx, const char* who) {
  int h_flag = 0;
  PCP_CUDA(cudaMemcpyAsync(&h_flag, ctx->flags.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  if (h_flag) return pcp_fail(ctx, PCP_ERR_NEIGHBOR_OVERFLOW, who);
  return PCP_OK;
}

// smallest stencil half-width whose guaranteed coverage (R − 1e-6)·cell holds the radius
int stencil_for_radius(const GridDesc& g, float r2) {
  double r = std::sqrt((double)r2) * (1.0 + 2e-6);
  int R = (int)std::ceil(r * g.inv_cell + 2e-6);
  return R < 1 ? 1 : R;
}

}  // namespace

extern "C" {

PCP_API int pcp_radius_pca(pcp_ctx* ctx, float radius, int min_pts, int center_mode, int carry_forward, pcp_feature68* out,
                           int out_is_device) {
  int rc = require_grid(ctx, "pcp_radius_pca");
  if (rc) return rc;
  if (!out || !(radius > 0.f) || (center_mode != PCP_CENTER_QUERY && center_mode != PCP_CENTER_MEAN))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_radius_pca: bad argument");
  const int64_t nq = ctx->n_query;
  SearchParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.mode = 1;
  sp.r2 = (float)((double)radius * (double)radius);  // PCL: static_cast<float>(r*r) in double
  sp.R0 = stencil_for_radius(ctx->grid, sp.r2);
  sp.min_pts = min_pts;
  sp.center_mode = center_mode;
  if ((rc = pick_out(ctx, ctx->out_feat, out, out_is_device, (size_t)nq, &sp.out))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->valid, (size_t)nq))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->flags, sizeof(int)))) return rc;
  sp.valid = (uint8_t*)ctx->valid.p;
  sp.overflow_flag = (int*)ctx->flags.p;
  PCP_CUDA(cudaMemsetAsync(ctx->flags.p, 0, sizeof(int), ctx->stream));
  if ((rc = pcp_search_launch(ctx, sp))) return rc;
  if (carry_forward && (rc = pcp_carry_forward_impl(ctx, sp.out, sp.valid, nq))) return rc;
  int h_flag = 0;
  pcp_tic(ctx, PCP_T_D2H);
  if ((rc = copy_back(ctx, out, sp.out, out_is_device, (size_t)nq))) return rc;
  pcp_toc(ctx, PCP_T_D2H);
  PCP_CUDA(cudaMemcpyAsync(&h_flag, ctx->flags.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  if (h_flag) return pcp_fail(ctx, PCP_ERR_NEIGHBOR_OVERFLOW, "pcp_radius_pca: a neighbourhood exceeded the per-query buffer");
  return PCP_OK;
}

PCP_API int pcp_knn_pca(pcp_ctx* ctx, int k, int center_mode, pcp_feature68* out, pcp_normal16* normals, int out_is_device) {
  return pcp_knn_pca_kth(ctx, k, center_mode, out, normals, nullptr, out_is_device);
}

PCP_API int pcp_knn_pca_kth(pcp_ctx* ctx, int k, int center_mode, pcp_feature68* out, pcp_normal16* normals, float* kth_d2,
                            int out_is_device) {
  int rc = require_grid(ctx, "pcp_knn_pca");
  if (rc) return rc;
  if (k < 1 || k > PCP_MAX_K) return pcp_fail(ctx, PCP_ERR_K_UNSUPPORTED, "pcp_knn_pca: k outside [1, PCP_MAX_K]");
  if ((!out && !normals) || (center_mode != PCP_CENTER_QUERY && center_mode != PCP_CENTER_MEAN))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_knn_pca: bad argument");
  const int64_t nq = ctx->n_query;
  SearchParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.mode = 0;
  sp.k = k;
  sp.R0 = 1;
  sp.center_mode = center_mode;
  if ((rc = knn_wide(ctx, sp))) return rc;
  if ((rc = pick_out(ctx, ctx->out_feat, out, out_is_device, (size_t)nq, &sp.out))) return rc;
  if ((rc = pick_out(ctx, ctx->out_norm, normals, out_is_device, (size_t)nq, &sp.normals))) return rc;
  if ((rc = pick_out(ctx, ctx->out_d2, kth_d2, out_is_device, (size_t)nq, &sp.kth_d2))) return rc;
  if ((rc = pcp_search_launch(ctx, sp))) return rc;
  pcp_tic(ctx, PCP_T_D2H);
  if ((rc = copy_back(ctx, out, sp.out, out_is_device, (size_t)nq))) return rc;
  if ((rc = copy_back(ctx, normals, sp.normals, out_is_device, (size_t)nq))) return rc;
  if ((rc = copy_back(ctx, kth_d2, sp.kth_d2, out_is_device, (size_t)nq))) return rc;
  pcp_toc(ctx, PCP_T_D2H);
  if (sp.overflow_flag) return knn_wide_check(ctx, "pcp_knn_pca");
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_knn_neighbors(pcp_ctx* ctx, int k, int32_t* idx, float* d2, int out_is_device) {
  int rc = require_grid(ctx, "pcp_knn_neighbors");
  if (rc) return rc;
  if (k < 1 || k > PCP_MAX_K) return pcp_fail(ctx, PCP_ERR_K_UNSUPPORTED, "pcp_knn_neighbors: k outside [1, PCP_MAX_K]");
  if (!idx) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_knn_neighbors: idx is NULL");
  const size_t cnt = (size_t)ctx->n_query * (size_t)k;
  SearchParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.mode = 0;
  sp.k = k;
  sp.R0 = 1;
  if ((rc = knn_wide(ctx, sp))) return rc;
  if ((rc = pick_out(ctx, ctx->out_idx, idx, out_is_device, cnt, &sp.nn_idx))) return rc;
  if ((rc = pick_out(ctx, ctx->out_d2, d2, out_is_device, cnt, &sp.nn_d2))) return rc;
  if ((rc = pcp_search_launch(ctx, sp))) return rc;
  if ((rc = copy_back(ctx, idx, sp.nn_idx, out_is_device, cnt))) return rc;
  if ((rc = copy_back(ctx, d2, sp.nn_d2, out_is_device, cnt))) return rc;
  if (sp.overflow_flag) return knn_wide_check(ctx, "pcp_knn_neighbors");
  if (!out_is_device) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_radius_neighbors(pcp_ctx* ctx, float radius, int32_t* counts, const int64_t* offsets, int32_t* idx, float* d2,
                                 int out_is_device) {
  int rc = require_grid(ctx, "pcp_radius_neighbors");
  if (rc) return rc;
  if (!(radius > 0.f) || (!counts && !idx) || (idx && !offsets))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_radius_neighbors: bad argument");
  const int64_t nq = ctx->n_query;
  SearchParams sp;
  memset(&sp, 0, sizeof(sp));
  sp.mode = 1;
  sp.r2 = (float)((double)radius * (double)radius);
  sp.R0 = stencil_for_radius(ctx->grid, sp.r2);
  sp.min_pts = 0x7FFFFFFF;  // no PCA
  if ((rc = pcp_reserve(ctx, ctx->flags, sizeof(int)))) return rc;
  sp.overflow_flag = (int*)ctx->flags.p;
  PCP_CUDA(cudaMemsetAsync(ctx->flags.p, 0, sizeof(int), ctx->stream));
  size_t total = 0;
  const int64_t* d_offsets = offsets;
  if (idx) {
    if (out_is_device) {
      int64_t last = 0;
      PCP_CUDA(cudaMemcpyAsync(&last, offsets + nq, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
      PCP_CUDA(cudaStreamSynchronize(ctx->stream));
      total = (size_t)last;
    } else {
      total = (size_t)offsets[nq];
      if ((rc = pcp_reserve(ctx, ctx->staging, (size_t)(nq + 1) * sizeof(int64_t)))) return rc;
      PCP_CUDA(cudaMemcpyAsync(ctx->staging.p, offsets, (size_t)(nq + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
      d_offsets = (const int64_t*)ctx->staging.p;
    }
  }
  sp.offsets = idx ? d_offsets : nullptr;
  if ((rc = pick_out(ctx, ctx->out_counts, counts, out_is_device, (size_t)nq, &sp.counts))) return rc;
  if ((rc = pick_out(ctx, ctx->out_idx, idx, out_is_device, total, &sp.nn_idx))) return rc;
  if ((rc = pick_out(ctx, ctx->out_d2, d2, out_is_device, total, &sp.nn_d2))) return rc;
  if ((rc = pcp_search_launch(ctx, sp))) return rc;
  if (idx) {
    // neighbourhoods beyond the on-chip buffer (r = 2-3 m on a dense scan: src/lidarOdometry.cpp:76,
    // src/buildingFacadeExtraction.cpp:83) raised the flag and left their list slots empty: filled through global memory
    int h_flag = 0;
    PCP_CUDA(cudaMemcpyAsync(&h_flag, ctx->flags.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    PCP_CUDA(cudaStreamSynchronize(ctx->stream));
    if (h_flag) {
      KArgs a;
      pcp_make_kargs(ctx, sp, a);
      if ((rc = pcp_radius_heavy_lists_impl(ctx, a, sp.r2, sp.R0, d_offsets, (int64_t)total, sp.nn_idx, sp.nn_d2))) return rc;
    }
  }
  if ((rc = copy_back(ctx, counts, sp.counts, out_is_device, (size_t)nq))) return rc;
  if ((rc = copy_back(ctx, idx, sp.nn_idx, out_is_device, total))) return rc;
  if ((rc = copy_back(ctx, d2, sp.nn_d2, out_is_device, total))) return rc;
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_segment_pca(pcp_ctx* ctx, const float* centres_xyz, const int64_t* seg_offsets, const int32_t* seg_idx, int64_t n_seg,
                            pcp_feature68* out) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_segment_pca: no cloud set");
  if (n_seg < 0 || (n_seg > 0 && (!centres_xyz || !seg_offsets || !seg_idx || !out)))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_segment_pca: bad argument");
  if (n_seg == 0) return PCP_OK;
  PCP_CUDA(cudaSetDevice(ctx->device));
  const size_t n_idx = (size_t)seg_offsets[n_seg];
  for (size_t i = 0; i < n_idx; ++i)
    if (seg_idx[i] < 0 || seg_idx[i] >= ctx->n) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_segment_pca: index out of range");
  const size_t b_cent = (size_t)n_seg * 3 * sizeof(float), b_off = (size_t)(n_seg + 1) * sizeof(int64_t), b_idx = n_idx * sizeof(int32_t);
  const size_t o_off = (b_cent + 15) & ~(size_t)15, o_idx = (o_off + b_off + 15) & ~(size_t)15;
  int rc = pcp_reserve(ctx, ctx->staging, o_idx + b_idx + 16);
  if (rc) return rc;
  if ((rc = pcp_reserve(ctx, ctx->out_feat, (size_t)n_seg * sizeof(pcp_feature68)))) return rc;
  char* st = (char*)ctx->staging.p;
  PCP_CUDA(cudaMemcpyAsync(st, centres_xyz, b_cent, cudaMemcpyHostToDevice, ctx->stream));
  PCP_CUDA(cudaMemcpyAsync(st + o_off, seg_offsets, b_off, cudaMemcpyHostToDevice, ctx->stream));
  if (b_idx) PCP_CUDA(cudaMemcpyAsync(st + o_idx, seg_idx, b_idx, cudaMemcpyHostToDevice, ctx->stream));
  rc = pcp_segment_pca_impl(ctx, (const float*)st, (const int64_t*)(st + o_off), (const int32_t*)(st + o_idx), n_seg,
                            (pcp_feature68*)ctx->out_feat.p);
  if (rc) return rc;
  PCP_CUDA(cudaMemcpyAsync(out, ctx->out_feat.p, (size_t)n_seg * sizeof(pcp_feature68), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

// ---- ring projection -------------------------------------------------------------------------

PCP_API int pcp_ring_calib_default(pcp_ring_calib* c) {
  if (!c) return PCP_ERR_INVALID_ARG;
  memset(c, 0, sizeof(*c));
  c->n_scan = 16;
  c->horizon_scan = 1800;
  c->ang_res_x = 0.2f;
  c->ang_res_y = 2.0f;
  c->ang_bottom = (float)(15.0 + 0.1);
  c->dist_lsb_cm = 0.2f;
  c->n_enabled = 16;
  for (int i = 0; i < 64; ++i) c->vert_deg[i] = 0.f;
  for (int i = 0; i < 32; ++i) {
    int j = i % 16;
    c->vert_deg[i] = (j % 2 == 0) ? (float)(-15 + j) : (float)(j);  // −15,1,−13,3,…,−1,15
  }
  return PCP_OK;
}

PCP_API int pcp_ring_calib_from_xml(const char* path, pcp_ring_calib* c) {
  if (!path || !c) return PCP_ERR_INVALID_ARG;
  FILE* f = fopen(path, "rb");
  if (!f) return PCP_ERR_IO;
  std::string s;
  char chunk[65536];
  size_t got;
  while ((got = fread(chunk, 1, sizeof(chunk), f)) > 0) s.append(chunk, got);
  fclose(f);
  pcp_ring_calib_default(c);
  auto tag_value = [&](const char* tag, size_t from, size_t* after) -> const char* {
    std::string open = std::string("<") + tag + ">";
    size_t p = s.find(open, from);
    if (p == std::string::npos) return nullptr;
    if (after) *after = p + open.size();
    return s.c_str() + p + open.size();
  };
  size_t pos = 0;
  const char* v = tag_value("distLSB_", 0, &pos);
  if (!v) return PCP_ERR_IO;
  c->dist_lsb_cm = (float)atof(v);
  // enabled_ : <count>64</count> then 64 <item>
  size_t en = s.find("<enabled_>");
  size_t en_end = s.find("</enabled_>");
  if (en == std::string::npos || en_end == std::string::npos) return PCP_ERR_IO;
  int enabled[64] = {0};
  int ne = 0;
  for (size_t p = s.find("<item>", en); p != std::string::npos && p < en_end && ne < 64; p = s.find("<item>", p + 6))
    enabled[ne++] = atoi(s.c_str() + p + 6);
  // per-laser vertical corrections in file order, with their ids
  int n_px = 0;
  size_t p = 0;
  for (;;) {
    size_t idp;
    const char* idv = tag_value("id_", p, &idp);
    if (!idv) break;
    size_t vp;
    const char* vv = tag_value("vertCorrection_", idp, &vp);
    if (!vv) break;
    int id = atoi(idv);
    if (id >= 0 && id < 64) c->vert_deg[id] = (float)atof(vv);
    p = vp;
    ++n_px;
  }
  if (n_px == 0) return PCP_ERR_IO;
  float lo = 1e9f, hi = -1e9f;
  int n_en = 0;
  for (int i = 0; i < 64; ++i)
    if (enabled[i]) {
      ++n_en;
      lo = fminf(lo, c->vert_deg[i]);
      hi = fmaxf(hi, c->vert_deg[i]);
    }
  if (n_en < 2) return PCP_ERR_IO;
  c->n_enabled = n_en;
  c->n_scan = n_en;
  c->ang_res_y = (hi - lo) / (float)(n_en - 1);
  c->ang_bottom = -lo + 0.1f;  // the reference's literal 15.0+0.1 (include/ringProjection.h:20)
  return PCP_OK;
}

PCP_API int pcp_ring_project(pcp_ctx* ctx, const pcp_ring_calib* calib, const int64_t* frame_offsets, int64_t n_frames,
                             float* range_img, int32_t* pt_index, int32_t* long_range, int out_is_device) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_ring_project: no cloud set");
  if (!calib || !range_img || !pt_index || n_frames < 1 || n_frames > 65535 || calib->n_scan < 1 || calib->horizon_scan < 1)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_project: bad argument");
  if (!frame_offsets && n_frames != 1) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_project: frame_offsets required");
  if (frame_offsets && (frame_offsets[0] < 0 || frame_offsets[n_frames] > ctx->n))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_project: frame_offsets outside the cloud");
  PCP_CUDA(cudaSetDevice(ctx->device));
  const size_t cells = (size_t)calib->n_scan * calib->horizon_scan * (size_t)n_frames;
  float* d_range;
  int32_t *d_index, *d_long;
  int rc;
  if ((rc = pick_out(ctx, ctx->ring_range, range_img, out_is_device, cells, &d_range))) return rc;
  if ((rc = pick_out(ctx, ctx->ring_idx, pt_index, out_is_device, cells, &d_index))) return rc;
  if ((rc = pick_out(ctx, ctx->ring_long, long_range, out_is_device, (size_t)n_frames, &d_long))) return rc;
  if ((rc = pcp_ring_project_impl(ctx, calib, frame_offsets, n_frames, d_range, d_index, d_long))) return rc;
  if ((rc = copy_back(ctx, range_img, d_range, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, pt_index, d_index, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, long_range, d_long, out_is_device, (size_t)n_frames))) return rc;
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_ring_seg_params_default(pcp_ring_seg_params* prm) {
  if (!prm) return PCP_ERR_INVALID_ARG;
  prm->ground_scan_ind = 7;
  prm->sensor_mount_angle = 0.0f;
  prm->segment_theta = 1.0472f;
  prm->segment_valid_point_num = 5;
  prm->segment_valid_line_num = 3;
  return PCP_OK;
}

PCP_API int pcp_ring_segment(pcp_ctx* ctx, const pcp_ring_calib* calib, const pcp_ring_seg_params* prm, const int64_t* frame_offsets,
                             int64_t n_frames, float* range_img, int32_t* pt_index, int32_t* long_range, int8_t* ground, int32_t* label,
                             float* col_max_ground, int32_t* above_idx, int32_t* above_count, int32_t* n_labels, int out_is_device) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_ring_segment: no cloud set");
  if (!calib || !prm || !label || n_frames < 1 || n_frames > 65535 || calib->n_scan < 1 || calib->horizon_scan < 1 ||
      prm->ground_scan_ind < 0 || prm->ground_scan_ind >= calib->n_scan)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_segment: bad argument");
  if (!frame_offsets && n_frames != 1) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_segment: frame_offsets required");
  if (frame_offsets && (frame_offsets[0] < 0 || frame_offsets[n_frames] > ctx->n))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_segment: frame_offsets outside the cloud");
  PCP_CUDA(cudaSetDevice(ctx->device));
  const size_t cells = (size_t)calib->n_scan * calib->horizon_scan * (size_t)n_frames;
  const size_t cols = (size_t)calib->horizon_scan * (size_t)n_frames;
  // every stage output lives on the device (the caller's buffer when it is a device pointer, else ctx-owned)
  float *d_range, *d_colmax;
  int32_t *d_index, *d_long, *d_label, *d_above, *d_acount, *d_nlab;
  int8_t* d_ground;
  int rc;
  auto need = [&](auto*& dev, DevBuf& own, auto* user, size_t count) -> int {
    using T = std::remove_pointer_t<std::remove_reference_t<decltype(dev)>>;
    if (user && out_is_device) {
      dev = user;
      return PCP_OK;
    }
    int r = pcp_reserve(ctx, own, count * sizeof(T));
    dev = (T*)own.p;
    return r;
  };
  if ((rc = need(d_range, ctx->ring_range, range_img, cells))) return rc;
  if ((rc = need(d_index, ctx->ring_idx, pt_index, cells))) return rc;
  if ((rc = need(d_long, ctx->ring_long, long_range, (size_t)n_frames))) return rc;
  if ((rc = need(d_ground, ctx->seg_ground, ground, cells))) return rc;
  if ((rc = need(d_label, ctx->seg_label, label, cells))) return rc;
  if ((rc = need(d_colmax, ctx->seg_colmax, col_max_ground, cols))) return rc;
  if ((rc = need(d_above, ctx->seg_above, above_idx, cells))) return rc;
  if ((rc = need(d_acount, ctx->seg_counts, above_count, (size_t)n_frames))) return rc;
  if ((rc = need(d_nlab, ctx->seg_nlabels, n_labels, (size_t)n_frames))) return rc;
  if ((rc = pcp_ring_project_impl(ctx, calib, frame_offsets, n_frames, d_range, d_index, d_long))) return rc;
  if ((rc = pcp_ring_segment_impl(ctx, calib, prm, frame_offsets ? (const int64_t*)ctx->ring_offsets.p : nullptr, n_frames, d_range,
                                  d_index, d_ground, d_label, d_colmax, d_above, d_acount, d_nlab)))
    return rc;
  if ((rc = copy_back(ctx, range_img, d_range, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, pt_index, d_index, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, long_range, d_long, out_is_device, (size_t)n_frames))) return rc;
  if ((rc = copy_back(ctx, ground, d_ground, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, label, d_label, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, col_max_ground, d_colmax, out_is_device, cols))) return rc;
  if ((rc = copy_back(ctx, above_idx, d_above, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, above_count, d_acount, out_is_device, (size_t)n_frames))) return rc;
  if ((rc = copy_back(ctx, n_labels, d_nlab, out_is_device, (size_t)n_frames))) return rc;
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_radius_count(pcp_ctx* ctx, float radius, int cap, int32_t* counts, int out_is_device) {
  int rc = require_grid(ctx, "pcp_radius_count");
  if (rc) return rc;
  if (!(radius > 0.f) || !counts) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_radius_count: bad argument");
  const float r2 = (float)((double)radius * (double)radius);
  int32_t* d_counts;
  if ((rc = pick_out(ctx, ctx->odom_counts, counts, out_is_device, (size_t)ctx->n_query, &d_counts))) return rc;
  if ((rc = pcp_radius_count_impl(ctx, r2, stencil_for_radius(ctx->grid, r2), cap, d_counts))) return rc;
  if ((rc = copy_back(ctx, counts, d_counts, out_is_device, (size_t)ctx->n_query))) return rc;
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_odom_params_default(pcp_odom_params* prm) {
  if (!prm) return PCP_ERR_INVALID_ARG;
  prm->n_rows = 16;
  prm->n_cols = 900;
  prm->vert_res_deg = 2.0;
  prm->hori_res_deg = 0.2;
  prm->radius = 3.0f;
  prm->min_neighbors = 6;
  prm->n_blocks = 5;
  prm->half_window = 9;
  prm->near_range = 1.0f;
  prm->min_valid = 9;
  prm->pick = 3;
  prm->min_block_valid = 9;
  return PCP_OK;
}

static bool odom_params_ok(const pcp_odom_params* p) {
  return p && p->n_rows >= 1 && p->n_cols >= 32 && p->n_blocks >= 1 && p->n_blocks <= p->n_cols && p->half_window >= 0 &&
         p->n_cols / p->n_blocks - p->half_window >= 1 && p->n_cols / p->n_blocks - p->half_window <= 256 && p->pick >= 0 &&
         p->min_block_valid >= 2 * p->pick && p->vert_res_deg > 0 && p->hori_res_deg > 0 && p->radius > 0.f;
}

PCP_API int pcp_odom_range_image(pcp_ctx* ctx, const pcp_odom_params* prm, const int64_t* frame_offsets, int64_t n_frames,
                                 float* range_img, float* ptid_img, int32_t* pt_index, int32_t* stats, double* angle_minmax,
                                 int out_is_device) {
  int rc = require_grid(ctx, "pcp_odom_range_image");
  if (rc) return rc;
  if (!odom_params_ok(prm) || !range_img || n_frames < 1 || n_frames > 65535)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_range_image: bad argument");
  if (ctx->d_gid) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_range_image: not defined for shards of a cloud");
  if (!frame_offsets && n_frames != 1) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_range_image: frame_offsets required");
  if (frame_offsets && (frame_offsets[0] < 0 || frame_offsets[n_frames] > ctx->n))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_range_image: frame_offsets outside the cloud");
  const size_t cells = (size_t)prm->n_rows * prm->n_cols * (size_t)n_frames;
  int64_t max_len = ctx->n;
  const int64_t* d_off = nullptr;
  if (frame_offsets) {
    max_len = 0;
    for (int64_t f = 0; f < n_frames; ++f) {
      if (frame_offsets[f + 1] < frame_offsets[f]) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_range_image: offsets decrease");
      max_len = std::max(max_len, frame_offsets[f + 1] - frame_offsets[f]);
    }
    if ((rc = pcp_reserve(ctx, ctx->ring_offsets, (n_frames + 1) * sizeof(int64_t)))) return rc;
    PCP_CUDA(cudaMemcpyAsync(ctx->ring_offsets.p, frame_offsets, (n_frames + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
    d_off = (const int64_t*)ctx->ring_offsets.p;
  }
  float *d_range, *d_ptid;
  int32_t *d_index, *d_stats;
  double* d_minmax;
  auto need = [&](auto*& dev, DevBuf& own, auto* user, size_t count) -> int {
    using T = std::remove_pointer_t<std::remove_reference_t<decltype(dev)>>;
    if (user && out_is_device) {
      dev = user;
      return PCP_OK;
    }
    int r = pcp_reserve(ctx, own, count * sizeof(T));
    dev = (T*)own.p;
    return r;
  };
  if ((rc = need(d_range, ctx->odom_range, range_img, cells))) return rc;
  if ((rc = need(d_index, ctx->odom_index, pt_index, cells))) return rc;
  if ((rc = need(d_stats, ctx->odom_stats, stats, 2 * (size_t)n_frames))) return rc;
  if ((rc = need(d_minmax, ctx->odom_minmax, angle_minmax, 4 * (size_t)n_frames))) return rc;
  d_ptid = nullptr;
  if (ptid_img && (rc = need(d_ptid, ctx->odom_ptid, ptid_img, cells))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->odom_counts, (size_t)ctx->n * sizeof(int32_t)))) return rc;
  int32_t* d_counts = (int32_t*)ctx->odom_counts.p;
  const float r2 = (float)((double)prm->radius * (double)prm->radius);
  if ((rc = pcp_radius_count_impl(ctx, r2, stencil_for_radius(ctx->grid, r2), prm->min_neighbors, d_counts))) return rc;
  if ((rc = pcp_odom_range_impl(ctx, prm, d_off, n_frames, max_len, d_counts, d_range, d_ptid, d_index, d_stats, d_minmax))) return rc;
  ctx->odom_last_range = d_range;
  ctx->odom_last_frames = n_frames;
  if ((rc = copy_back(ctx, range_img, d_range, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, ptid_img, d_ptid, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, pt_index, d_index, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, stats, d_stats, out_is_device, 2 * (size_t)n_frames))) return rc;
  if ((rc = copy_back(ctx, angle_minmax, d_minmax, out_is_device, 4 * (size_t)n_frames))) return rc;
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_odom_smoothness(pcp_ctx* ctx, const pcp_odom_params* prm, const float* range_img, int range_is_device,
                                int64_t n_frames, float* smooth_img, float* type_img, int out_is_device) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!odom_params_ok(prm) || !type_img || n_frames < 1 || n_frames > 65535)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_smoothness: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  const size_t cells = (size_t)prm->n_rows * prm->n_cols * (size_t)n_frames;
  const float* d_range = range_img;
  int rc;
  if (!range_img) {
    if (!ctx->odom_last_range || ctx->odom_last_frames != n_frames)
      return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_odom_smoothness: no range images of that many frames in this context");
    d_range = ctx->odom_last_range;
  } else if (!range_is_device) {
    if ((rc = pcp_reserve(ctx, ctx->odom_range, cells * sizeof(float)))) return rc;
    PCP_CUDA(cudaMemcpyAsync(ctx->odom_range.p, range_img, cells * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    d_range = (const float*)ctx->odom_range.p;
    ctx->odom_last_range = d_range;
    ctx->odom_last_frames = n_frames;
  }
  float *d_smooth, *d_type;
  if ((rc = pick_out(ctx, ctx->odom_smooth, smooth_img, out_is_device, cells, &d_smooth))) return rc;
  if ((rc = pick_out(ctx, ctx->odom_type, type_img, out_is_device, cells, &d_type))) return rc;
  if ((rc = pcp_odom_smooth_impl(ctx, prm, d_range, n_frames, d_smooth, d_type))) return rc;
  if ((rc = copy_back(ctx, smooth_img, d_smooth, out_is_device, cells))) return rc;
  if ((rc = copy_back(ctx, type_img, d_type, out_is_device, cells))) return rc;
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_dbscan(pcp_ctx* ctx, float radius, int min_pts, int32_t* labels, int32_t* n_clusters, int out_is_device) {
  return pcp_dbscan_d(ctx, (double)radius, min_pts, labels, n_clusters, out_is_device);
}

PCP_API int pcp_dbscan_d(pcp_ctx* ctx, double radius, int min_pts, int32_t* labels, int32_t* n_clusters, int out_is_device) {
  int rc = require_grid(ctx, "pcp_dbscan");
  if (rc) return rc;
  if (!(radius > 0.0) || min_pts < 1 || !labels || !n_clusters) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_dbscan: bad argument");
  if (ctx->d_gid) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_dbscan: not defined for shards of a cloud");
  const float r2 = (float)(radius * radius);  // PCL: static_cast<float>(r * r) with the caller's double r
  int32_t* d_labels;
  if ((rc = pick_out(ctx, ctx->db_labels, labels, out_is_device, (size_t)ctx->n, &d_labels))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->flags, sizeof(int)))) return rc;
  if ((rc = pcp_dbscan_impl(ctx, r2, stencil_for_radius(ctx->grid, r2), min_pts, d_labels, (int32_t*)ctx->flags.p))) return rc;
  if ((rc = copy_back(ctx, labels, d_labels, out_is_device, (size_t)ctx->n))) return rc;
  PCP_CUDA(cudaMemcpyAsync(n_clusters, ctx->flags.p, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_cluster_boxes(pcp_ctx* ctx, const int32_t* labels, int labels_is_device, int32_t n_clusters, float* centroid_xyz,
                              float* min_xyz, float* max_xyz, int32_t* counts) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_cluster_boxes: no cloud set");
  if (!labels || n_clusters < 0 || (n_clusters > 0 && (!centroid_xyz || !min_xyz || !max_xyz)))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_cluster_boxes: bad argument");
  if (n_clusters == 0) return PCP_OK;
  PCP_CUDA(cudaSetDevice(ctx->device));
  int rc;
  const int32_t* d_labels = labels;
  if (!labels_is_device) {
    if ((rc = pcp_reserve(ctx, ctx->db_labels, (size_t)ctx->n * sizeof(int32_t)))) return rc;
    PCP_CUDA(cudaMemcpyAsync(ctx->db_labels.p, labels, (size_t)ctx->n * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    d_labels = (const int32_t*)ctx->db_labels.p;
  }
  const size_t nc = (size_t)n_clusters;
  if ((rc = pcp_reserve(ctx, ctx->db_out, nc * 10 * sizeof(float)))) return rc;
  float* o = (float*)ctx->db_out.p;
  if ((rc = pcp_cluster_boxes_impl(ctx, d_labels, n_clusters, o, o + 3 * nc, o + 6 * nc, (int32_t*)(o + 9 * nc)))) return rc;
  PCP_CUDA(cudaMemcpyAsync(centroid_xyz, o, nc * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaMemcpyAsync(min_xyz, o + 3 * nc, nc * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaMemcpyAsync(max_xyz, o + 6 * nc, nc * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  if (counts) PCP_CUDA(cudaMemcpyAsync(counts, o + 9 * nc, nc * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_mean_shift(pcp_ctx* ctx, float radius, int min_pts, int iterations, const void* start_xyzi, int start_is_device,
                           void* out_xyzi, int32_t* counts, int out_is_device) {
  int rc = require_grid(ctx, "pcp_mean_shift");
  if (rc) return rc;
  if (!(radius > 0.f) || !out_xyzi) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_mean_shift: bad argument");
  if (ctx->d_gid || ctx->grid.n_seg != 1) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_mean_shift: single unsharded cloud only");
  const size_t n = (size_t)ctx->n;
  const float r2 = (float)((double)radius * (double)radius);
  const int R = stencil_for_radius(ctx->grid, r2);
  if (iterations < 1) iterations = 1;  // do { ... } while (times < iterTimes)
  if ((rc = pcp_reserve(ctx, ctx->ms_a, n * sizeof(float4)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->ms_b, n * sizeof(float4)))) return rc;
  int32_t* d_counts;
  if ((rc = pick_out(ctx, ctx->odom_counts, counts, out_is_device, n, &d_counts))) return rc;
  const float4* cur = ctx->d_cloud;
  if (start_xyzi) {
    if (start_is_device) {
      cur = (const float4*)start_xyzi;
    } else {
      PCP_CUDA(cudaMemcpyAsync(ctx->ms_b.p, start_xyzi, n * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
      cur = (const float4*)ctx->ms_b.p;
    }
  }
  float4* bufs[2] = {(float4*)ctx->ms_a.p, (float4*)ctx->ms_b.p};
  for (int it = 0; it < iterations; ++it) {
    const bool last = it == iterations - 1;
    float4* nxt = (last && out_is_device) ? (float4*)out_xyzi : bufs[it & 1];
    if ((rc = pcp_mean_shift_impl(ctx, cur, r2, R, min_pts, nxt, last ? d_counts : nullptr))) return rc;
    cur = nxt;
  }
  if (!out_is_device) PCP_CUDA(cudaMemcpyAsync(out_xyzi, cur, n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
  if ((rc = copy_back(ctx, counts, d_counts, out_is_device, n))) return rc;
  if (!out_is_device && !ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

static int set_cloud_text(pcp_ctx* ctx, const void* text, int64_t n_bytes, int text_is_device, int64_t* n_points, float* out_xyzi,
                          int64_t out_capacity, int out_is_device, int xyz, const char* who) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (n_bytes < 0 || (n_bytes > 0 && !text) || !n_points || (out_xyzi && out_capacity < 0)) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, who);
  PCP_CUDA(cudaSetDevice(ctx->device));
  *n_points = 0;
  int rc;
  if (n_bytes == 0) {
    if (!xyz) return PCP_OK;
    // xyz2pc on an empty file: eof() is not set before the first getline, so ONE (zero) point is pushed
    if ((rc = pcp_reserve(ctx, ctx->txt_out, sizeof(float4)))) return rc;
    PCP_CUDA(cudaMemsetAsync(ctx->txt_out.p, 0, sizeof(float4), ctx->stream));
    *n_points = 1;
  } else {
    const unsigned char* d_text = (const unsigned char*)text;
    if (!text_is_device) {
      if ((rc = pcp_reserve(ctx, ctx->txt_in, (size_t)n_bytes))) return rc;
      PCP_CUDA(cudaMemcpyAsync(ctx->txt_in.p, text, (size_t)n_bytes, cudaMemcpyHostToDevice, ctx->stream));
      d_text = (const unsigned char*)ctx->txt_in.p;
    }
    int64_t n = 0;
    int too_long = 0;
    if ((rc = pcp_parse_txt_impl(ctx, d_text, n_bytes, &n, xyz, &too_long))) return rc;
    if (xyz) {
      PCP_CUDA(cudaStreamSynchronize(ctx->stream));
      if (too_long)
        return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_cloud_xyz: a line of 256 characters or more (the reference's getline(line, 256) loop never ends on it)");
    }
    *n_points = n;
  }
  const int64_t n = *n_points;
  if (n > 0 && (rc = pcp_set_cloud_device(ctx, ctx->txt_out.p, n))) return rc;
  if (out_xyzi) {
    const size_t m = (size_t)(n < out_capacity ? n : out_capacity);
    if (m) PCP_CUDA(cudaMemcpyAsync(out_xyzi, ctx->txt_out.p, m * 4 * sizeof(float), out_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                                    ctx->stream));
  }
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_set_cloud_txt(pcp_ctx* ctx, const void* text, int64_t n_bytes, int text_is_device, int64_t* n_points, float* out_xyzi,
                              int64_t out_capacity, int out_is_device) {
  return set_cloud_text(ctx, text, n_bytes, text_is_device, n_points, out_xyzi, out_capacity, out_is_device, 0, "pcp_set_cloud_txt: bad argument");
}

PCP_API int pcp_set_cloud_xyz(pcp_ctx* ctx, const void* text, int64_t n_bytes, int text_is_device, int64_t* n_points, float* out_xyzi,
                              int64_t out_capacity, int out_is_device) {
  return set_cloud_text(ctx, text, n_bytes, text_is_device, n_points, out_xyzi, out_capacity, out_is_device, 1, "pcp_set_cloud_xyz: bad argument");
}

}  // extern "C"
