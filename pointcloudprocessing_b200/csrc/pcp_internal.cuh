// pcp_internal.cuh — context object and shared helpers of libpcp.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "pcp.h"

#define PCP_WARP 32

// Uniform grid over the cloud's bounding box.  Cell (c0,c1,c2) are coordinates along the
// permuted axes ax[0] (fastest in the linear cell index, so the 2R+1 cells of a stencil
// row are one contiguous run of the sorted point array), ax[1], ax[2] (slowest = the
// longest box axis, so three consecutive slabs stay L2-resident while queries sweep).
struct GridDesc {
  double origin[3];   // box minimum along x,y,z
  double inv_cell;    // 1 / cell edge
  float cell;         // cell edge
  int ax[3];          // permuted axis -> x/y/z
  int dim[3];         // cells along permuted axes
  uint32_t n_cells;   // n_seg*seg_cells; cell id n_cells is the trash cell (non-finite points)
  // Segmented grid (frame batches, pcp_set_frames): every frame has its own copy of the dim0×dim1×dim2
  // box, frame f's cells are ids [f*seg_cells, (f+1)*seg_cells) — the frame is folded into the slowest
  // cell coordinate (row z' = f*dim2 + c2), so a stencil clamped to [0, dim2) never leaves its frame.
  uint32_t seg_cells; // dim0*dim1*dim2
  int n_seg;          // 1 for a plain cloud
};

// a query's folded slowest coordinate -> (frame row offset, coordinate inside the frame)
__host__ __device__ __forceinline__ int pcp_unfold_z(const GridDesc& g, int& c2) {
  if (g.n_seg == 1) return 0;
  const int f = c2 / g.dim[2];
  c2 -= f * g.dim[2];
  return f * g.dim[2];
}

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  bool listed = false;  // registered in pcp_ctx::owned (pcp_destroy frees every buffer that was ever reserved)
};

// Environment knobs (A/B timing and tests), read ONCE in pcp_create — never on the launch path.
struct PcpKnobs {
  int knn_impl = 0;        // PCP_KNN_IMPL: 0 = automatic (knn_tile6.cu for single clouds, knn_tile.cu for frame batches) + work list, 6 = "tile6" / 4 = "tile4" force one: the round-1
                           // tile kernel (knn_tile.cu), 3 = "v3": the warp-per-query kernel for every point
  int knn_groups = 0;      // PCP_KNN_GROUPS: stencil rows served per tile (0 = automatic)
  int radius_impl = 0;     // PCP_RADIUS_IMPL: 0 = automatic, 1 = "tile" (32-entry lanes), 2 = "tile64", 3 = "v3"
  int radius_t6 = -1;      // PCP_RADIUS_T6=1/0: fixed-radius tiles through knn_tile6.cu (T6_BIN_RADIUS) / through the round-1 kernel (-1 = automatic)
  int wl_blocks = 0;       // PCP_WL_BLOCKS / PCP_WL_TILE: work-list kernel grid and tile (0 = default)
  int wl_tile = 0;
  double grid_shift = 0.0; // PCP_GRID_SHIFT: lower the grid origin by this fraction of a cell (diagnostic; results unchanged)
  int grid_shift_axis = -1;// PCP_GRID_SHIFT_AXIS: 0 / 1 / 2 = x / y / z only
  int t6_bin = -1;         // PCP_T6_BIN=sqrt|log: bin mode of the k-NN tile kernel's selection passes (-1 = the launcher's choice)
  int t6_refine_min = 0;   // PCP_T6_REFINE_MIN: lanes of a tile that must want a refinement pass for it to run (0 = default)
  int t6_no_order = 0;     // PCP_T6_ORDER=0: warps take tiles in cell order (A/B of tile_order_kernel)
  int flat_priority = 0;   // PCP_PRIO=flat: the search kernels run at the stream's priority (A/B of the two-level scheme below)
  int shard_full_sort = 0; // PCP_SHARD_SORT (A/B, tests): "rank" = rank a shard's global ids even when they are small enough to serve as ord (2),
                           // "full" = rank them by sorting all of them even when the own ids ascend (1)
};

struct pcp_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  // Launch priority of the long search kernels (a per-launch attribute; = the stream's own priority by default).
  // Measured with three contexts in flight (PCP_PRIO=two: stream at the device's highest priority, search kernels at
  // the lowest, so that one context's short grid kernels overtake another's search): 434 M points/s end to end against
  // 587 M with everything at one priority — contexts that share the SMs evenly finish together and their result copies
  // no longer overlap the next context's compute.  Kept as a knob; only the tiny bounding-box kernels run at high priority.
  int prio_search = 0;
  int prio_high = 0;
  bool host_async = false;  // pcp_set_host_async: host-pointer calls only enqueue
  int64_t launches = 0;
  PcpKnobs knobs;
  std::vector<DevBuf*> owned;  // every DevBuf pcp_reserve ever allocated for this context
  float t_ms[PCP_T_COUNT] = {0};
  cudaEvent_t ev[2 * PCP_T_COUNT] = {nullptr};
  bool ev_used[PCP_T_COUNT] = {false};

  // cloud (packed float4 x,y,z,intensity on the device)
  const float4* d_cloud = nullptr;
  const int32_t* d_gid = nullptr;  // optional global ids (sharded)
  int64_t n = 0, n_query = 0;
  DevBuf own_cloud;  // when the cloud came from the host
  int64_t n_frames = 1;            // pcp_set_frames: the cloud is n_frames independent clouds laid end to end
  std::vector<int64_t> h_frames;   // host copy of the n_frames + 1 offsets
  DevBuf frames;                   // device copy (int64)

  // grid
  bool grid_ok = false;
  GridDesc grid;
  DevBuf bbox;        // 6 ordered-uint encodings (+ the shard id flag)
  bool bbox_enqueued = false;  // the bounding-box kernels of the CURRENT cloud are already in the stream (pcp_set_cloud_host in
                               // pcp_set_host_async mode enqueues them behind the upload, so that a caller who uploads the next
                               // cloud early finds the box ready when it builds that cloud's grid); reset by every pcp_set_cloud_*
  DevBuf occ_hist;    // automatic cell: point-weighted histogram of the cell occupancy
  unsigned long long* h_occ = nullptr;  // its page-locked host copy
  cudaEvent_t occ_ev = nullptr;         // recorded behind an asynchronous copy into h_occ
  bool occ_pending = false;
  unsigned occ_tick = 0;
  double cell_scale = 1.0;              // learned correction of the first-guess cell for clouds like the last one
  int64_t scale_n = 0;
  int scale_seg = 0, scale_k = 0;
  DevBuf keys_in, keys_sorted, vals_in, vals_sorted, sort_tmp;
  DevBuf sorted_pts;  // float4 {x,y,z, ord bits} in cell order (ord = rank in index order)
  DevBuf inv_pos;     // uint32 ord -> sorted position (inv_size entries: n, or the largest id + 1 of a shard whose ids serve as ord)
  size_t inv_size = 0;
  bool gid_direct = false;  // shard: ord IS the global id (no ranking; ord -> id is the identity)
  DevBuf gid_sorted, gid_tmp, ord_of_li, li_of_ord;  // sharded clouds: ord <-> local index <-> global id
  DevBuf sorted_slot; // int32 output slot (−1 for halo points) per sorted position
  DevBuf cell_end;    // uint32 E[c] = #points with cell id < c, n_cells + 2 entries
  DevBuf scan_tmp;

  // outputs / scratch
  DevBuf tile_perm;   // knn_tile6.cu: tile served by every warp (tiles of similar work share a CTA)
  DevBuf work_list, work_count;  // k-NN queries deferred by the tile kernel (sorted positions)
  DevBuf out_feat, out_norm, out_idx, out_d2, out_counts, flags, valid, last_valid, staging;
  DevBuf ring_range, ring_idx, ring_long, ring_offsets, ring_key;
  DevBuf obb;         // pcp_bbox_pca: sums, extents and the result
  DevBuf heavy_keys, heavy_keys2, heavy_seg;  // radius_heavy.cu: keys of the heavy lists (in / sorted), segment bounds
  DevBuf halo_done;   // block counter of halo_pad_signal_kernel
  DevBuf dec_tables, dec_packed, dec_out;  // pcp_decode_packets: azimuth cos/sin tables, per-block scan, staged outputs
  DevBuf odom_counts, odom_range, odom_ptid, odom_index, odom_stats, odom_minmax, odom_smooth, odom_type;  // lidarOdometry images
  DevBuf db_parent, db_root, db_flag, db_rank, db_boxes, db_labels, db_out, ms_a, ms_b;  // DBSCAN / mean shift scratch
  DevBuf txt_in, txt_rank, txt_starts, txt_vals, txt_last, txt_out;  // pcp_set_cloud_txt
  const float* odom_last_range = nullptr;  // device images of the last pcp_odom_range_image
  int64_t odom_last_frames = 0;
  DevBuf seg_ground, seg_label, seg_colmax, seg_above, seg_counts, seg_nlabels;  // pcp_ring_segment outputs
};

int pcp_fail(pcp_ctx* ctx, int status, const char* what, cudaError_t ce = cudaSuccess);
int pcp_reserve(pcp_ctx* ctx, DevBuf& b, size_t bytes);
void pcp_tic(pcp_ctx* ctx, int stage);
void pcp_toc(pcp_ctx* ctx, int stage);
void pcp_collect_timings(pcp_ctx* ctx);

#define PCP_CUDA(call)                                               \
  do {                                                               \
    cudaError_t e__ = (call);                                        \
    if (e__ != cudaSuccess) return pcp_fail(ctx, PCP_ERR_CUDA, #call, e__); \
  } while (0)

#define PCP_LAUNCH_CHECK(name)                                       \
  do {                                                               \
    ctx->launches++;                                                 \
    cudaError_t e__ = cudaGetLastError();                            \
    if (e__ != cudaSuccess) return pcp_fail(ctx, PCP_ERR_CUDA, name, e__); \
  } while (0)

// launch of a long-running search kernel on the context's stream at the low priority (see pcp_ctx::prio_search)
template <typename... Params, typename... Args>
inline cudaError_t pcp_launch_search(pcp_ctx* ctx, void (*kernel)(Params...), dim3 grid, dim3 block, size_t smem, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = ctx->stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributePriority;
  at[0].val.priority = ctx->prio_search;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(args)...);
}

// implemented in grid.cu / search_pca.cu / ring.cu
int pcp_grid_build_impl(pcp_ctx* ctx, float cell, int k_hint);
int pcp_bbox_enqueue(pcp_ctx* ctx);  // grid.cu: bounding-box reduction of the current cloud into ctx->bbox (no host wait)

struct SearchParams {
  int mode;           // 0 = k-NN, 1 = radius
  int k;              // k-NN
  float r2;           // radius: (float)((double)r*(double)r)
  int R0;             // initial stencil half-width in cells
  int min_pts;        // radius: gate (13 in the reference)
  int center_mode;    // PCP_CENTER_*
  pcp_feature68* out; // device, n_query records (may be null)
  pcp_normal16* normals; // device (may be null)
  int32_t* nn_idx;    // device neighbour lists (may be null)
  float* nn_d2;
  int32_t* counts;    // radius: per-query neighbour count (may be null)
  const int64_t* offsets; // radius CSR offsets for nn_idx / nn_d2
  float* kth_d2;      // k-NN: squared distance of the k-th neighbour per output slot (may be null)
  uint8_t* valid;     // per output slot: 1 when the record was computed (may be null)
  int* overflow_flag; // device int: set when a radius neighbourhood exceeded the warp buffer
};
int pcp_search_launch(pcp_ctx* ctx, const SearchParams& sp);  // radius (and the v3 warp-per-query k-NN, A/B only)
struct KArgs;
void pcp_make_kargs(pcp_ctx* ctx, const SearchParams& sp, KArgs& a);  // search_pca.cu
// radius_heavy.cu: lists of neighbourhoods beyond the 512-entry on-chip buffer (count-then-fill through global memory)
int pcp_radius_heavy_lists_impl(pcp_ctx* ctx, const KArgs& a, float r2, int R, const int64_t* d_offsets, int64_t total, int32_t* d_idx,
                                float* d_d2);
int pcp_knn_tile_launch(pcp_ctx* ctx, const SearchParams& sp, const KArgs& a, int radius_cap);  // k-NN tile kernel (knn_tile.cu)
int pcp_knn_tile6_launch(pcp_ctx* ctx, const SearchParams& sp, const KArgs& a, int radius_cap = 32);  // k-NN tile kernel (knn_tile6.cu): k <= PCP_T6_KMAX
#define PCP_T6_K32 24   // 32-member lanes up to this k, 64-member lanes above
#define PCP_T6_KMAX 32   // (the work-list kernel that finishes the deferred lanes keeps one key per lane)
int pcp_carry_forward_impl(pcp_ctx* ctx, pcp_feature68* d_out, const uint8_t* d_valid, int64_t n);
