// grid.cu — device build of the cell-sorted uniform grid that replaces the FLANN
// KDTreeSingleIndex build inside pcl::KdTreeFLANN::setInputCloud (reference call site
// src/pcaAnalysis.cpp:14-15).  Stages: bbox reduce → cell keys → radix sort (key, local
// index) → gather to cell order + cell-end marks → max-scan into the lower-bound table.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cfloat>
#include <cmath>
#include <cstring>

#include <cstdio>

#include "pcp_internal.cuh"

namespace {

__device__ __forceinline__ uint32_t enc_ordered(float f) {
  uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
inline float dec_ordered(uint32_t u) {
  uint32_t b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
  float f;
  memcpy(&f, &b, 4);
  return f;
}

__device__ __forceinline__ bool finite3(float4 p) { return isfinite(p.x) && isfinite(p.y) && isfinite(p.z); }

// bbox[0..2] = min xyz, bbox[3..5] = max xyz (ordered-uint encoding); grid-stride, one
// float4 per thread per step (coalesced 16-B loads), warp-shuffle then one atomic per warp.
__global__ void __launch_bounds__(256) bbox_kernel(const float4* __restrict__ pts, int64_t n,
                                                   uint32_t* __restrict__ bbox) {
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float4 p = __ldg(pts + i);
    if (finite3(p)) {
      lo[0] = fminf(lo[0], p.x); hi[0] = fmaxf(hi[0], p.x);
      lo[1] = fminf(lo[1], p.y); hi[1] = fmaxf(hi[1], p.y);
      lo[2] = fminf(lo[2], p.z); hi[2] = fmaxf(hi[2], p.z);
    }
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[d] = fminf(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
      hi[d] = fmaxf(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
    }
  }
  // block-level combine, then six atomics per block (thousands of warps hammering six words was
  // most of this kernel's time)
  __shared__ float s_lo[8][3], s_hi[8][3];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      s_lo[warp][d] = lo[d];
      s_hi[warp][d] = hi[d];
    }
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    const int d = threadIdx.x;
    float l = s_lo[0][d], h = s_hi[0][d];
#pragma unroll
    for (int w = 1; w < 8; ++w) {
      l = fminf(l, s_lo[w][d]);
      h = fmaxf(h, s_hi[w][d]);
    }
    atomicMin(bbox + d, enc_ordered(l));
    atomicMax(bbox + 3 + d, enc_ordered(h));
  }
}

__global__ void bbox_init_kernel(uint32_t* bbox) {
  if (threadIdx.x < 3) bbox[threadIdx.x] = 0xFFFFFFFFu;
  else if (threadIdx.x < 8) bbox[threadIdx.x] = 0u;  // [6]: "the shard's own global ids are not strictly ascending"
}

// Shards, one pass over all n ids (own, halo, padding = INT_MAX): flag[0] |= 1 unless the own ids are strictly
// ascending (a slab cut out of a cloud in index order: then only the halo entries need sorting to rank every id);
// flag[0] |= 2 for a negative id; flag[1] = the largest real id (small ids serve as `ord` directly, no ranking at all).
__global__ void __launch_bounds__(256) gid_scan_kernel(const int32_t* __restrict__ gid, int64_t n, int64_t n_own, uint32_t* __restrict__ flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  uint32_t bad = 0, top = 0;
  if (i < n) {
    const int32_t g = __ldg(gid + i);
    if (g < 0) bad |= 2u;
    if (i < n_own && (g == 0x7fffffff || (i + 1 < n_own && __ldg(gid + i + 1) <= g))) bad |= 1u;
    if (g >= 0 && g != 0x7fffffff) top = (uint32_t)g;
  }
  bad = __reduce_or_sync(0xffffffffu, bad);
  top = __reduce_max_sync(0xffffffffu, top);
  if ((threadIdx.x & 31) == 0) {
    if (bad) atomicOr(flag, bad);
    if (top) atomicMax(flag + 1, top);
  }
}

// Cell coordinates in double: (p − origin) is exact for fp32 inputs, so a neighbour within
// (R − 1e-6) cells of a query in true distance is never more than R cells away in index.
__device__ __forceinline__ uint32_t cell_of(const GridDesc& g, float4 p) {
  const float c[3] = {p.x, p.y, p.z};
  uint32_t id = 0;
#pragma unroll
  for (int a = 2; a >= 0; --a) {
    int axis = g.ax[a];
    double t = ((double)c[axis] - g.origin[axis]) * g.inv_cell;
    int ci = (int)floor(t);
    ci = max(0, min(ci, g.dim[a] - 1));
    id = id * (uint32_t)g.dim[a] + (uint32_t)ci;
  }
  return id;
}

// frames != nullptr: frame f holds the points [frames[f], frames[f+1]); its cells follow frame f-1's.
__global__ void __launch_bounds__(256) keys_kernel(const float4* __restrict__ pts, int64_t n, GridDesc g,
                                                   const int64_t* __restrict__ frames, uint32_t* __restrict__ keys,
                                                   int32_t* __restrict__ vals) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 p = __ldg(pts + i);
  uint32_t key = g.n_cells;
  if (finite3(p)) {
    key = cell_of(g, p);
    if (frames) {
      int lo = 0, hi = g.n_seg;  // last f with frames[f] <= i
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(frames + mid) <= i) lo = mid; else hi = mid;
      }
      key += (uint32_t)lo * g.seg_cells;
    }
  }
  keys[i] = key;
  vals[i] = (int32_t)i;
}

// Gather into cell order; mark E[key+1] = pos+1 at the last point of every cell.
// .w of the sorted point carries `ord`, the point's rank in index order (SURVEY Q7 tie rule):
// the local index for a plain cloud, the rank of its global id for a shard.
__global__ void __launch_bounds__(256)
    gather_kernel(const float4* __restrict__ pts, const int32_t* __restrict__ ord_of_li, const uint32_t* __restrict__ keys_sorted,
                  const int32_t* __restrict__ vals_sorted, int64_t n, int64_t n_query, float4* __restrict__ sorted,
                  int32_t* __restrict__ slot, uint32_t* __restrict__ inv, uint32_t inv_size, uint32_t* __restrict__ cell_end) {
  int64_t pos = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (pos >= n) return;
  int32_t li = vals_sorted[pos];
  float4 p = __ldg(pts + li);
  int32_t ord = ord_of_li ? __ldg(ord_of_li + li) : li;
  p.w = __int_as_float(ord);
  sorted[pos] = p;
  if ((uint32_t)ord < inv_size) inv[ord] = (uint32_t)pos;  // (a shard's padding entries carry id INT_MAX when the ids serve as ord)
  slot[pos] = li < n_query ? li : -1;
  uint32_t k = keys_sorted[pos];
  if (pos == n - 1 || keys_sorted[pos + 1] != k) cell_end[k + 1] = (uint32_t)(pos + 1);
}

__global__ void __launch_bounds__(256) iota_kernel(int32_t* __restrict__ v, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) v[i] = (int32_t)i;
}
// Rank of every global id when the own ids ascend and the halo ids hs[0..m) are sorted: an own point's rank is its
// index plus the halo ids below it, a halo entry's rank its sorted position plus the own ids at or below it — what
// a stable sort of all n ids gives (own entries come first in local order), from a sort of the m halo entries only.
__global__ void __launch_bounds__(256) rank_own_kernel(const int32_t* __restrict__ gid, int64_t n_own, const uint32_t* __restrict__ hs, int m,
                                                       int32_t* __restrict__ ord_of_li, int32_t* __restrict__ gid_of_ord) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_own) return;
  const uint32_t g = (uint32_t)__ldg(gid + i);
  int lo = 0, hi = m;  // first halo id >= g
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(hs + mid) < g) lo = mid + 1; else hi = mid;
  }
  const int32_t ord = (int32_t)(i + lo);
  ord_of_li[i] = ord;
  gid_of_ord[ord] = (int32_t)g;
}
__global__ void __launch_bounds__(256) rank_halo_kernel(const int32_t* __restrict__ gid, int64_t n_own, const uint32_t* __restrict__ hs,
                                                        const int32_t* __restrict__ hv, int m, int32_t* __restrict__ ord_of_li,
                                                        int32_t* __restrict__ gid_of_ord) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= m) return;
  const uint32_t g = __ldg(hs + r);
  int64_t lo = 0, hi = n_own;  // first own id > g
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if ((uint32_t)__ldg(gid + mid) <= g) lo = mid + 1; else hi = mid;
  }
  const int32_t ord = (int32_t)(r + lo);
  ord_of_li[n_own + __ldg(hv + r)] = ord;
  gid_of_ord[ord] = (int32_t)g;
}
__global__ void __launch_bounds__(256) invert_kernel(const int32_t* __restrict__ li_of_ord, int32_t* __restrict__ ord_of_li, int64_t n) {
  int64_t o = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (o < n) ord_of_li[li_of_ord[o]] = (int32_t)o;
}

// Point-weighted histogram of the cell occupancy for the automatic cell size: hist[m] = number of points
// living in a cell of m points (m clamped to OCC_BINS - 1).  The host takes its median — the occupancy a
// typical point sees; the mean would be hijacked by a handful of cells holding thousands of returns.
constexpr int OCC_BINS = 1024;
__global__ void __launch_bounds__(256) occupancy_kernel(const uint32_t* __restrict__ E, uint32_t n_cells,
                                                        unsigned long long* __restrict__ hist) {
  __shared__ uint32_t sh[OCC_BINS];
  for (int i = threadIdx.x; i < OCC_BINS; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < n_cells; c += gridDim.x * blockDim.x) {
    const uint32_t m = __ldg(E + c + 1) - __ldg(E + c);
    if (m) atomicAdd(&sh[min(m, (uint32_t)OCC_BINS - 1)], m);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < OCC_BINS; i += blockDim.x)
    if (sh[i]) atomicAdd(hist + i, (unsigned long long)sh[i]);
}

struct MaxOp {
  __device__ __forceinline__ uint32_t operator()(uint32_t a, uint32_t b) const { return a > b ? a : b; }
};

int bits_for(uint32_t v) {
  int b = 1;
  while (b < 32 && (v >> b) != 0) ++b;
  return b;
}

}  // namespace

// Bounding box of the current cloud (+ the "own ids ascend" flag of a shard) into ctx->bbox, at the device's highest
// launch priority: the two kernels are tiny and a host thread will wait for their result.
int pcp_bbox_enqueue(pcp_ctx* ctx) {
  const int64_t n = ctx->n;
  cudaStream_t st = ctx->stream;
  int rc;
  if (n <= 0) return PCP_OK;
  if ((rc = pcp_reserve(ctx, ctx->bbox, 8 * sizeof(uint32_t)))) return rc;
  pcp_tic(ctx, PCP_T_BBOX);
  bbox_init_kernel<<<1, 32, 0, st>>>((uint32_t*)ctx->bbox.p);
  PCP_LAUNCH_CHECK("bbox_init_kernel");
  {
    const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 4);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributePriority;
    at[0].val.priority = ctx->prio_high;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    PCP_CUDA(cudaLaunchKernelEx(&cfg, bbox_kernel, ctx->d_cloud, n, (uint32_t*)ctx->bbox.p));
    PCP_LAUNCH_CHECK("bbox_kernel");
  }
  pcp_toc(ctx, PCP_T_BBOX);
  if (ctx->d_gid) {  // rides on the bounding box's read-back: no host wait of its own
    gid_scan_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(ctx->d_gid, n, ctx->n_query, (uint32_t*)ctx->bbox.p + 6);
    PCP_LAUNCH_CHECK("gid_scan_kernel");
  }
  ctx->bbox_enqueued = true;
  return PCP_OK;
}

int pcp_grid_build_impl(pcp_ctx* ctx, float cell, int k_hint) {
  const int64_t n = ctx->n;
  const int n_seg = (int)ctx->n_frames;
  ctx->grid_ok = false;
  if (n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_build_grid: empty cloud");
  if (n >= (int64_t)0x7FFFFF00) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_build_grid: n exceeds int32 indexing");
  cudaStream_t st = ctx->stream;
  int rc;

  // ---- bbox (already in the stream when the upload enqueued it)
  if (!ctx->bbox_enqueued && (rc = pcp_bbox_enqueue(ctx))) return rc;
  ctx->bbox_enqueued = false;  // consumed: a second build of the same cloud (automatic cell) reads the same values, a new cloud resets it
  uint32_t hb[8];
  PCP_CUDA(cudaMemcpyAsync(hb, ctx->bbox.p, sizeof(hb), cudaMemcpyDeviceToHost, st));
  PCP_CUDA(cudaStreamSynchronize(st));
  double lo[3], hi[3], ext[3];
  for (int d = 0; d < 3; ++d) {
    lo[d] = dec_ordered(hb[d]);
    hi[d] = dec_ordered(hb[3 + d]);
    if (!(hi[d] >= lo[d])) { lo[d] = 0; hi[d] = 0; }  // no finite point at all
    ext[d] = hi[d] - lo[d];
  }

  // ---- grid geometry
  GridDesc g;
  // axis order: ax[2] (slowest) = longest extent, ax[0] (fastest, the run axis) = next, ax[1] = shortest
  int order[3] = {0, 1, 2};
  for (int i = 0; i < 3; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (ext[order[j]] > ext[order[i]]) std::swap(order[i], order[j]);
  g.ax[2] = order[0];
  g.ax[0] = order[1];
  g.ax[1] = order[2];
  bool auto_cell = !(cell > 0.f);
  const int kk = k_hint > 0 ? k_hint : 20;
  const double occ_target = 0.9 * (double)kk;
  // median of the point-weighted occupancy histogram in ctx->h_occ; 0 when it is empty
  auto occ_median = [&]() -> int {
    unsigned long long total = 0, cum = 0;
    for (int i = 1; i < OCC_BINS; ++i) total += ctx->h_occ[i];
    if (total == 0) return 0;
    int med = 1;
    for (; med < OCC_BINS - 1; ++med) {
      cum += ctx->h_occ[med];
      if (2 * cum >= total) break;
    }
    return med;
  };
  bool learned = false;  // a similar cloud was built before on this context: its cell correction is reused
  float cell0 = 1.f;     // the first guess
  if (auto_cell) {
    // surface-like density estimate: n points over the two largest extents; aim at about
    // 1.9*k points in an average cell (measured optimum of the tile kernel on the urban cloud: fewer
    // uncovered queries beat the larger candidate set) so a 3×3×3 stencil usually covers the k-NN ball.
    double area = std::max(ext[order[0]], 1e-3) * std::max(ext[order[1]], 1e-3);
    cell0 = (float)std::sqrt(1.9 * kk * area / ((double)n / (double)n_seg));  // per frame: the frames share the box
    // k close to the 32-entry lane capacity: the threshold bin overflows the lane's column more and more often as
    // cells grow (measured cliff, 1 M-point urban cloud: k = 32 2.94 ms at this guess vs 2.30 ms at 0.85 x, k = 30
    // 2.31 vs ~2.1 ms; k <= 25 is best at the plain guess), so the guess is taken 15 % smaller there
    if (kk > 26) cell0 *= 0.85f;
    if (!(cell0 > 0.f)) cell0 = 1.f;
    learned = ctx->scale_n > 0 && ctx->scale_seg == n_seg && ctx->scale_k == kk && (double)n >= 0.8 * (double)ctx->scale_n &&
              (double)n <= 1.25 * (double)ctx->scale_n;
    if (!learned) {
      ctx->cell_scale = 1.0;
      ctx->occ_pending = false;
    } else if (ctx->occ_pending && cudaEventQuery(ctx->occ_ev) == cudaSuccess) {
      // the occupancy of an earlier build of this stream of clouds has arrived: steer the correction
      ctx->occ_pending = false;
      const int med = occ_median();
      if (med > 0 && ((double)med > 1.6 * occ_target || (double)med < 0.5 * occ_target)) ctx->cell_scale *= std::sqrt(occ_target / (double)med);
    }
    cell = (float)((double)cell0 * ctx->cell_scale);
    if (!(cell > 0.f)) cell = cell0;
    if (!ctx->h_occ) {
      PCP_CUDA(cudaHostAlloc((void**)&ctx->h_occ, OCC_BINS * sizeof(unsigned long long), cudaHostAllocDefault));
      PCP_CUDA(cudaEventCreateWithFlags(&ctx->occ_ev, cudaEventDisableTiming));
    }
  }
  // ---- sharded clouds: ord = rank of the global id (ties are broken by global index) — or the id itself when a dense
  // ord -> position table over the ids is affordable (1 GB = 268 M ids; 64 MB stay L2-resident, a larger table costs the
  // work-list kernels a DRAM sector per neighbour of a deferred query, far less than ranking 10^7 ids per build: the tile
  // kernels never read it): nothing to rank, nothing to sort
  ctx->gid_direct = false;
  ctx->inv_size = (size_t)n;
  if (ctx->d_gid && (hb[6] & 2u)) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_build_grid: negative global id");
  if (ctx->d_gid && ctx->knobs.shard_full_sort == 0 && ((size_t)hb[7] + 1) * sizeof(uint32_t) <= ((size_t)1 << 30)) {
    ctx->gid_direct = true;
    ctx->inv_size = (size_t)hb[7] + 1;
  } else if (ctx->d_gid) {
    if ((rc = pcp_reserve(ctx, ctx->gid_sorted, n * sizeof(int32_t)))) return rc;
    if ((rc = pcp_reserve(ctx, ctx->gid_tmp, n * sizeof(int32_t)))) return rc;
    if ((rc = pcp_reserve(ctx, ctx->ord_of_li, n * sizeof(int32_t)))) return rc;
    const int nbi = (int)((n + 255) / 256);
    const int64_t n_own = ctx->n_query;
    const int64_t m = n - n_own;  // halo entries (padding included: ids 0x7fffffff, ranked last in arrival order)
    if (n_own > 0 && (hb[6] & 1u) == 0u && ctx->knobs.shard_full_sort != 1) {
      // the own ids ascend (a slab cut from a cloud in index order): sort the halo entries only and rank by two
      // binary searches — the full-width sort of all n ids was the largest fixed cost of a sharded build
      int32_t* hv_in = (int32_t*)ctx->gid_tmp.p;    // 0 .. m-1
      if ((rc = pcp_reserve(ctx, ctx->li_of_ord, (size_t)std::max<int64_t>(n, 2 * m) * sizeof(int32_t)))) return rc;
      uint32_t* hs = (uint32_t*)ctx->li_of_ord.p;   // the halo ids, sorted
      int32_t* hv = (int32_t*)ctx->li_of_ord.p + m; // their halo-local indices
      if (m > 0) {
        iota_kernel<<<(int)((m + 255) / 256), 256, 0, st>>>(hv_in, m);
        PCP_LAUNCH_CHECK("iota_kernel");
        size_t tb = 0;
        PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, (const uint32_t*)(ctx->d_gid + n_own), hs, (const int32_t*)hv_in, hv, (int)m, 0, 32, st));
        if ((rc = pcp_reserve(ctx, ctx->sort_tmp, tb))) return rc;
        PCP_CUDA(cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tb, (const uint32_t*)(ctx->d_gid + n_own), hs, (const int32_t*)hv_in, hv, (int)m, 0, 32, st));
        ctx->launches += 5;
      }
      rank_own_kernel<<<(int)((n_own + 255) / 256), 256, 0, st>>>(ctx->d_gid, n_own, hs, (int)m, (int32_t*)ctx->ord_of_li.p, (int32_t*)ctx->gid_sorted.p);
      PCP_LAUNCH_CHECK("rank_own_kernel");
      if (m > 0) {
        rank_halo_kernel<<<(int)((m + 255) / 256), 256, 0, st>>>(ctx->d_gid, n_own, hs, hv, (int)m, (int32_t*)ctx->ord_of_li.p, (int32_t*)ctx->gid_sorted.p);
        PCP_LAUNCH_CHECK("rank_halo_kernel");
      }
    } else {
      if ((rc = pcp_reserve(ctx, ctx->li_of_ord, n * sizeof(int32_t)))) return rc;
      iota_kernel<<<nbi, 256, 0, st>>>((int32_t*)ctx->gid_tmp.p, n);
      PCP_LAUNCH_CHECK("iota_kernel");
      size_t tb = 0;
      PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, (const uint32_t*)ctx->d_gid, (uint32_t*)ctx->gid_sorted.p,
                                               (const int32_t*)ctx->gid_tmp.p, (int32_t*)ctx->li_of_ord.p, (int)n, 0, 32, st));
      if ((rc = pcp_reserve(ctx, ctx->sort_tmp, tb))) return rc;
      PCP_CUDA(cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tb, (const uint32_t*)ctx->d_gid, (uint32_t*)ctx->gid_sorted.p,
                                               (const int32_t*)ctx->gid_tmp.p, (int32_t*)ctx->li_of_ord.p, (int)n, 0, 32, st));
      ctx->launches += 5;
      invert_kernel<<<nbi, 256, 0, st>>>((const int32_t*)ctx->li_of_ord.p, (int32_t*)ctx->ord_of_li.p, n);
      PCP_LAUNCH_CHECK("invert_kernel");
    }
  }

  // diagnostic knob: lower the grid origin by a fraction of a cell (all axes, or one); results are
  // unchanged, only where the cell faces fall relative to the cloud's layers
  const double origin_shift = ctx->knobs.grid_shift;
  const int shift_axis = ctx->knobs.grid_shift_axis;
  for (int attempt = 0;; ++attempt) {
  const double max_cells = 1.0e9;
  for (;;) {
    double total = 1.0;
    for (int a = 0; a < 3; ++a) {
      double cnt = std::floor(ext[g.ax[a]] / (double)cell + (shift_axis < 0 || shift_axis == g.ax[a] ? origin_shift : 0.0)) + 1.0;
      if (cnt > 2.0e9) cnt = 2.0e9;
      g.dim[a] = (int)cnt;
      total *= cnt;
    }
    if (total * n_seg <= max_cells) {
      g.seg_cells = (uint32_t)total;
      g.n_cells = g.seg_cells * (uint32_t)n_seg;
      g.n_seg = n_seg;
      break;
    }
    if (!auto_cell) return pcp_fail(ctx, PCP_ERR_GRID_TOO_LARGE, "pcp_build_grid: dense cell table above 1e9 cells; use a larger cell");
    cell *= 1.26f;
  }
  g.cell = cell;
  g.inv_cell = 1.0 / (double)cell;
  for (int d = 0; d < 3; ++d) g.origin[d] = lo[d] - (shift_axis < 0 || shift_axis == d ? origin_shift : 0.0) * (double)cell;
  ctx->grid = g;

  // ---- keys
  if ((rc = pcp_reserve(ctx, ctx->keys_in, n * sizeof(uint32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->keys_sorted, (n + 1) * sizeof(uint32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->vals_in, n * sizeof(int32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->vals_sorted, n * sizeof(int32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->sorted_pts, n * sizeof(float4)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->sorted_slot, n * sizeof(int32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->inv_pos, ctx->inv_size * sizeof(uint32_t)))) return rc;
  const size_t table_entries = (size_t)g.n_cells + 2;
  if ((rc = pcp_reserve(ctx, ctx->cell_end, table_entries * sizeof(uint32_t)))) return rc;
  const int nb = (int)((n + 255) / 256);
  pcp_tic(ctx, PCP_T_KEYS);
  keys_kernel<<<nb, 256, 0, st>>>(ctx->d_cloud, n, g, n_seg > 1 ? (const int64_t*)ctx->frames.p : nullptr, (uint32_t*)ctx->keys_in.p,
                                  (int32_t*)ctx->vals_in.p);
  PCP_LAUNCH_CHECK("keys_kernel");
  pcp_toc(ctx, PCP_T_KEYS);

  // ---- radix sort of (cell id, local index); stable, so equal cells keep index order
  const int end_bit = bits_for(g.n_cells);
  size_t tmp_bytes = 0;
  PCP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint32_t*)ctx->keys_in.p, (uint32_t*)ctx->keys_sorted.p,
                                           (const int32_t*)ctx->vals_in.p, (int32_t*)ctx->vals_sorted.p, (int)n, 0, end_bit, st));
  if ((rc = pcp_reserve(ctx, ctx->sort_tmp, tmp_bytes))) return rc;
  pcp_tic(ctx, PCP_T_SORT);
  PCP_CUDA(cub::DeviceRadixSort::SortPairs(ctx->sort_tmp.p, tmp_bytes, (const uint32_t*)ctx->keys_in.p, (uint32_t*)ctx->keys_sorted.p,
                                           (const int32_t*)ctx->vals_in.p, (int32_t*)ctx->vals_sorted.p, (int)n, 0, end_bit, st));
  ctx->launches += 4;  // onesweep: histogram + a few digit passes (CUB-internal kernels)
  pcp_toc(ctx, PCP_T_SORT);

  // ---- gather + cell table
  pcp_tic(ctx, PCP_T_TABLE);
  PCP_CUDA(cudaMemsetAsync(ctx->cell_end.p, 0, table_entries * sizeof(uint32_t), st));
  gather_kernel<<<nb, 256, 0, st>>>(ctx->d_cloud, ctx->d_gid ? (ctx->gid_direct ? ctx->d_gid : (const int32_t*)ctx->ord_of_li.p) : nullptr,
                                    (const uint32_t*)ctx->keys_sorted.p, (const int32_t*)ctx->vals_sorted.p, n, ctx->n_query,
                                    (float4*)ctx->sorted_pts.p, (int32_t*)ctx->sorted_slot.p, (uint32_t*)ctx->inv_pos.p,
                                    (uint32_t)std::min<size_t>(ctx->inv_size, 0xFFFFFFFFu), (uint32_t*)ctx->cell_end.p);
  PCP_LAUNCH_CHECK("gather_kernel");
  size_t scan_bytes = 0;
  uint32_t* E = (uint32_t*)ctx->cell_end.p;
  PCP_CUDA(cub::DeviceScan::InclusiveScan(nullptr, scan_bytes, E, E, MaxOp(), (int)table_entries, st));
  if ((rc = pcp_reserve(ctx, ctx->scan_tmp, scan_bytes))) return rc;
  PCP_CUDA(cub::DeviceScan::InclusiveScan(ctx->scan_tmp.p, scan_bytes, E, E, MaxOp(), (int)table_entries, st));
  ctx->launches += 2;
  pcp_toc(ctx, PCP_T_TABLE);
  // ---- automatic cell: check the occupancy a typical point sees (point-weighted median of the cell
  // counts) against ~0.9 k, the measured optimum of the tile kernel on both the urban cloud and VLP-16
  // frames.  The first guess assumes n points spread evenly over the two largest extents; a cloud that is
  // dense in places and empty elsewhere (a sensor-centred scan) lands far above the target and is rebuilt
  // ONCE with cell * sqrt(target / occupancy) (surface-like scaling; a build costs ~2 % of the search).
  // The correction is remembered: the next cloud of similar size on this context starts from it, and its
  // own occupancy comes back asynchronously (no host wait), steering the clouds after it.
  if (!auto_cell) break;
  {
    if ((rc = pcp_reserve(ctx, ctx->occ_hist, OCC_BINS * sizeof(unsigned long long)))) return rc;
    unsigned long long* d_occ = (unsigned long long*)ctx->occ_hist.p;
    if (learned && ctx->occ_pending) break;  // an earlier measurement is still in flight
    if (learned && (ctx->occ_tick++ & 15) != 0) break;  // steady state: look again every 16th cloud only
    PCP_CUDA(cudaMemsetAsync(d_occ, 0, OCC_BINS * sizeof(unsigned long long), st));
    const int ob = (int)std::min<uint64_t>(((uint64_t)g.n_cells + 4095) / 4096, 148 * 2);  // few blocks: each flushes 1024 bins
    occupancy_kernel<<<ob, 256, 0, st>>>((const uint32_t*)ctx->cell_end.p, g.n_cells, d_occ);
    PCP_LAUNCH_CHECK("occupancy_kernel");
    PCP_CUDA(cudaMemcpyAsync(ctx->h_occ, d_occ, OCC_BINS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    if (learned) {
      PCP_CUDA(cudaEventRecord(ctx->occ_ev, st));
      ctx->occ_pending = true;
      break;
    }
    PCP_CUDA(cudaStreamSynchronize(st));
    const int med = occ_median();
    const bool in_band = med == 0 || ((double)med <= 1.6 * occ_target && (double)med >= 0.5 * occ_target);
    if (in_band || attempt > 0) {
      ctx->scale_n = n;
      ctx->scale_seg = n_seg;
      ctx->scale_k = kk;
      ctx->cell_scale = (double)g.cell / (double)cell0;
      break;
    }
    cell = (float)((double)cell * std::sqrt(occ_target / (double)med));
    if (!(cell > 0.f)) break;
  }
  }
  ctx->grid_ok = true;
  return PCP_OK;
}
