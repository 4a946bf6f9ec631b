// consumers.cu — the two radius-search consumers next to the PCA stage.
//
// DBSCAN — replaces DynaTools::clusterDBSCAN (include/dbscanCLuster.h:256-300).  The reference is a
// sequential seed loop whose expansion list grows while it is walked; its result has a closed form that
// does not depend on the walk (checked against the literal restatement on the CPU side):
//   * core(i)  = |{j : d2(i,j) < (float)(r*r)}| >= minpts (the point itself counts);
//   * clusters = connected components of the core points (edges: core pairs within r); the seed of a
//     cluster is its smallest core index, and ids run 1.. in the order of the seeds;
//   * a non-core point b joins, among the clusters owning a core point within r of it, the one with the
//     smallest seed — if that seed is < b; otherwise the seed loop has already visited b and called it
//     noise (state -2), which later clusters do not undo (:294 tests `== -1` only).
// So: capped neighbour count -> lock-free union-find that always hooks the larger root under the smaller
// (root = smallest index = seed) -> scan over the seed flags -> one labelling pass.  All integer work:
// bit-exact.
//
// Mean shift — replaces the sweep of buildingFacadeExtractor::meanShiftClustering
// (src/buildingFacadeExtraction.cpp:72-123).  The kd-tree keeps the ORIGINAL positions (FLANN copies
// them at setInputCloud) while the query and the differences use the CURRENT ones, so the search runs
// around arbitrary positions in the grid of the original cloud; one warp per point, the shift is the
// double-precision mean of the fp32 differences (the reference adds fp32 `diff / n` terms in (d2, index)
// order: equal to ~1e-7 relative, compared with a tolerance).
#include <cfloat>
#include <cmath>

#include <cub/device/device_scan.cuh>

#include "search_common.cuh"

namespace {

struct Stencil {
  int lo0, hi0, lo1, hi1, lo2, hi2, zoff;
};

__device__ __forceinline__ Stencil stencil_of_cell(const GridDesc& g, uint32_t cid, int R) {
  const int c0 = (int)(cid % (uint32_t)g.dim[0]);
  const uint32_t t = cid / (uint32_t)g.dim[0];
  const int c1 = (int)(t % (uint32_t)g.dim[1]);
  int c2 = (int)(t / (uint32_t)g.dim[1]);
  Stencil s;
  s.zoff = pcp_unfold_z(g, c2);
  s.lo0 = max(c0 - R, 0), s.hi0 = min(c0 + R, g.dim[0] - 1);
  s.lo1 = max(c1 - R, 0), s.hi1 = min(c1 + R, g.dim[1] - 1);
  s.lo2 = max(c2 - R, 0), s.hi2 = min(c2 + R, g.dim[2] - 1);
  return s;
}

// ---- union-find over point indices; roots are the smallest index of their set ----
__device__ __forceinline__ int uf_find(int* parent, int x) {
  int p = parent[x];
  while (p != x) {  // path halving (benign races: parents only ever decrease along a path)
    const int gp = parent[p];
    if (gp != p) parent[x] = gp;
    x = p;
    p = gp;
  }
  return x;
}
__device__ __forceinline__ void uf_union(int* parent, int a, int b) {
  int ra = uf_find(parent, a), rb = uf_find(parent, b);
  while (ra != rb) {
    if (ra < rb) {
      const int t = ra;
      ra = rb;
      rb = t;
    }  // ra > rb: hook ra under rb
    const int old = atomicCAS(parent + ra, ra, rb);
    if (old == ra) return;
    ra = uf_find(parent, old);
    rb = uf_find(parent, rb);
  }
}

__global__ void __launch_bounds__(256) dbscan_init_kernel(int* __restrict__ parent, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) parent[i] = (int)i;
}

// one thread per cell-sorted point: a core point unions with every core neighbour of smaller index
__global__ void __launch_bounds__(128) dbscan_union_kernel(KArgs a, int R, float r2, int min_pts, const int32_t* __restrict__ counts,
                                                           int* parent) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= a.n) return;
  const uint32_t cid = __ldg(&a.keys[p]);
  if (cid >= a.g.n_cells) return;
  const float4 Q = __ldg(&a.sorted[p]);
  const int i = __float_as_int(Q.w);
  if (counts[i] < min_pts) return;
  const Stencil s = stencil_of_cell(a.g, cid, R);
#pragma unroll 1
  for (int e2 = s.lo2; e2 <= s.hi2; ++e2)
#pragma unroll 1
    for (int e1 = s.lo1; e1 <= s.hi1; ++e1) {
      const uint32_t base = ((uint32_t)(s.zoff + e2) * (uint32_t)a.g.dim[1] + (uint32_t)e1) * (uint32_t)a.g.dim[0];
      const uint32_t beg = __ldg(&a.E[base + s.lo0]), end = __ldg(&a.E[base + s.hi0 + 1]);
#pragma unroll 1
      for (uint32_t j = beg; j < end; ++j) {
        const float4 P = __ldg(&a.sorted[j]);
        const int jn = __float_as_int(P.w);
        if (jn < i && dist2_nofma(Q.x, Q.y, Q.z, P) < r2 && counts[jn] >= min_pts) uf_union(parent, i, jn);
      }
    }
}

// root[i] = seed of i's cluster for core points (-1 otherwise); seed_flag[i] = 1 at the seeds
__global__ void __launch_bounds__(256) dbscan_flatten_kernel(int* __restrict__ parent, const int32_t* __restrict__ counts, int min_pts,
                                                             int64_t n, int* __restrict__ root, int* __restrict__ seed_flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int r = -1;
  if (counts[i] >= min_pts) {
    r = (int)i;
    for (int p = parent[r]; p != r; p = parent[r]) r = p;
  }
  root[i] = r;
  seed_flag[i] = r == (int)i ? 1 : 0;
}

__global__ void __launch_bounds__(128) dbscan_label_kernel(KArgs a, int R, float r2, const int* __restrict__ root,
                                                           const int* __restrict__ seed_rank, int32_t* __restrict__ labels) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= a.n) return;
  const float4 Q = __ldg(&a.sorted[p]);
  const int i = __float_as_int(Q.w);
  const uint32_t cid = __ldg(&a.keys[p]);
  if (cid >= a.g.n_cells) {  // non-finite point: its search finds nothing -> noise
    labels[i] = -2;
    return;
  }
  const int ri = root[i];
  if (ri >= 0) {
    labels[i] = seed_rank[ri] + 1;
    return;
  }
  const Stencil s = stencil_of_cell(a.g, cid, R);
  int best = 0x7FFFFFFF;
#pragma unroll 1
  for (int e2 = s.lo2; e2 <= s.hi2; ++e2)
#pragma unroll 1
    for (int e1 = s.lo1; e1 <= s.hi1; ++e1) {
      const uint32_t base = ((uint32_t)(s.zoff + e2) * (uint32_t)a.g.dim[1] + (uint32_t)e1) * (uint32_t)a.g.dim[0];
      const uint32_t beg = __ldg(&a.E[base + s.lo0]), end = __ldg(&a.E[base + s.hi0 + 1]);
#pragma unroll 1
      for (uint32_t j = beg; j < end; ++j) {
        const float4 P = __ldg(&a.sorted[j]);
        if (dist2_nofma(Q.x, Q.y, Q.z, P) < r2) {
          const int rj = root[__float_as_int(P.w)];
          if (rj >= 0 && rj < best) best = rj;
        }
      }
    }
  // reached by the cluster with the smallest seed, unless the seed loop got to this point first (state -2)
  labels[i] = best < i ? seed_rank[best] + 1 : -2;
}

// ---- per-cluster centroid / min / max / count ----
__device__ __forceinline__ uint32_t f2ord(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

struct BoxAcc {
  double sum[3];
  uint32_t lo[3], hi[3];
  int32_t count;
  int32_t pad;
};

__global__ void __launch_bounds__(256) boxes_init_kernel(BoxAcc* __restrict__ acc, int n_clusters) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_clusters) return;
  BoxAcc b;
  for (int k = 0; k < 3; ++k) {
    b.sum[k] = 0.0;
    b.lo[k] = f2ord(FLT_MAX);
    b.hi[k] = f2ord(-FLT_MAX);
  }
  b.count = 0;
  b.pad = 0;
  acc[c] = b;
}

__global__ void __launch_bounds__(256) boxes_accum_kernel(const float4* __restrict__ pts, const int32_t* __restrict__ labels, int64_t n,
                                                          int n_clusters, BoxAcc* __restrict__ acc) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = labels[i] - 1;
  if (c < 0 || c >= n_clusters) return;
  const float4 p = __ldg(pts + i);
  BoxAcc* b = acc + c;
  atomicAdd(&b->sum[0], (double)p.x);
  atomicAdd(&b->sum[1], (double)p.y);
  atomicAdd(&b->sum[2], (double)p.z);
  atomicMin(&b->lo[0], f2ord(p.x));
  atomicMin(&b->lo[1], f2ord(p.y));
  atomicMin(&b->lo[2], f2ord(p.z));
  atomicMax(&b->hi[0], f2ord(p.x));
  atomicMax(&b->hi[1], f2ord(p.y));
  atomicMax(&b->hi[2], f2ord(p.z));
  atomicAdd(&b->count, 1);
}

__global__ void __launch_bounds__(256) boxes_finish_kernel(const BoxAcc* __restrict__ acc, int n_clusters, float* __restrict__ centroid,
                                                           float* __restrict__ lo, float* __restrict__ hi, int32_t* __restrict__ counts) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_clusters) return;
  const BoxAcc b = acc[c];
  for (int k = 0; k < 3; ++k) {
    centroid[3 * c + k] = b.count ? (float)(b.sum[k] / (double)b.count) : 0.f;
    lo[3 * c + k] = ord2f(b.lo[k]);
    hi[3 * c + k] = ord2f(b.hi[k]);
  }
  if (counts) counts[c] = b.count;
}

// ---- mean shift: one warp per point, search around its CURRENT position in the grid of the ORIGINAL cloud ----
__device__ __forceinline__ bool query_range(const GridDesc& g, int a, double v, int R, int& lo, int& hi) {
  const int axis = g.ax[a];
  double t = (v - g.origin[axis]) * g.inv_cell;
  t = fmin(fmax(t, -1.0e9), 1.0e9);
  const int ci = (int)floor(t);
  lo = ci - R;
  if (lo == g.dim[a]) lo = g.dim[a] - 1;  // points exactly on the box's upper face live in the last cell
  lo = max(lo, 0);
  hi = min(ci + R, g.dim[a] - 1);
  return lo <= hi;
}

#define MS_WARPS 4
__global__ void __launch_bounds__(MS_WARPS * 32) mean_shift_kernel(KArgs a, const float4* __restrict__ cur, int R, float r2, int min_pts,
                                                                   float4* __restrict__ next, int32_t* __restrict__ counts) {
  const int lane = threadIdx.x & 31;
  const int64_t i = blockIdx.x * (int64_t)MS_WARPS + (threadIdx.x >> 5);
  if (i >= a.n) return;
  const float4 q = __ldg(cur + i);
  const GridDesc& g = a.g;
  int lo0, hi0, lo1, hi1, lo2, hi2;
  const float qc[3] = {q.x, q.y, q.z};
  bool ok = isfinite(q.x) && isfinite(q.y) && isfinite(q.z);
  ok = ok && query_range(g, 0, (double)qc[g.ax[0]], R, lo0, hi0);
  ok = ok && query_range(g, 1, (double)qc[g.ax[1]], R, lo1, hi1);
  ok = ok && query_range(g, 2, (double)qc[g.ax[2]], R, lo2, hi2);
  double sx = 0.0, sy = 0.0, sz = 0.0;
  int cnt = 0;
  if (ok) {
#pragma unroll 1
    for (int e2 = lo2; e2 <= hi2; ++e2)
#pragma unroll 1
      for (int e1 = lo1; e1 <= hi1; ++e1) {
        const uint32_t base = ((uint32_t)e2 * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
        const uint32_t beg = __ldg(&a.E[base + lo0]), end = __ldg(&a.E[base + hi0 + 1]);
#pragma unroll 1
        for (uint32_t j = beg + lane; j < end; j += 32) {
          const float4 P = __ldg(&a.sorted[j]);
          if (dist2_nofma(q.x, q.y, q.z, P) < r2) {
            const float4 C = __ldg(cur + __float_as_int(P.w));  // the neighbour's CURRENT position (:96-98)
            sx += (double)__fsub_rn(C.x, q.x);
            sy += (double)__fsub_rn(C.y, q.y);
            sz += (double)__fsub_rn(C.z, q.z);
            ++cnt;
          }
        }
      }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sx += __shfl_xor_sync(PCP_FULL, sx, o);
    sy += __shfl_xor_sync(PCP_FULL, sy, o);
    sz += __shfl_xor_sync(PCP_FULL, sz, o);
    cnt += __shfl_xor_sync(PCP_FULL, cnt, o);
  }
  if (lane == 0) {
    float4 o = q;
    if (cnt >= min_pts) {  // (:81-85)
      o.x = __fadd_rn(q.x, (float)(sx / (double)cnt));
      o.y = __fadd_rn(q.y, (float)(sy / (double)cnt));
      o.z = __fadd_rn(q.z, (float)(sz / (double)cnt));
    }
    next[i] = o;
    if (counts) counts[i] = cnt;
  }
}

KArgs make_kargs(pcp_ctx* ctx) {
  KArgs a;
  a.sorted = (const float4*)ctx->sorted_pts.p;
  a.keys = (const uint32_t*)ctx->keys_sorted.p;
  a.slot = (const int32_t*)ctx->sorted_slot.p;
  a.E = (const uint32_t*)ctx->cell_end.p;
  a.inv = (const uint32_t*)ctx->inv_pos.p;
  a.gid_of_ord = nullptr;
  a.g = ctx->grid;
  a.n = ctx->n;
  memset(&a.sp, 0, sizeof(a.sp));
  return a;
}

}  // namespace

int pcp_radius_count_impl(pcp_ctx* ctx, float r2, int R, int cap, int32_t* d_counts);

// labels: n int32 on the device; *d_n_clusters: one int32 on the device
int pcp_dbscan_impl(pcp_ctx* ctx, float r2, int R, int min_pts, int32_t* d_labels, int32_t* d_n_clusters) {
  const int64_t n = ctx->n;
  int rc;
  if ((rc = pcp_reserve(ctx, ctx->odom_counts, (size_t)n * sizeof(int32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->db_parent, (size_t)n * sizeof(int)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->db_root, (size_t)n * sizeof(int)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->db_flag, (size_t)(n + 1) * sizeof(int)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->db_rank, (size_t)(n + 1) * sizeof(int)))) return rc;
  int32_t* counts = (int32_t*)ctx->odom_counts.p;
  int *parent = (int*)ctx->db_parent.p, *root = (int*)ctx->db_root.p, *flag = (int*)ctx->db_flag.p, *rank = (int*)ctx->db_rank.p;
  if ((rc = pcp_radius_count_impl(ctx, r2, R, min_pts, counts))) return rc;
  const KArgs a = make_kargs(ctx);
  const unsigned b256 = (unsigned)((n + 255) / 256), b128 = (unsigned)((n + 127) / 128);
  dbscan_init_kernel<<<b256, 256, 0, ctx->stream>>>(parent, n);
  PCP_LAUNCH_CHECK("dbscan_init_kernel");
  dbscan_union_kernel<<<b128, 128, 0, ctx->stream>>>(a, R, r2, min_pts, counts, parent);
  PCP_LAUNCH_CHECK("dbscan_union_kernel");
  dbscan_flatten_kernel<<<b256, 256, 0, ctx->stream>>>(parent, counts, min_pts, n, root, flag);
  PCP_LAUNCH_CHECK("dbscan_flatten_kernel");
  PCP_CUDA(cudaMemsetAsync(flag + n, 0, sizeof(int), ctx->stream));
  size_t tmp = 0;
  PCP_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, flag, rank, (int)(n + 1), ctx->stream));
  if ((rc = pcp_reserve(ctx, ctx->scan_tmp, tmp))) return rc;
  PCP_CUDA(cub::DeviceScan::ExclusiveSum(ctx->scan_tmp.p, tmp, flag, rank, (int)(n + 1), ctx->stream));
  ctx->launches += 2;
  dbscan_label_kernel<<<b128, 128, 0, ctx->stream>>>(a, R, r2, root, rank, d_labels);
  PCP_LAUNCH_CHECK("dbscan_label_kernel");
  PCP_CUDA(cudaMemcpyAsync(d_n_clusters, rank + n, sizeof(int32_t), cudaMemcpyDeviceToDevice, ctx->stream));
  return PCP_OK;
}

int pcp_cluster_boxes_impl(pcp_ctx* ctx, const int32_t* d_labels, int n_clusters, float* d_centroid, float* d_lo, float* d_hi,
                           int32_t* d_counts) {
  int rc;
  if ((rc = pcp_reserve(ctx, ctx->db_boxes, (size_t)n_clusters * sizeof(BoxAcc)))) return rc;
  BoxAcc* acc = (BoxAcc*)ctx->db_boxes.p;
  const unsigned bc = (unsigned)((n_clusters + 255) / 256);
  boxes_init_kernel<<<bc, 256, 0, ctx->stream>>>(acc, n_clusters);
  PCP_LAUNCH_CHECK("boxes_init_kernel");
  boxes_accum_kernel<<<(unsigned)((ctx->n + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_cloud, d_labels, ctx->n, n_clusters, acc);
  PCP_LAUNCH_CHECK("boxes_accum_kernel");
  boxes_finish_kernel<<<bc, 256, 0, ctx->stream>>>(acc, n_clusters, d_centroid, d_lo, d_hi, d_counts);
  PCP_LAUNCH_CHECK("boxes_finish_kernel");
  return PCP_OK;
}

int pcp_mean_shift_impl(pcp_ctx* ctx, const float4* d_cur, float r2, int R, int min_pts, float4* d_next, int32_t* d_counts) {
  const KArgs a = make_kargs(ctx);
  const unsigned blocks = (unsigned)((ctx->n + MS_WARPS - 1) / MS_WARPS);
  mean_shift_kernel<<<blocks, MS_WARPS * 32, 0, ctx->stream>>>(a, d_cur, R, r2, min_pts, d_next, d_counts);
  PCP_LAUNCH_CHECK("mean_shift_kernel");
  return PCP_OK;
}
