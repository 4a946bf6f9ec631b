// radius_heavy.cu — neighbour LISTS of radius neighbourhoods larger than the 512-entry on-chip buffer of the
// warp-per-query kernel (search_pca.cu): count-then-fill through global memory.
//
// The reference's own radii produce such lists: radiusSearch(i, 3.0) at src/lidarOdometry.cpp:76 and main.cpp:117,
// radiusSearch(i, 2.0) at src/buildingFacadeExtraction.cpp:83 — thousands of hits per query on a dense scan.
// pcp_radius_neighbors is a two-call protocol (counts, then lists at caller-built offsets), so every list's place and
// size is known when it is filled: a heavy query's hits are appended unsorted to a key array at its offset
// (heavy_fill_kernel, one warp per heavy query, keys bits(d2) << 32 | ord as everywhere), the heavy segments are
// sorted by cub::DeviceSegmentedSort — (d2, index) order, the reference's (FLANN DistanceIndex) — and written out as
// index / distance (heavy_emit_kernel).  Lists of up to 512 hits keep coming from the on-chip path.
#include <cub/device/device_segmented_sort.cuh>

#include "pcp_internal.cuh"
#include "search_common.cuh"

namespace {

constexpr int HEAVY_MIN = 513;  // the on-chip path holds RADIUS_CAP = 512 hits (search_pca.cu)

// one warp per sorted position; only queries whose list is longer than the on-chip buffer do anything
__global__ void __launch_bounds__(256) heavy_fill_kernel(KArgs a, float r2, int R, const int64_t* __restrict__ offsets, u64* __restrict__ keys,
                                                         int64_t* __restrict__ seg_beg, int64_t* __restrict__ seg_end, int* __restrict__ n_seg) {
  const int64_t p = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (p >= a.n) return;
  const int slot = __ldg(&a.slot[p]);
  if (slot < 0) return;
  const int64_t beg = offsets[slot], end = offsets[slot + 1];
  if (end - beg < HEAVY_MIN) return;
  const GridDesc& g = a.g;
  const uint32_t cid = __ldg(&a.keys[p]);
  if (cid >= g.n_cells) return;
  const float4 Q = __ldg(&a.sorted[p]);
  const int c0 = (int)(cid % (uint32_t)g.dim[0]);
  const uint32_t tq = cid / (uint32_t)g.dim[0];
  const int c1 = (int)(tq % (uint32_t)g.dim[1]);
  int c2 = (int)(tq / (uint32_t)g.dim[1]);
  const int zoff = pcp_unfold_z(g, c2);
  const int lo0 = max(c0 - R, 0), hi0 = min(c0 + R, g.dim[0] - 1);
  const int lo1 = max(c1 - R, 0), hi1 = min(c1 + R, g.dim[1] - 1);
  const int lo2 = max(c2 - R, 0), hi2 = min(c2 + R, g.dim[2] - 1);
  int64_t w = beg;  // next free entry of this query's segment (same on every lane)
  for (int e2 = lo2; e2 <= hi2; ++e2)
    for (int e1 = lo1; e1 <= hi1; ++e1) {
      const uint32_t base = ((uint32_t)(zoff + e2) * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
      const uint32_t b = __ldg(&a.E[base + lo0]), e = __ldg(&a.E[base + hi0 + 1]);
      for (uint32_t j0 = b; j0 < e; j0 += 32) {
        const uint32_t j = j0 + (uint32_t)lane;
        bool hit = false;
        u64 key = 0;
        if (j < e) {
          const float4 P = __ldg(&a.sorted[j]);
          const float d2 = dist2_nofma(Q.x, Q.y, Q.z, P);
          hit = d2 < r2;  // strict, as FLANN's RadiusResultSet
          key = ((u64)__float_as_uint(d2) << 32) | (u64)__float_as_uint(P.w);
        }
        const unsigned m = __ballot_sync(PCP_FULL, hit);
        if (hit) {
          const int64_t at = w + __popc(m & ((1u << lane) - 1u));
          if (at < end) keys[at] = key;  // (the count of the first call is the same predicate: always true)
        }
        w += __popc(m);
      }
    }
  if (lane == 0) {
    const int s = atomicAdd(n_seg, 1);
    seg_beg[s] = beg;
    seg_end[s] = min(w, end);
  }
}

// sorted keys of the heavy segments -> the caller's index / distance arrays
__global__ void __launch_bounds__(256) heavy_emit_kernel(const u64* __restrict__ keys, const int64_t* __restrict__ seg_beg,
                                                         const int64_t* __restrict__ seg_end, int n_seg, const int32_t* __restrict__ gid_of_ord,
                                                         int32_t* __restrict__ idx, float* __restrict__ d2) {
  const int s = blockIdx.x;
  if (s >= n_seg) return;
  const int64_t b = seg_beg[s], e = seg_end[s];
  for (int64_t i = b + threadIdx.x; i < e; i += blockDim.x) {
    const u64 key = keys[i];
    const int ord = (int)(uint32_t)key;
    idx[i] = gid_of_ord ? __ldg(&gid_of_ord[ord]) : ord;
    if (d2) d2[i] = __uint_as_float((uint32_t)(key >> 32));
  }
}

}  // namespace

// Fills the lists of every query with more than 512 hits.  d_offsets: n_query + 1 entries (device); d_idx / d_d2: the
// caller's arrays (device), total = offsets[n_query] entries.  Synchronises once (the number of heavy segments).
int pcp_radius_heavy_lists_impl(pcp_ctx* ctx, const KArgs& a, float r2, int R, const int64_t* d_offsets, int64_t total, int32_t* d_idx,
                                float* d_d2) {
  if (total <= 0) return PCP_OK;
  if (total >= (int64_t)0x7FFFFFF0) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_radius_neighbors: more than 2^31 neighbours in one call");
  int rc;
  const int64_t nq = ctx->n_query;
  const int64_t max_seg = total / HEAVY_MIN + 1;  // every heavy segment holds at least HEAVY_MIN entries
  if ((rc = pcp_reserve(ctx, ctx->heavy_keys, (size_t)total * sizeof(u64)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->heavy_keys2, (size_t)total * sizeof(u64)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->heavy_seg, (size_t)(2 * max_seg) * sizeof(int64_t) + 16))) return rc;
  (void)nq;
  int* n_seg = (int*)ctx->heavy_seg.p;
  int64_t* seg_beg = (int64_t*)((char*)ctx->heavy_seg.p + 16);
  int64_t* seg_end = seg_beg + max_seg;
  PCP_CUDA(cudaMemsetAsync(n_seg, 0, sizeof(int), ctx->stream));
  u64* keys = (u64*)ctx->heavy_keys.p;
  u64* keys2 = (u64*)ctx->heavy_keys2.p;
  heavy_fill_kernel<<<(unsigned)((ctx->n * 32 + 255) / 256), 256, 0, ctx->stream>>>(a, r2, R, d_offsets, keys, seg_beg, seg_end, n_seg);
  PCP_LAUNCH_CHECK("heavy_fill_kernel");
  int h_seg = 0;
  PCP_CUDA(cudaMemcpyAsync(&h_seg, n_seg, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  if (h_seg <= 0) return PCP_OK;
  size_t tmp = 0;
  PCP_CUDA(cub::DeviceSegmentedSort::SortKeys(nullptr, tmp, keys, keys2, (int)total, h_seg, seg_beg, seg_end, ctx->stream));
  if ((rc = pcp_reserve(ctx, ctx->sort_tmp, tmp))) return rc;
  PCP_CUDA(cub::DeviceSegmentedSort::SortKeys(ctx->sort_tmp.p, tmp, keys, keys2, (int)total, h_seg, seg_beg, seg_end, ctx->stream));
  ctx->launches += 3;
  heavy_emit_kernel<<<h_seg, 256, 0, ctx->stream>>>(keys2, seg_beg, seg_end, h_seg, a.gid_of_ord, d_idx, d_d2);
  PCP_LAUNCH_CHECK("heavy_emit_kernel");
  return PCP_OK;
}
