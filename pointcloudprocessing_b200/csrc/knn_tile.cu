// knn_tile.cu — k-NN + PCA fast path: one warp per TILE of 32 consecutive cell-sorted queries, the
// tile's candidate set staged ONCE in shared memory, one lane per query, uniform control flow.
//
// Replaces, per point, nearestKSearch + computeCovarianceMatrixNormalized + SelfAdjointEigenSolver
// + the structure features (src/smoothing.cpp:239-246, src/pcaAnalysis.cpp:44-87, :29-31).
//
// The 32 queries of a tile lie in one stencil row of the grid (consecutive points of the cell-sorted
// array), so the union of their (2R+1)^3 stencils is (2R+1)^2 contiguous runs of that array.  The warp
// copies those runs into shared memory as float4 (coalesced global reads, every point read once per
// tile instead of once per query) and then every lane runs the SAME loop over the staged candidates
// (broadcast LDS.128, no divergence, unrolled):
//   pass 1  d2 (FLANN L2_Simple arithmetic, no FMA) -> 64 linear bins over [0, cover^2) in the lane's
//           own histogram column (bin 64 = beyond the radius the stencil is guaranteed to cover);
//   scan    the bin holding the k-th smallest; the lane succeeds when that bin is inside the cover
//           radius and at most 32 candidates lie at or below it;
//   pass 2  candidates at or below the bin are appended as keys bits(d2)<<32 | ord to the lane's column;
//   sort    32-input odd-even merge network on registers -> the (d2, index) order of the reference;
//   then the ordered moment, the fp64 Jacobi eigen-solve and the 68-byte record, all per lane.
// Lanes that do not succeed (sparse neighbourhood: k-th neighbour beyond the cover radius; a tile that
// crosses a row end; an over-full bin) are appended to a work list that the warp-per-query kernel of
// search_pca.cu finishes with its expanding-shell search.  Results are identical either way.
#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "eigen3.cuh"
#ifndef PCP_TILE_SYNC
#define PCP_TILE_SYNC 1
#endif
#include "pcp_internal.cuh"
#include "search_common.cuh"
#include "knn_finish.cuh"
#include "tile_common.cuh"

#ifdef PCP_STATS
__device__ unsigned long long g_tile_stats[16];
#define TSTAT(i, v) atomicAdd(&g_tile_stats[i], (unsigned long long)(v))
#else
#define TSTAT(i, v) do { } while (0)
#endif

namespace {

#ifndef PCP_TA_WARPS
#define PCP_TA_WARPS 4
#endif
constexpr int TA_WARPS = PCP_TA_WARPS;  // warps per CTA (2 / 4 / 8 measured: see DESIGN §6)
constexpr int TA_CH = 128;         // staged candidates per chunk
// CAP (template parameter: 24 or 32 for k-NN, 32 or 64 for radius search) = candidates a lane can hold = width of the sorting network;
// the histogram has BINS = 2*CAP - 1 linear d2 bins inside the cover radius + 1 overflow bin, so that it
// occupies exactly the rows the candidate keys use afterwards.
constexpr int TA_REC_WORDS = 17;   // 68-byte record
constexpr int TA_MAXGROUPS = 1;    // stencil rows (row groups) served per tile by default; lanes of further rows are
                                   // deferred (measured: a second pass for a few lanes costs more than deferring them)

template <int CAP>
struct __align__(16) TaSmem {
  float4 cand[TA_CH];                    // staged candidates of the current chunk
  // One 32-bit column per lane, two uses: pass 1 counts into hist[bin][lane] (bin 64 = beyond the cover
  // radius, BINS = 2*CAP-1); pass 2 stores candidate j's key as lo word plane[j][lane] and hi word plane[CAP + j][lane].
  // Both live in the lane's OWN column, so a lane of a later row group can histogram while the lanes
  // already served keep their candidates.  The record staging reuses the space at the end.
  union U {
    uint32_t plane[2 * CAP][32];
    float rec[32][TA_REC_WORDS];
  } u;
  uint32_t run_beg[TA_MAXROWS];          // first sorted position of each stencil row's run
  uint32_t run_off[TA_MAXROWS + 1];      // exclusive prefix of the run lengths
};

// copies candidates [cb, cb + cnt) of the tile's virtual candidate sequence (the stencil rows' runs laid
// end to end) into sm.cand: row by row, the 32 lanes reading consecutive float4 of the run
template <int CAP>
__device__ __forceinline__ void stage_chunk(const float4* __restrict__ sorted, TaSmem<CAP>& sm, int nrows, uint32_t cb, int cnt,
                                            int lane) {
  const uint32_t ce = cb + (uint32_t)cnt;
#pragma unroll 1
  for (int r = 0; r < nrows; ++r) {
    const uint32_t off = sm.run_off[r], end = sm.run_off[r + 1];
    const uint32_t lo = max(off, cb), hi = min(end, ce);
    if (lo >= hi) continue;
    const float4* __restrict__ src = sorted + sm.run_beg[r] - off;  // src[v] = candidate v
#pragma unroll 1
    for (uint32_t v = lo + (uint32_t)lane; v < hi; v += 32) sm.cand[v - cb] = __ldg(&src[v]);
  }
  __syncwarp();
}

// MODE 0: k nearest (two passes: histogram select, then collect).  MODE 1: radius search — membership is
// d2 < r2 itself, so there is no selection pass; a lane is served when its neighbourhood fits its column.
template <int CAP, int MODE>
__global__ void __launch_bounds__(TA_WARPS * 32, (CAP == 64 ? 12 : 20) / TA_WARPS) knn_tile_kernel(KArgs a, int R0, float grow, int max_groups, uint32_t* __restrict__ work_count,
                                                                 uint32_t* __restrict__ work_list) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int BINS = 2 * CAP - 1;
  TaSmem<CAP>& sm = reinterpret_cast<TaSmem<CAP>*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int64_t pos0 = ((int64_t)blockIdx.x * TA_WARPS + (threadIdx.x >> 5)) * 32;
  const bool live_warp = pos0 < a.n;  // a tail warp stays (inactive) for the CTA barriers below
  const int64_t p = live_warp ? pos0 + lane : a.n;
  const GridDesc& g = a.g;
  const SearchParams& sp = a.sp;
  const float4* __restrict__ sorted = a.sorted;
  const int k = sp.k;
  const int my_slot = p < a.n ? __ldg(&a.slot[p]) : -1;  // -1: tail lane or halo point (neighbour only)
  const float4 Q = p < a.n ? __ldg(&sorted[p]) : make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t cid = p < a.n ? __ldg(&a.keys[p]) : g.n_cells;
  const bool active = my_slot >= 0;
  const bool finite = active && cid < g.n_cells;
  uint32_t* const hcol = &sm.u.plane[0][lane];
  uint32_t* wp = hcol;       // next free entry of this lane's candidate column (lo word; hi word CAP rows on)
  const uint32_t my_row = cid / (uint32_t)g.dim[0];
  const int c0 = (int)(cid - my_row * (uint32_t)g.dim[0]);
  bool pending = finite;     // not yet served by a row group
  bool ok = false;           // the lane's k nearest are in its column
  bool far = false;          // k-NN: the k-th neighbour lies beyond the block of half-width 1 (told to the work-list kernel)

  // ---- a tile usually lies in ONE stencil row; where it crosses row ends, each row is a group of its own
#pragma unroll 1
  for (int grp = 0; grp < max_groups; ++grp) {
    const unsigned pmask = __ballot_sync(PCP_FULL, pending);
    if (pmask == 0) break;
    // the group's row: the one most pending lanes share (a tile that crosses a row end keeps its larger part)
    const unsigned peers = __match_any_sync(PCP_FULL, pending ? my_row : 0xFFFFFFFFu);
    const int votes = pending ? __popc(peers) : 0;
    const int best_votes = __reduce_max_sync(PCP_FULL, votes);
    const unsigned lead = __ballot_sync(PCP_FULL, pending && votes == best_votes);
    const uint32_t row0 = __shfl_sync(PCP_FULL, my_row, __ffs(lead) - 1);
    const bool in_grp = pending && my_row == row0;
    pending = pending && !in_grp;
    const int c1 = (int)(row0 % (uint32_t)g.dim[1]);
    int c2 = (int)(row0 / (uint32_t)g.dim[1]);
    const int zoff = pcp_unfold_z(g, c2);  // frame batches: the frame's first z row; c2 is now frame-local
    const int c0min = __reduce_min_sync(PCP_FULL, in_grp ? c0 : 0x7fffffff);
    const int c0max = __reduce_max_sync(PCP_FULL, in_grp ? c0 : -1);
    int R = R0, nrows = 0, bstar = -1, cum_star = 0;
    uint32_t total = 0;
    float Sb = 0.f, noff = 0.f;
    if (MODE == 1) {
      // radius: the stencil of half-width R0 covers the radius by construction (stencil_for_radius)
      bool whole;
      float cover_cells;
      total = tile_setup(a, sm.run_beg, sm.run_off, lane, R, c0min, c0max, c1, c2, zoff, c0, Q, nrows, whole, cover_cells);
      Sb = 1.f;
      bstar = 0;
    } else {
    // ---- pass 1 (+ on-demand growth of the stencil): histogram of d2 over the staged candidates.
    // A lane's bin coordinate is fma(d2, Sb, noff): first BINS linear bins over [0, cover^2); a lane whose
    // threshold bin holds too many candidates for its column (d_k << cover: dense scan lines next to a
    // sparse surrounding, e.g. the near field of a VLP-16 frame) gets ONE refinement pass with the window
    // narrowed to that bin (candidates below the window count in bin 0 and are all collected).
    bool hist_on = in_grp, fresh = true, refined = false, whole = false;
#pragma unroll 1
    for (;;) {
      if (fresh) {
        float cover_cells;
        total = tile_setup(a, sm.run_beg, sm.run_off, lane, R, c0min, c0max, c1, c2, zoff, c0, Q, nrows, whole, cover_cells);
        Sb = (float)BINS / (cover_cells * cover_cells * g.cell * g.cell * (1.f - 1e-6f));
        noff = 0.f;
        hist_on = in_grp;
      }
      if (hist_on) {
#pragma unroll
        for (int b = 0; b <= BINS; ++b) hcol[b * 32] = 0;
      }
      __syncwarp();
#pragma unroll 1
      for (uint32_t cb = 0; cb < total; cb += TA_CH) {
        const int cnt = (int)min((uint32_t)TA_CH, total - cb);
        stage_chunk(sorted, sm, nrows, cb, cnt, lane);
        // four candidates per step: the loads and the distance arithmetic of a group are independent, and
        // the histogram update is a fire-and-forget shared atomic (own column: no conflicts, no read-back)
        int j = 0;
#pragma unroll 2
        for (; j + 4 <= cnt; j += 4) {
          const float4 P0 = sm.cand[j], P1 = sm.cand[j + 1], P2 = sm.cand[j + 2], P3 = sm.cand[j + 3];
          const uint32_t b0 = __float2uint_rz(fminf(__fmaf_rn(dist2_nofma(Q.x, Q.y, Q.z, P0), Sb, noff), (float)BINS));
          const uint32_t b1 = __float2uint_rz(fminf(__fmaf_rn(dist2_nofma(Q.x, Q.y, Q.z, P1), Sb, noff), (float)BINS));
          const uint32_t b2 = __float2uint_rz(fminf(__fmaf_rn(dist2_nofma(Q.x, Q.y, Q.z, P2), Sb, noff), (float)BINS));
          const uint32_t b3 = __float2uint_rz(fminf(__fmaf_rn(dist2_nofma(Q.x, Q.y, Q.z, P3), Sb, noff), (float)BINS));
          if (hist_on) {
            atomicAdd(&hcol[b0 * 32], 1u);
            atomicAdd(&hcol[b1 * 32], 1u);
            atomicAdd(&hcol[b2 * 32], 1u);
            atomicAdd(&hcol[b3 * 32], 1u);
          }
        }
        for (; j < cnt; ++j) {
          const uint32_t b0 = __float2uint_rz(fminf(__fmaf_rn(dist2_nofma(Q.x, Q.y, Q.z, sm.cand[j]), Sb, noff), (float)BINS));
          if (hist_on) atomicAdd(&hcol[b0 * 32], 1u);
        }
        __syncwarp();
      }
      // the bin holding the k-th smallest, and how many candidates lie at or below it
      if (hist_on) {
        int cum = 0;
        bstar = -1;
#pragma unroll 8
        for (int b = 0; b < BINS; ++b) {
          cum += (int)hcol[b * 32];
          if (bstar < 0 && cum >= k) {
            bstar = b;
            cum_star = cum;
          }
        }
      }
      if (refined) break;
      {
        // threshold bin over-full: narrow the lane's window to [bstar - 0.02, bstar + 1.03) old bins and
        // count again (negative coordinates convert to bin 0; the window holds the k-th by construction)
        const bool refine = hist_on && bstar >= 0 && cum_star > CAP;
        if (__any_sync(PCP_FULL, refine)) {
          TSTAT(9, refine ? 1 : 0);
          refined = true;
          fresh = false;
          hist_on = refine;
          if (refine) {
            const float z = (float)BINS * (1.f / 1.05f);
            noff = -((float)bstar - 0.02f) * z;
            Sb *= z;
          }
          continue;
        }
      }
      // lanes whose k-th neighbour lies beyond the cover radius: widen the stencil when a pass over the
      // wider union costs less than finishing them one by one (~2000 vs ~28 instructions per unit)
      const int n_unc = __popc(__ballot_sync(PCP_FULL, in_grp && bstar < 0));
      if (n_unc == 0 || R >= TA_MAXR || whole) break;
      const int span = c0max - c0min + 1;
      const float wider = (float)total * (float)((span + 2 * R + 2) * (2 * R + 3)) / (float)((span + 2 * R) * (2 * R + 1));
      if ((float)n_unc * grow < wider) break;
      TSTAT(7, lane == 0 ? 1 : 0);
      ++R;
    }
    }
    TSTAT(2, lane == 0 ? total : 0);
    TSTAT(5, (in_grp && bstar < 0) ? 1 : 0);                          // deferred: k-th beyond the cover radius
    TSTAT(6, (in_grp && bstar >= 0 && cum_star > CAP) ? 1 : 0);    // deferred: threshold bin too full
    bool grp_ok = in_grp && bstar >= 0 && cum_star <= CAP;
    if (MODE == 0 && in_grp && bstar < 0 && R >= 1) far = true;  // d_k > the lane's cover radius >= one cell
    // the collect test is d2 * Sb < thr_bin: the bin bound (k-NN) or, with Sb = 1, r2 itself (radius: strict
    // d2 < r2 as FLANN's RadiusResultSet); other lanes collect nothing
    const float thr_bin = grp_ok ? (MODE == 1 ? sp.r2 : (float)(bstar + 1)) : -3.0e38f;
    uint32_t* const wend = hcol + CAP * 32;
    __syncwarp();
    // ---- pass 2: collect the candidates at or below the threshold bin (same d2 arithmetic, same bins)
    if (__any_sync(PCP_FULL, grp_ok)) {
      const bool resident = MODE == 0 && total <= (uint32_t)TA_CH;  // one chunk: still staged from pass 1
      // A candidate farther from the bounding box of the served queries than the largest threshold
      // distance is nobody's neighbour: each chunk is compacted in place (one candidate per lane, ballot
      // + rank) before the per-lane loop, which then runs over about half of it.
      const float big = 3.0e38f;
      float bx0 = grp_ok ? Q.x : big, bx1 = grp_ok ? Q.x : -big, by0 = grp_ok ? Q.y : big, by1 = grp_ok ? Q.y : -big;
      float bz0 = grp_ok ? Q.z : big, bz1 = grp_ok ? Q.z : -big;
      float tmax = grp_ok ? (thr_bin - noff) / Sb : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        bx0 = fminf(bx0, __shfl_xor_sync(PCP_FULL, bx0, o));
        bx1 = fmaxf(bx1, __shfl_xor_sync(PCP_FULL, bx1, o));
        by0 = fminf(by0, __shfl_xor_sync(PCP_FULL, by0, o));
        by1 = fmaxf(by1, __shfl_xor_sync(PCP_FULL, by1, o));
        bz0 = fminf(bz0, __shfl_xor_sync(PCP_FULL, bz0, o));
        bz1 = fmaxf(bz1, __shfl_xor_sync(PCP_FULL, bz1, o));
        tmax = fmaxf(tmax, __shfl_xor_sync(PCP_FULL, tmax, o));
      }
      tmax *= 1.001f;  // slack for the rounding of thr_bin / Sb and of the box distance
#pragma unroll 1
      for (uint32_t cb = 0; cb < total; cb += TA_CH) {
        int n_ch = (int)min((uint32_t)TA_CH, total - cb);
        if (!resident) stage_chunk(sorted, sm, nrows, cb, n_ch, lane);
        {
          int kept = 0;
#pragma unroll 1
          for (int i0 = 0; i0 < n_ch; i0 += 32) {
            const int i = i0 + lane;
            float4 P = make_float4(0.f, 0.f, 0.f, 0.f);
            bool keep = false;
            if (i < n_ch) {
              P = sm.cand[i];
              const float dx = fmaxf(fmaxf(bx0 - P.x, P.x - bx1), 0.f), dy = fmaxf(fmaxf(by0 - P.y, P.y - by1), 0.f);
              const float dz = fmaxf(fmaxf(bz0 - P.z, P.z - bz1), 0.f);
              keep = dx * dx + dy * dy + dz * dz <= tmax;
            }
            const unsigned km = __ballot_sync(PCP_FULL, keep);
            __syncwarp();  // every lane has read its candidate before any slot is overwritten
            if (keep) sm.cand[kept + __popc(km & ((1u << lane) - 1u))] = P;
            kept += __popc(km);
          }
          __syncwarp();
          TSTAT(8, lane == 0 ? kept : 0);
          n_ch = kept;
        }
        int j = 0;
#pragma unroll 2
        for (; j + 4 <= n_ch; j += 4) {
          const float4 P0 = sm.cand[j], P1 = sm.cand[j + 1], P2 = sm.cand[j + 2], P3 = sm.cand[j + 3];
          const float e0 = dist2_nofma(Q.x, Q.y, Q.z, P0), e1 = dist2_nofma(Q.x, Q.y, Q.z, P1);
          const float e2 = dist2_nofma(Q.x, Q.y, Q.z, P2), e3 = dist2_nofma(Q.x, Q.y, Q.z, P3);
          if (__fmaf_rn(e0, Sb, noff) < thr_bin) { if (MODE == 0 || wp < wend) { wp[0] = __float_as_uint(P0.w); wp[CAP * 32] = __float_as_uint(e0); } wp += 32; }
          if (__fmaf_rn(e1, Sb, noff) < thr_bin) { if (MODE == 0 || wp < wend) { wp[0] = __float_as_uint(P1.w); wp[CAP * 32] = __float_as_uint(e1); } wp += 32; }
          if (__fmaf_rn(e2, Sb, noff) < thr_bin) { if (MODE == 0 || wp < wend) { wp[0] = __float_as_uint(P2.w); wp[CAP * 32] = __float_as_uint(e2); } wp += 32; }
          if (__fmaf_rn(e3, Sb, noff) < thr_bin) { if (MODE == 0 || wp < wend) { wp[0] = __float_as_uint(P3.w); wp[CAP * 32] = __float_as_uint(e3); } wp += 32; }
        }
        for (; j < n_ch; ++j) {
          const float4 P0 = sm.cand[j];
          const float e0 = dist2_nofma(Q.x, Q.y, Q.z, P0);
          if (__fmaf_rn(e0, Sb, noff) < thr_bin) { if (MODE == 0 || wp < wend) { wp[0] = __float_as_uint(P0.w); wp[CAP * 32] = __float_as_uint(e0); } wp += 32; }
        }
        __syncwarp();
      }
    }
    if (MODE == 1) grp_ok = grp_ok && wp <= wend;  // a larger neighbourhood than the column holds: deferred
    ok = ok || grp_ok;
  }
  const int cnt = (int)(wp - hcol) >> 5;
  TSTAT(0, finite ? 1 : 0);
  TSTAT(1, (finite && !ok) ? 1 : 0);
  TSTAT(3, ok ? cnt : 0);
  TSTAT(4, pending ? 1 : 0);  // deferred: more row groups in the tile than the kernel serves
  // ---- deferred lanes -> work list (sorted positions), finished by the expanding-shell kernel
  {
    const bool defer = finite && !ok;
    const unsigned dm = __ballot_sync(PCP_FULL, defer);
    if (dm != 0) {
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(work_count, (uint32_t)__popc(dm));
      base = __shfl_sync(PCP_FULL, base, 0);
      if (defer) work_list[base + __popc(dm & ((1u << lane) - 1u))] = (uint32_t)p | (far ? 0x80000000u : 0u);
    }
  }
  // The sorting network and the finish stage are ~50 KB of straight-line code walked once per tile: the
  // warps of a CTA re-align here so that they fetch it together (one miss serves four warps) instead of
  // evicting each other's lines (PCP_TILE_SYNC=0 at build time to compare).
#if PCP_TILE_SYNC
  __syncthreads();
#endif
  // ---- order the collected keys by (d2, index)
  u64 keys[CAP];
#pragma unroll
  for (int j = 0; j < CAP; ++j) keys[j] = j < cnt ? (((u64)hcol[(CAP + j) * 32] << 32) | (u64)hcol[j * 32]) : PCP_KEY_INF;
  if (CAP == 32) sort32_regs(reinterpret_cast<u64(&)[32]>(keys));
  else sort64_regs(reinterpret_cast<u64(&)[64]>(keys));
  // the ordered keys go back to the lane's column: everything after this point walks the list with rolled loops
#pragma unroll
  for (int j = 0; j < CAP; ++j) {
    hcol[j * 32] = (uint32_t)keys[j];
    hcol[(CAP + j) * 32] = (uint32_t)(keys[j] >> 32);
  }
  const int n_nb = ok ? (MODE == 1 ? cnt : k) : 0;
  const bool emit = active && (ok || !finite);  // non-finite queries get their empty record here
  __syncwarp();
#if PCP_TILE_SYNC
  __syncthreads();
#endif
  if (!live_warp) return;
  knn_finish(a, hcol, hcol + CAP * 32, 32, n_nb, MODE == 1 ? sp.min_pts : 1, Q, emit ? my_slot : -1, sm.u.rec, lane);
}

}  // namespace

#ifdef PCP_STATS
extern "C" __attribute__((visibility("default"))) int pcp_debug_tile_stats(unsigned long long* out) {
  unsigned long long z[16] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_tile_stats, sizeof(z));
  cudaMemcpyToSymbol(g_tile_stats, z, sizeof(z));
  return 0;
}
#endif

// Launches the tile kernel over the whole cloud; deferred queries are left in ctx->work_list /
// ctx->work_count (device) for pcp_search_launch_worklist.
int pcp_knn_tile_launch(pcp_ctx* ctx, const SearchParams& sp, const KArgs& a, int radius_cap) {
  int rc;
  if ((rc = pcp_reserve(ctx, ctx->work_list, (size_t)ctx->n * sizeof(uint32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->work_count, 2 * sizeof(uint32_t)))) return rc;  // [0] list length, [1] ticket counter of the work-list kernel
  PCP_CUDA(cudaMemsetAsync(ctx->work_count.p, 0, 2 * sizeof(uint32_t), ctx->stream));
  const int64_t tiles = (ctx->n + 31) / 32;
  const int blocks = (int)((tiles + TA_WARPS - 1) / TA_WARPS);
  const float grow = 0.f;  // never widen a tile's stencil (measured slower on the urban cloud)
  // Stencil rows served per tile.  One on big clouds (a second row group costs the urban cloud 5-10 %);
  // up to four when the rows are short — frame batches: 26 k points over a few hundred rows, where 8-30 %
  // of the lanes sit in another row than their tile's majority (measured 8.8 -> 7.1 ms per 128 frames).
  const double pts_per_row = (double)ctx->n / ((double)ctx->grid.n_seg * (double)ctx->grid.dim[1] * (double)ctx->grid.dim[2]);
  int max_groups = pts_per_row < 256.0 ? 4 : TA_MAXGROUPS;
  if (ctx->knobs.knn_groups > 0) max_groups = ctx->knobs.knn_groups;
  uint32_t* wc = (uint32_t*)ctx->work_count.p;
  uint32_t* wl = (uint32_t*)ctx->work_list.p;
  if (sp.mode == 1 && radius_cap == 64) {
    const size_t smem = sizeof(TaSmem<64>) * TA_WARPS;
    PCP_CUDA(cudaFuncSetAttribute(knn_tile_kernel<64, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PCP_CUDA(pcp_launch_search(ctx, knn_tile_kernel<64, 1>, dim3(blocks), dim3(TA_WARPS * 32), smem, a, sp.R0, 0.f, max_groups, wc, wl));
  } else if (sp.mode == 1) {
    const size_t smem = sizeof(TaSmem<32>) * TA_WARPS;
    PCP_CUDA(cudaFuncSetAttribute(knn_tile_kernel<32, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PCP_CUDA(pcp_launch_search(ctx, knn_tile_kernel<32, 1>, dim3(blocks), dim3(TA_WARPS * 32), smem, a, sp.R0, 0.f, max_groups, wc, wl));
  } else {
    const size_t smem = sizeof(TaSmem<32>) * TA_WARPS;
    PCP_CUDA(cudaFuncSetAttribute(knn_tile_kernel<32, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PCP_CUDA(pcp_launch_search(ctx, knn_tile_kernel<32, 0>, dim3(blocks), dim3(TA_WARPS * 32), smem, a, sp.R0, grow, max_groups, wc, wl));
  }
  PCP_LAUNCH_CHECK("knn_tile_kernel");
  return PCP_OK;
}
