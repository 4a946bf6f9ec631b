// knn_pca.cu — fused k-nearest search + ordered moment + 3×3 eigen + features, ONE THREAD PER QUERY.
//
// Replaces, per point, what pcl::NormalEstimation::setKSearch (src/smoothing.cpp:239-246) and the
// pcaAnalysis loop body (src/pcaAnalysis.cpp:44-87, :29-31) do on the CPU: the k nearest neighbours
// in FLANN's (d2, index) order, pcl::computeCovarianceMatrixNormalized in its accumulation order,
// the symmetric 3×3 eigen-decomposition and the three structure features — without the neighbour
// list ever leaving the SM.
//
// Work decomposition (v4).  A warp owns 32 consecutive cell-sorted points and lane l solves query l
// from start to finish; the lanes of a warp are spatial neighbours, so they walk (almost) the same
// stencil rows and read (almost) the same candidate addresses at the same time: the float4 loads
// coalesce into a few sectors and the loops stay in step without any shuffle, ballot or warp sync.
// Selection is exact and two-pass:
//   pass 1  quantises every candidate's d2 monotonically to 22 bits and counts it into a 64-bin
//           histogram (one shared-memory column per lane, conflict free); a bin scan finds the bin
//           holding the k-th smallest; if that bin is too full the pass repeats on the bin alone
//           (radix select, 6 bits per round) — in practice one round;
//   pass 2  re-walks the rows (now clipped by the threshold) and appends the <= 32 candidates below
//           the threshold as 64-bit keys (bits(d2) << 32 | ord) to the lane's column;
//   sort    a 32-input odd-even merge network on registers orders them by (d2, index).
// Coverage: the answer is final when the k-th distance lies inside the scanned block; otherwise the
// lane scans the shell out to ceil(d_k / cell) with the current k-th key as the bound.  Lanes whose
// candidate set cannot be narrowed to 32 (heavy exact ties) take an exact insertion path.
#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "eigen3.cuh"
#include "pcp_internal.cuh"
#include "search_common.cuh"
#include "knn_finish.cuh"

#ifdef PCP_STATS
__device__ unsigned long long g_tq_stats[16];
#define TQSTAT(i, v) atomicAdd(&g_tq_stats[i], (unsigned long long)(v))
#else
#define TQSTAT(i, v) do { } while (0)
#endif

namespace {

constexpr int TQ_WARPS = 4;       // warps per CTA
constexpr int TQ_BINS = 64;       // histogram bins per selection round
constexpr int TQ_CAP = 32;        // candidates a lane can hold (sorting-network width)
constexpr int TQ_QBITS = 22;      // quantised d2 width
constexpr int TQ_REC_WORDS = 17;  // 68-byte record

// one warp's shared memory: the three views are used in disjoint phases
union TqSmem {
  uint32_t hist[TQ_BINS][32];     // pass 1: hist[bin][lane]
  u64 pend[TQ_CAP][32];           // pass 2: pend[j][lane]
  float rec[32][TQ_REC_WORDS];    // record staging
};

// A lane's query in grid terms.
struct QCell {
  int c0, c1, c2;     // own cell along the permuted axes
  float f0, f1, f2;   // fractional position inside it, [0,1]
};

// Visits every point of the cells of the block of half-width R around the query's cell that are NOT
// in the block of half-width Rp (Rp < 0: whole block).  With a finite `bound2` rows are dropped and
// runs clipped by their distance to the query, never excluding a point with d2 <= bound2 (the slack
// terms are those of the v3 kernel, proven by the brute-force parity tests).
template <class F>
__device__ __forceinline__ void scan_region(const KArgs& a, const QCell& q, int Rp, int R, float bound2, F&& fn) {
  const GridDesc& g = a.g;
  const float4* __restrict__ sorted = a.sorted;
  const uint32_t* __restrict__ E = a.E;
  const int lo0 = max(q.c0 - R, 0), hi0 = min(q.c0 + R, g.dim[0] - 1);
  const int lo1 = max(q.c1 - R, 0), hi1 = min(q.c1 + R, g.dim[1] - 1);
  const int lo2 = max(q.c2 - R, 0), hi2 = min(q.c2 + R, g.dim[2] - 1);
  const bool prune = bound2 < 3.0e38f;
  const float inv_c2 = prune ? (float)(g.inv_cell * g.inv_cell) * bound2 * 1.0002f : 0.f;  // bound in cells^2, widened
#pragma unroll 1
  for (int e2 = lo2; e2 <= hi2; ++e2) {
    const int d2c = e2 - q.c2;
    const float g2 = cell_gap(d2c, q.f2);
#pragma unroll 1
    for (int e1 = lo1; e1 <= hi1; ++e1) {
      const int d1c = e1 - q.c1;
      int xlo = lo0, xhi = hi0;
      if (prune) {
        const float g1 = cell_gap(d1c, q.f1);
        const float rest = inv_c2 - (g1 * g1 + g2 * g2) * 0.9998f;
        if (rest < 0.f) continue;
        const float hx = sqrtf(rest) + 4e-3f;  // cells; offsets relative to c0 stay exact
        xlo = max(xlo, q.c0 + (int)floorf(q.f0 - hx));
        xhi = min(xhi, q.c0 + (int)floorf(q.f0 + hx));
        if (xlo > xhi) continue;
      }
      const uint32_t base = ((uint32_t)e2 * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
      uint32_t bA, eA, bB = 0, eB = 0;
      if (Rp >= 0 && abs(d1c) <= Rp && abs(d2c) <= Rp) {
        // this row crosses the already-scanned block: keep the parts left and right of it
        const int ahi = min(xhi, q.c0 - Rp - 1), blo = max(xlo, q.c0 + Rp + 1);
        bA = eA = 0;
        if (xlo <= ahi) {
          bA = __ldg(&E[base + xlo]);
          eA = __ldg(&E[base + ahi + 1]);
        }
        if (blo <= xhi) {
          bB = __ldg(&E[base + blo]);
          eB = __ldg(&E[base + xhi + 1]);
        }
      } else {
        bA = __ldg(&E[base + xlo]);
        eA = __ldg(&E[base + xhi + 1]);
      }
#pragma unroll 1
      for (uint32_t i = bA; i < eA; ++i) fn(__ldg(&sorted[i]));
#pragma unroll 1
      for (uint32_t i = bB; i < eB; ++i) fn(__ldg(&sorted[i]));
    }
  }
}

// monotone 22-bit quantisation of d2: bits(min(d2*S, 2^22-1) + 2^23) = 0x4B000000 + round(...)
__device__ __forceinline__ uint32_t quant_d2(float d2, float S) {
  return __float_as_uint(fminf(d2 * S, 4194303.f) + 8388608.f) - 0x4B000000u;
}

// number of points in the region scan_region(..., bound2 = inf) would visit (two table reads per row)
__device__ __forceinline__ uint32_t count_region(const KArgs& a, const QCell& q, int Rp, int R) {
  const GridDesc& g = a.g;
  const uint32_t* __restrict__ E = a.E;
  const int lo0 = max(q.c0 - R, 0), hi0 = min(q.c0 + R, g.dim[0] - 1);
  const int lo1 = max(q.c1 - R, 0), hi1 = min(q.c1 + R, g.dim[1] - 1);
  const int lo2 = max(q.c2 - R, 0), hi2 = min(q.c2 + R, g.dim[2] - 1);
  uint32_t total = 0;
#pragma unroll 1
  for (int e2 = lo2; e2 <= hi2; ++e2) {
#pragma unroll 1
    for (int e1 = lo1; e1 <= hi1; ++e1) {
      const uint32_t base = ((uint32_t)e2 * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
      if (Rp >= 0 && abs(e1 - q.c1) <= Rp && abs(e2 - q.c2) <= Rp) {
        const int ahi = min(hi0, q.c0 - Rp - 1), blo = max(lo0, q.c0 + Rp + 1);
        if (lo0 <= ahi) total += __ldg(&E[base + ahi + 1]) - __ldg(&E[base + lo0]);
        if (blo <= hi0) total += __ldg(&E[base + hi0 + 1]) - __ldg(&E[base + blo]);
      } else {
        total += __ldg(&E[base + hi0 + 1]) - __ldg(&E[base + lo0]);
      }
    }
  }
  return total;
}

// fractional position of Q inside its cell c along the x/y/z axis `ax` (selects instead of dynamic
// indexing, so the kernel arguments stay in the constant bank)
__device__ __forceinline__ float frac_in_cell(const GridDesc& g, int ax, const float4 Q, int c) {
  const double v = ax == 0 ? (double)Q.x : (ax == 1 ? (double)Q.y : (double)Q.z);
  const double o = ax == 0 ? g.origin[0] : (ax == 1 ? g.origin[1] : g.origin[2]);
  return fminf(fmaxf((float)((v - o) * g.inv_cell - (double)c), 0.f), 1.f);
}

// After a scan of the block of half-width R: is the list final, and if not how far to go next.
// `dk` = squared distance of the current k-th neighbour (only when found).  Returns true when done.
__device__ __forceinline__ bool coverage_step(const GridDesc& g, const QCell& q, int max_dim, bool found, float dk, int& Rp, int& R) {
  const bool whole = (q.c0 - R <= 0) && (q.c0 + R >= g.dim[0] - 1) && (q.c1 - R <= 0) && (q.c1 + R >= g.dim[1] - 1) &&
                     (q.c2 - R <= 0) && (q.c2 + R >= g.dim[2] - 1);
  if (whole) return true;
  int Rn;
  if (found) {
    if (dk <= cover2_of(R, g.cell)) return true;
    Rn = (int)fmin(ceil(sqrt((double)dk) * (1.0 + 2e-6) * g.inv_cell + 2e-6), 1.0e9);
    Rn = max(R + 1, Rn);
  } else {
    Rn = 2 * R;
  }
  Rp = R;
  R = min(Rn, max_dim);
  return false;
}

// Exact streaming path of one lane (always correct, slow): sorted insertion into the lane's column.
// Taken only when the histogram selection cannot narrow the candidate set to TQ_CAP entries.
__device__ __noinline__ int slow_knn(const KArgs& a, const QCell& q, float4 Q, u64* col, int k, int max_dim) {
  int cnt = 0, Rp = -1, R = a.sp.R0;
  u64 thr = PCP_KEY_INF;
  float bound = 3.4e38f;
  for (;;) {
    scan_region(a, q, Rp, R, bound, [&](const float4 P) {
      const float d2v = dist2_nofma(Q.x, Q.y, Q.z, P);
      const u64 key = ((u64)__float_as_uint(d2v) << 32) | (u64)__float_as_uint(P.w);
      if (key < thr) {
        int j = cnt < k ? cnt : k - 1;
        while (j > 0 && col[(j - 1) * 32] > key) {
          col[j * 32] = col[(j - 1) * 32];
          --j;
        }
        col[j * 32] = key;
        if (cnt < k) ++cnt;
        if (cnt >= k) thr = col[(k - 1) * 32];
      }
    });
    const bool found = cnt >= k;
    const float dk = __uint_as_float((uint32_t)(thr >> 32));
    if (found) bound = dk;
    if (coverage_step(a.g, q, max_dim, found, dk, Rp, R)) break;
  }
  return cnt;
}

__global__ void __launch_bounds__(TQ_WARPS * 32) knn_pca_kernel(KArgs a, int force_slow) {
  __shared__ TqSmem smem_all[TQ_WARPS];
  TqSmem& sm = smem_all[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  uint32_t* const hcol = &sm.hist[0][lane];  // this lane's histogram column, stride 32
  u64* const pcol = &sm.pend[0][lane];       // this lane's candidate column, stride 32
  const int64_t pos0 = ((int64_t)blockIdx.x * TQ_WARPS + (threadIdx.x >> 5)) * 32;
  if (pos0 >= a.n) return;  // whole warp
  const int64_t p = pos0 + lane;
  const GridDesc& g = a.g;
  const SearchParams& sp = a.sp;
  const int k = sp.k;
  const int max_dim = max(g.dim[0], max(g.dim[1], g.dim[2]));
  const int my_slot = p < a.n ? __ldg(&a.slot[p]) : -1;  // -1: tail lane or halo point (neighbour only)
  const float4 Q = p < a.n ? __ldg(&a.sorted[p]) : make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t cid = p < a.n ? __ldg(&a.keys[p]) : g.n_cells;
  const bool searching = my_slot >= 0 && cid < g.n_cells;  // non-finite queries have no neighbours

  QCell qc;
  qc.c0 = qc.c1 = qc.c2 = 0;
  qc.f0 = qc.f1 = qc.f2 = 0.f;
  if (searching) {
    qc.c0 = (int)(cid % (uint32_t)g.dim[0]);
    const uint32_t tq = cid / (uint32_t)g.dim[0];
    qc.c1 = (int)(tq % (uint32_t)g.dim[1]);
    qc.c2 = (int)(tq / (uint32_t)g.dim[1]);
    // fractional position inside the own cell (same double arithmetic as the key kernel)
    qc.f0 = frac_in_cell(g, g.ax[0], Q, qc.c0);
    qc.f1 = frac_in_cell(g, g.ax[1], Q, qc.c1);
    qc.f2 = frac_in_cell(g, g.ax[2], Q, qc.c2);
  }

  u64 best[TQ_CAP];
#pragma unroll
  for (int j = 0; j < TQ_CAP; ++j) best[j] = PCP_KEY_INF;
  int have = 0;              // valid entries of best[] (ascending)
  bool done = !searching;
  bool slow = searching && force_slow != 0;
  int Rp = -1, R = sp.R0;
  float bound = 3.4e38f;     // inclusive d2 bound once k neighbours are known
  u64 thr = PCP_KEY_INF;     // key of the current k-th neighbour

#pragma unroll 1
  for (;;) {
    if (!__any_sync(PCP_FULL, !done && !slow)) break;
    const bool run = !done && !slow;
    if (lane == 0) TQSTAT(1, 1);   // warp-level outer iterations
    if (run) TQSTAT(2, 1);         // lane-level outer iterations
    // ---- how many points does this scan see?  few: take them all; many and no bound yet: select
    // (the entries already held take part in the selection: a shell point may be nearer than they are)
    const int k_need = k, cap = TQ_CAP;
    uint32_t n_reg = 0;
    if (run && have < k) n_reg = count_region(a, qc, Rp, R);
    bool sel = run && have < k && n_reg + (uint32_t)have > (uint32_t)cap;
    const bool by_quant = sel;     // pass-2 predicate: quantised threshold (else: key < thr)
    uint32_t uT = 0xFFFFFFFFu;     // quantised threshold, exclusive
    float S = 0.f;
    bool full_range = Rp >= 0;
    if (sel) {
      // farthest a point of the block can be from the query; first scans start from a density guess
      const float ext = (float)(R + 1) * g.cell;
      const float corner2 = 3.f * ext * ext * 1.001f;
      float range2 = corner2;
      if (!full_range) {
        const float side = (float)(2 * R + 1) * g.cell;
        range2 = fminf(corner2, 0.96f * side * side * (float)k / (float)(n_reg + (uint32_t)have));
      }
      S = 4194303.f / range2;
    }
    uint32_t ulo = 0, span = 1u << TQ_QBITS;
    int shift = TQ_QBITS - 6, below = 0;
#pragma unroll 1
    while (__any_sync(PCP_FULL, sel)) {
      if (lane == 0) TQSTAT(3, 1);  // warp-level histogram passes
      if (sel) {
        TQSTAT(4, 1);               // lane-level histogram passes
#pragma unroll
        for (int b = 0; b < TQ_BINS; ++b) hcol[b * 32] = 0;
        if (have > 0) {
#pragma unroll
          for (int j = 0; j < TQ_CAP; ++j) {
            if (j < have) {
              const uint32_t t = quant_d2(__uint_as_float((uint32_t)(best[j] >> 32)), S) - ulo;
              if (t < span) hcol[(t >> shift) * 32] += 1;
            }
          }
        }
        scan_region(a, qc, Rp, R, 3.4e38f, [&](const float4 P) {
          const float d2v = dist2_nofma(Q.x, Q.y, Q.z, P);
          const uint32_t t = quant_d2(d2v, S) - ulo;
          if (t < span) hcol[(t >> shift) * 32] += 1;
          TQSTAT(5, 1);             // candidates seen by histogram passes
        });
        int cum = below, bstar = -1, cum_star = 0, cum_prev = 0;
#pragma unroll 8
        for (int b = 0; b < TQ_BINS; ++b) {
          const int c = (int)hcol[b * 32];
          if (bstar < 0 && cum + c >= k_need) {
            bstar = b;
            cum_prev = cum;
            cum_star = cum + c;
          }
          cum += c;
        }
        if (bstar < 0) {
          // cannot happen (n_reg > cap >= k_need candidates were counted) — take everything
          uT = 0xFFFFFFFFu;
          sel = false;
          slow = true;
        } else if (!full_range && ulo == 0 && shift == TQ_QBITS - 6 && bstar == TQ_BINS - 1) {
          // the k-th lies beyond the guessed range (the last bin also holds the clamped values): redo on
          // the full range, where nothing clamps
          const float ext = (float)(R + 1) * g.cell;
          S = 4194303.f / (3.f * ext * ext * 1.001f);
          full_range = true;
          TQSTAT(6, 1);             // range guess too small
        } else if (cum_star <= cap) {
          uT = ulo + ((uint32_t)(bstar + 1) << shift);
          sel = false;
        } else if (shift == 0) {
          slow = true;  // > cap candidates share one quantised value: exact ties, insertion path
          sel = false;
        } else {
          TQSTAT(7, 1);             // threshold bin too full: narrowing round
          below = cum_prev;
          ulo += (uint32_t)bstar << shift;
          span = 1u << shift;
          shift = max(shift - 6, 0);
        }
      }
    }
    // ---- pass 2: collect
    const bool col = run && !slow;
    int cnt = 0;
    if (col) {
      // held entries stay while they can still be among the k nearest (a prefix: best[] is ascending)
      const int keep = have < k ? have : k;
#pragma unroll
      for (int j = 0; j < TQ_CAP; ++j) {
        if (j < keep && (!by_quant || quant_d2(__uint_as_float((uint32_t)(best[j] >> 32)), S) < uT)) {
          pcol[j * 32] = best[j];
          cnt = j + 1;
        }
      }
      float cb = bound;
      if (by_quant && uT != 0xFFFFFFFFu) cb = fminf(cb, ((float)uT + 1.f) / S * 1.00001f);
      scan_region(a, qc, Rp, R, cb, [&](const float4 P) {
        const float d2v = dist2_nofma(Q.x, Q.y, Q.z, P);
        const u64 key = ((u64)__float_as_uint(d2v) << 32) | (u64)__float_as_uint(P.w);
        const bool take = by_quant ? quant_d2(d2v, S) < uT : key < thr;
        if (take) {
          if (cnt < TQ_CAP) pcol[cnt * 32] = key;
          ++cnt;
        }
        TQSTAT(8, 1);               // candidates seen by collect passes
      });
      TQSTAT(9, cnt);               // entries handed to the sorting network
      if (cnt > TQ_CAP) slow = true;  // more than cap points inside the k-th distance of the inner block
    }
    if (col && !slow) {
#pragma unroll
      for (int j = 0; j < TQ_CAP; ++j) best[j] = j < cnt ? pcol[j * 32] : PCP_KEY_INF;
      sort32_regs(best);
      have = cnt;
      u64 kth = PCP_KEY_INF;
#pragma unroll
      for (int j = 0; j < TQ_CAP; ++j)
        if (j == k - 1) kth = best[j];
      const bool found = kth != PCP_KEY_INF;
      const float dk = __uint_as_float((uint32_t)(kth >> 32));
      if (found) {
        thr = kth;
        bound = dk;
      }
      done = coverage_step(g, qc, max_dim, found, dk, Rp, R);
    }
  }
  // ---- lanes the selection could not serve: exact insertion path, then reload
  if (searching) TQSTAT(0, 1);     // queries
  if (slow) {
    TQSTAT(10, 1);                 // lanes on the exact insertion path
    const int cnt = slow_knn(a, qc, Q, pcol, k, max_dim);
#pragma unroll
    for (int j = 0; j < TQ_CAP; ++j) best[j] = j < cnt ? pcol[j * 32] : PCP_KEY_INF;
    have = cnt;
  }
  const int n_nb = have < k ? have : k;

#pragma unroll
  for (int j = 0; j < TQ_CAP; ++j) pcol[j * 32] = best[j];
  knn_finish(a, reinterpret_cast<uint32_t*>(pcol), reinterpret_cast<uint32_t*>(pcol) + 1, 64, n_nb, 1, Q, my_slot, sm.rec, lane);
}

}  // namespace

#ifdef PCP_STATS
extern "C" __attribute__((visibility("default"))) int pcp_debug_tq_stats(unsigned long long* out) {
  unsigned long long z[16] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_tq_stats, sizeof(z));
  cudaMemcpyToSymbol(g_tq_stats, z, sizeof(z));
  return 0;
}
#endif

int pcp_knn_launch(pcp_ctx* ctx, const SearchParams& sp) {
  KArgs a;
  a.sorted = (const float4*)ctx->sorted_pts.p;
  a.keys = (const uint32_t*)ctx->keys_sorted.p;
  a.slot = (const int32_t*)ctx->sorted_slot.p;
  a.E = (const uint32_t*)ctx->cell_end.p;
  a.inv = (const uint32_t*)ctx->inv_pos.p;
  a.gid_of_ord = ctx->d_gid ? (const int32_t*)ctx->gid_sorted.p : nullptr;
  a.g = ctx->grid;
  a.n = ctx->n;
  a.sp = sp;
  // test hook: PCP_KNN_FORCE_EXACT=1 sends every lane through the insertion path (same results, slow)
  const char* dbg = getenv("PCP_KNN_FORCE_EXACT");
  const int force_slow = (dbg && dbg[0] == '1') ? 1 : 0;
  const int64_t tiles = (ctx->n + 31) / 32;
  const int blocks = (int)((tiles + TQ_WARPS - 1) / TQ_WARPS);
  pcp_tic(ctx, PCP_T_SEARCH_PCA);
  knn_pca_kernel<<<blocks, TQ_WARPS * 32, 0, ctx->stream>>>(a, force_slow);
  PCP_LAUNCH_CHECK("knn_pca_kernel");
  pcp_toc(ctx, PCP_T_SEARCH_PCA);
  return PCP_OK;
}
