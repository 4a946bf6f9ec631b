// knn_tile6.cu — k-NN / fixed-radius search + PCA fast path: one warp per tile of 32 consecutive cell-sorted queries, one
// lane per query.
//
// Replaces, per point, nearestKSearch (src/smoothing.cpp:239-246) or radiusSearch (src/pcaAnalysis.cpp:25) +
// computeCovarianceMatrixNormalized + SelfAdjointEigenSolver + the structure features (src/pcaAnalysis.cpp:44-87, :29-31).
// The stages below describe the k-NN form; the radius form (T6_BIN_RADIUS) drops pass 1 and the scan, see the bin modes.
//
// The kernel is bound by latency and by what a warp keeps resident (shared memory, registers), not by arithmetic
// (ncu, profiles/r02a_*: 9 % of a warp's cycles issue), so every stage is built to hold little and wait little:
//   staging   the union of the tile's stencils (contiguous runs of the cell-sorted array) is copied to shared memory
//             128 candidates at a time by the TMA unit: one 1-D bulk copy (cp.async.bulk.shared.global, SASS UBLKCP) per
//             piece of a run, completion counted in bytes by an mbarrier, two buffers — the copy of chunk c + 1 is in
//             flight while the warp works on chunk c and costs it neither registers nor issue slots; the chunk is then
//             rewritten in place in a LOCAL frame (p' = p - centre) together with |p'|^2 (and, in pass 2 only, the
//             candidates' sorted positions).  -DT6_TMA=0 keeps the earlier register staging (four LDG.128 in flight
//             per lane) for A/B runs;
//   pass 1    bin coordinate Y = S |p' - q'|^2 + off = S |p'|^2 + (S |q'|^2 + off) - 2 S q'.p': one broadcast LDS.128 +
//             four FFMA with per-lane constants (the last one saturating to [0, 1]: FFMA.SAT); the bin is
//             round(NB sqrt~(Y)) — one MUFU.SQRT and one more FFMA that adds 2^23 so that the float's low bits ARE the
//             bin (bins uniform in distance) — or, in the geometric mode, the top bits of the float Y itself (shift +
//             clamp, no MUFU); one IMAD forms the address and one shared reduction counts into the lane's own
//             histogram column (CAP rows: CAP - 1 bins inside the cover radius + 1 beyond);
//   scan      the bin b* holding the k-th smallest (one refinement pass for an over-full bin);
//   pass 2    the same four FFMA + ONE float compare Y < edge(b*) (geometric bins: exact, the bins are read off Y's own
//             bits; sqrt bins re-evaluate the bin: MUFU + FFMA + compare); members append their SORTED POSITION to the
//             lane's column (the histogram rows are dead by then: the column is the kernel's only per-lane storage);
//   resolve   every lane reads its <= CAP members from the sorted array, computes the reference distance (FLANN
//             L2_Simple arithmetic) and forms 32-BIT keys prefix(d2) | slot; a register sorting network orders them
//             (one min/max pair per compare-exchange, CAP registers) and the column is permuted into that order;
//   check     the inexact bin coordinate is within eps_Y of the exact one (bound below), so a point that was NOT
//             collected (Y_computed >= y*) has d2 >= t_safe := (y* - eps_Y - off) / S; the lane's result stands only if
//             its k-th distance lies below t_safe, otherwise it goes to the work list like any lane the tile cannot
//             serve;
//   finish    knn_finish_pos (knn_finish.cuh) walks the column, recomputes the full keys bits(d2) << 32 | ord, repairs
//             the order inside groups of equal prefix (exact ties), accumulates the moment in that order — the
//             (d2, index) order of the reference, so the moment is bitwise the reference's — then eigen-solve,
//             features and the 68-byte record.
// Results are identical to the round-1 kernels and to the CPU restatement the tests compare with.
//
// Error bound of the bin coordinate.  Floats p', q' (local coordinates, each rounded once: relative u = 2^-24),
// pp = |p'|^2 and qq = |q'|^2 by FFMA chains (<= 3 u relative), A_i = -2 S q'_i (u), K = fl(S qq + off) (u), four
// chained FFMA (u of the partial sum each, |partial| <= S (|p'| + |q'|)^2 + |off|), and the local frame itself
// (| |p'-q'|^2 - |p-q|^2 | <= 2 u (|p'| + |q'|)^2):  |Y_computed - Y_exact| <= u (10 S (|p'| + |q'|)^2 + 5 |off|).
// A candidate with |p'| > |q'| + 2 c (c = the lane's cover radius) is farther than 2 c from the query and cannot
// matter (the k-th distance is checked to be < c); for the others |p'| + |q'| <= 2 |q'| + 2 c.  The kernel uses
// eps_Y = 1.01 u (12 S (2 |q'| + 2 c)^2 + 12 |off| + 4), which also covers the float evaluation of t_safe itself.
#include <cub/block/block_radix_sort.cuh>

#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "eigen3.cuh"
#include "pcp_internal.cuh"
#include "search_common.cuh"
#include "knn_finish.cuh"
#include "tile_common.cuh"

#ifdef PCP_STATS
__device__ unsigned long long g_tile6_stats[16];
#define T6STAT(i, v) atomicAdd(&g_tile6_stats[i], (unsigned long long)(v))
#else
#define T6STAT(i, v) do { } while (0)
#endif

namespace {

#ifndef T6_NWARPS
#define T6_NWARPS 4
#endif
constexpr int T6_WARPS = T6_NWARPS;     // warps per CTA
#ifndef T6_CHUNK
#define T6_CHUNK 128
#endif
constexpr int T6_CH = T6_CHUNK;  // staged candidates per chunk (a multiple of 32)
constexpr int T6_NL = T6_CH / 32;  // loads a lane has in flight while a chunk is staged
// CAP (template parameter, 32 or 64) = members a lane can hold = rows of its column = width of the sorting network; the
// histogram uses the same rows: NB = CAP - 1 bins inside the cover radius, bin NB = at or beyond it.
constexpr float T6_MAGIC = 8388608.f;        // 2^23: float_as_uint(x + 2^23) - 0x4B000000 = round-to-nearest-even(x), 0 <= x < 2^22
constexpr uint32_t T6_MAGIC_BITS = 0x4B000000u;
constexpr int T6_REC_WORDS = 17;
#ifndef T6_SYNC
#define T6_SYNC 1     // CTA barriers around the straight-line sorting network (instruction-cache alignment of the four warps)
#endif
#ifndef T6_TMA
#define T6_TMA 1      // stage the candidate chunks with 1-D TMA bulk copies + mbarrier, double-buffered (0: LDG.128 -> STS.128 through registers)
#endif
// Bin modes of the selection passes (template parameter BIN; chosen per k in the launcher, PCP_T6_BIN forces one):
//   T6_BIN_SQRT   both passes evaluate round(NB sqrt~(Y)) (MUFU.SQRT): bins uniform in distance (the round-2 form, kept for A/B)
//   T6_BIN_LOG    bins = top bits of the float Y itself (geometric): integer ops, no MUFU; pass 2 is the compare Y < edge
//   T6_BIN_RADIUS fixed-radius search (radiusSearch, src/pcaAnalysis.cpp:25): the threshold is known, so pass 1 and the scan
//                 disappear; pass 2 collects the superset Y < 1 + eps (Y = d2 / r2 in the tile-local frame), the resolve step
//                 keeps the members whose REFERENCE distance satisfies d2 < r2 (strict, FLANN's arithmetic) — exact by
//                 construction, no proof needed; lanes with more hits than the column holds go to the work list
constexpr int T6_BIN_SQRT = 0, T6_BIN_LOG = 1, T6_BIN_RADIUS = 2;
// Measured and dropped: sqrt bins in pass 1 with the compare Y < y* in pass 2, y* found by bisection with the pass-1
// instructions (needs a guard against a non-monotone MUFU and was the fastest form for no k); no shared reduction for
// candidates beyond the cover radius — `if (Y < 1) red...` — 1.16 -> 1.28 ms: the reductions are cheap, the divergence
// around them is not.
#ifndef T6_ORDER
#define T6_ORDER 1    // warps of a CTA take tiles of similar candidate counts (tile_order_kernel): they reach the CTA barriers together
#endif
#ifndef T6_MINB
#define T6_MINB 6     // CTAs per SM the 32-member kernel is compiled for (register budget 65536 / (128 * T6_MINB) = 80, no spills; shared memory allows 6).  With TMA staging 6 beats 5 (1.26 vs 1.28 ms); with register staging 5 was best
#endif

template <int CAP>
struct __align__(16) T6Smem {
  float4 cand[T6_TMA ? 2 : 1][T6_CH];    // staged candidates of the current chunk (and, with TMA staging, the next one in flight): local x, y, z and |p'|^2
  uint32_t cpos[T6_CH];                  // their sorted positions (pass 2)
  // One 32-bit column per lane, three uses: pass 1 counts into plane[bin][lane]; pass 2 stores the sorted position of
  // member j in plane[j][lane]; after the sort the same rows hold the positions in neighbour order.  The record
  // staging (32 x 17 floats) reuses the space at the end.
  union U {
    uint32_t plane[CAP][32];
    float rec[32][T6_REC_WORDS];
  } u;
  uint32_t run_beg[TA_MAXROWS];
  uint32_t run_off[TA_MAXROWS + 1];
#if T6_TMA
  unsigned long long bar[2];             // one mbarrier per chunk buffer (transaction bytes of the bulk copies)
#endif
};

__device__ __forceinline__ float fma_sat(float a, float b, float c) {
  float r;
  asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

// Shared-memory accesses of the two hot loops, with explicit 32-bit shared addresses and immediate offsets: one
// address register serves the four candidates of a step (and a second one their positions).
template <int OFF>
__device__ __forceinline__ float4 lds_f4(const uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4+%5];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr), "n"(OFF));
  return v;
}
template <int OFF>
__device__ __forceinline__ uint32_t lds_u32(const uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFF));
  return v;
}
__device__ __forceinline__ void sts_u32(const uint32_t addr, const uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
// histogram update: a shared reduction into the lane's own column (no conflicts, no read-back); lanes that are not
// counting add 0
__device__ __forceinline__ void red_add_u32(const uint32_t addr, const uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

#if T6_TMA
// ---- 1-D TMA bulk copies global -> shared, completion on an mbarrier (PTX ISA: cp.async.bulk, mbarrier) ----
__device__ __forceinline__ void mbar_init(const uint32_t bar, const uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// one arrival (the barrier's only one) + the bytes the phase waits for
__device__ __forceinline__ void mbar_arrive_expect_tx(const uint32_t bar, const uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// src and dst 16-byte aligned, bytes a multiple of 16; the copy's bytes are counted on `bar`
__device__ __forceinline__ void bulk_g2s(const uint32_t dst, const void* src, const uint32_t bytes, const uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// waits for the phase of the given parity; a wait that does not end (a lost copy) traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(const uint32_t bar, const uint32_t parity) {
  uint32_t done = 0;
#pragma unroll 1
  for (uint32_t spins = 0; !done; ++spins) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spins > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#endif

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));  // one MUFU.SQRT; relative error below 2^-22 (PTX ISA), denormals -> 0
  return r;
}

// the bin coordinate of one staged candidate for this lane: Y = S d2 + off clamped to [0, 1] (four chained FFMA, the last
// one saturating)
template <bool SAT = true>
__device__ __forceinline__ float bin_y(const float4 P, const float S, const float K, const float A0, const float A1, const float A2) {
  float y = __fmaf_rn(P.w, S, K);
  y = __fmaf_rn(P.x, A0, y);
  y = __fmaf_rn(P.y, A1, y);
  return SAT ? fma_sat(P.z, A2, y) : __fmaf_rn(P.z, A2, y);
}
// Geometric bins read off the float itself: code(Y) = bits(Y) >> SH keeps the exponent and the top 23 - SH mantissa bits
// (SH = 21: four bins per octave of Y, a ratio of 2^(1/8) = 9 % in DISTANCE, for 32-row columns; SH = 20: eight per octave
// for 64-row columns); Y <= 1 has code <= C1 = code(1.0), the bin is code - C0 clamped at 0, C0 = C1 - NB: bin NB <=> Y = 1
// (at or beyond the cover radius), bin 0 = everything below 2^(-NB / bins per octave).  Monotone in Y, exact integer
// arithmetic: "bin(Y) <= b" is the float compare Y < y_edge(b), which is all pass 2 needs (no MUFU in either pass).
template <int CAP>
struct LogBin {
  static constexpr int SH = CAP == 32 ? 21 : 20;
  static constexpr uint32_t C1 = 127u << (23 - SH);
  static constexpr uint32_t C0 = C1 - (uint32_t)(CAP - 1);
  // upper edge (exclusive) of bin b, 0 <= b < NB
  static __device__ __forceinline__ float edge(const int b) { return __uint_as_float((C0 + (uint32_t)b + 1u) << SH); }
  // byte offset of plane[bin][0] relative to plane[-C0][0] (the caller folds the constant into its base address)
  static __device__ __forceinline__ uint32_t row(const float y) { return max(__float_as_uint(y) >> SH, C0) * 128u; }
};
// the bin as the float 2^23 + round(NB sqrt(Y)).  Bins are uniform in DISTANCE, not in squared distance: on a surface the
// members per bin grow with the bin index instead of being constant, so a k-th neighbour far inside the cover radius
// (dense scan lines of a VLP-16 ring: d_k ~ cover / 20) still meets a bin narrow enough for the lane's column.
template <int NB>
__device__ __forceinline__ float bin_z(const float4 P, const float S, const float K, const float A0, const float A1, const float A2) {
  return __fmaf_rn(sqrt_approx(bin_y(P, S, K, A0, A1, A2)), (float)NB, T6_MAGIC);
}

// The sorted positions of the candidates [cb, cb + cnt) of the tile's virtual candidate sequence (the stencil rows' runs
// laid end to end): lane l takes the candidates cb + l + 32 i, i < T6_NL, and finds their stencil rows by counting the
// row ends at or below each index (the rows are few).
template <int CAP>
__device__ __forceinline__ void chunk_positions(const T6Smem<CAP>& sm, const int nrows, const uint32_t cb, const int lane, uint32_t (&pos)[T6_NL]) {
  uint32_t v[T6_NL];
  int row[T6_NL];
#pragma unroll
  for (int i = 0; i < T6_NL; ++i) {
    v[i] = cb + (uint32_t)(lane + 32 * i);
    row[i] = 0;
  }
#pragma unroll 4
  for (int r = 1; r < nrows; ++r) {
    const uint32_t off = sm.run_off[r];  // broadcast
#pragma unroll
    for (int i = 0; i < T6_NL; ++i) row[i] += (off <= v[i]) ? 1 : 0;
  }
#pragma unroll
  for (int i = 0; i < T6_NL; ++i) pos[i] = sm.run_beg[row[i]] + (v[i] - sm.run_off[row[i]]);
}

#if T6_TMA
// Starts the copy of candidates [cb, cb + cnt) into chunk buffer `buf`: the chunk is a few contiguous pieces of the
// cell-sorted array (one per stencil row that overlaps it); lane r issues the piece of row r as ONE bulk copy.  The
// buffer's mbarrier completes when all cnt * 16 bytes have landed.  The proxy fence orders the warp's earlier
// (generic) reads and writes of the buffer before the copy unit's writes.
template <int CAP>
__device__ __forceinline__ void chunk_issue(const float4* __restrict__ sorted, T6Smem<CAP>& sm, const int nrows, const uint32_t cb, const int cnt,
                                            const int lane, const int buf) {
  const uint32_t bar = smem_addr(&sm.bar[buf]);
  fence_proxy_async();
  if (lane == 0) mbar_arrive_expect_tx(bar, (uint32_t)cnt * 16u);
  __syncwarp();
#pragma unroll 1
  for (int r = lane; r < nrows; r += 32) {
    const uint32_t o0 = sm.run_off[r], o1 = sm.run_off[r + 1];
    const uint32_t s = max(o0, cb), e = min(o1, cb + (uint32_t)cnt);
    if (s < e) bulk_g2s(smem_addr(&sm.cand[buf][s - cb]), sorted + sm.run_beg[r] + (s - o0), (e - s) * 16u, bar);
  }
}
// Waits for chunk buffer `buf` and rewrites it in place in the local frame: x - cx, y - cy, z - cz and |p'|^2.
// WITH_POS (pass 2): the candidates' sorted positions go to sm.cpos.  CONV = false: the chunk was converted before
// (a tile whose candidates fit one chunk keeps it from pass 1), only the positions are wanted.
template <int CAP, bool CONV, bool WITH_POS>
__device__ __forceinline__ void chunk_land(T6Smem<CAP>& sm, const int nrows, const uint32_t cb, const int cnt, const int lane, const int buf,
                                           uint32_t& parity, const float cx, const float cy, const float cz) {
  if (CONV) {
    mbar_wait(smem_addr(&sm.bar[buf]), (parity >> buf) & 1u);
    parity ^= 1u << buf;
    float4* c = sm.cand[buf];
#pragma unroll
    for (int i = 0; i < T6_NL; ++i) {
      if (lane + 32 * i < cnt) {
        const float4 P = c[lane + 32 * i];
        const float x = __fsub_rn(P.x, cx), y = __fsub_rn(P.y, cy), z = __fsub_rn(P.z, cz);
        c[lane + 32 * i] = make_float4(x, y, z, __fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x))));
      }
    }
  }
  if (WITH_POS) {
    uint32_t pos[T6_NL];
    chunk_positions(sm, nrows, cb, lane, pos);
#pragma unroll
    for (int i = 0; i < T6_NL; ++i)
      if (lane + 32 * i < cnt) sm.cpos[lane + 32 * i] = pos[i];
  }
  __syncwarp();
}
#else
// Register staging: copies candidates [cb, cb + cnt) into sm.cand[0] / sm.cpos, coordinates relative to (cx, cy, cz):
// every lane issues its T6_NL loads back to back — one round trip to L2 per chunk instead of one per row piece —
// and converts and stores when they land.
template <int CAP>
__device__ __forceinline__ void stage6(const float4* __restrict__ sorted, T6Smem<CAP>& sm, const int nrows, const uint32_t cb, const int cnt,
                                       const int lane, const float cx, const float cy, const float cz) {
  uint32_t pos[T6_NL];
  float4 P[T6_NL];
  chunk_positions(sm, nrows, cb, lane, pos);
#pragma unroll
  for (int i = 0; i < T6_NL; ++i)
    if (lane + 32 * i < cnt) P[i] = __ldg(&sorted[pos[i]]);
#pragma unroll
  for (int i = 0; i < T6_NL; ++i) {
    if (lane + 32 * i < cnt) {
      const float x = __fsub_rn(P[i].x, cx), y = __fsub_rn(P[i].y, cy), z = __fsub_rn(P[i].z, cz);
      sm.cand[0][lane + 32 * i] = make_float4(x, y, z, __fmaf_rn(z, z, __fmaf_rn(y, y, __fmul_rn(x, x))));
      sm.cpos[lane + 32 * i] = pos[i];
    }
  }
  __syncwarp();
}
#endif

// The four warps of a CTA meet at two barriers (below), so a CTA is as slow as its slowest tile: 17 % of the kernel's
// warp time was spent waiting there (ncu, profiles/r02f_search_ncu_tma_sqrt_bins.txt) because tiles next to each other in
// cell order differ by +-30 % in candidates.  This kernel estimates every tile's candidate count (first and last
// query's cells, two table reads per stencil row) and sorts each window of 256 consecutive tiles by it: warp w of CTA b
// then serves tile perm[4 b + w], four tiles of nearly equal work, still from one neighbourhood of the cloud.
#ifndef T6_WIN_ITEMS
#define T6_WIN_ITEMS 1
#endif
constexpr int T6_WIN_THREADS = 256;
constexpr int T6_WIN = T6_WIN_THREADS * T6_WIN_ITEMS;  // tiles per window
__global__ void __launch_bounds__(T6_WIN_THREADS) tile_order_kernel(KArgs a, int R, int64_t n_tiles, uint32_t* __restrict__ perm) {
  using Sort = cub::BlockRadixSort<uint32_t, T6_WIN_THREADS, T6_WIN_ITEMS, uint32_t>;
  __shared__ typename Sort::TempStorage tmp;
  const GridDesc& g = a.g;
  uint32_t w[T6_WIN_ITEMS], id[T6_WIN_ITEMS];
#pragma unroll
  for (int it = 0; it < T6_WIN_ITEMS; ++it) {
    const int64_t t = (int64_t)blockIdx.x * T6_WIN + threadIdx.x * T6_WIN_ITEMS + it;
    w[it] = 0xFFFFFFFFu;
    id[it] = (uint32_t)t;
    if (t < n_tiles) {
      const int64_t p0 = t * 32, p1 = min(p0 + 31, a.n - 1);
      const uint32_t cid0 = __ldg(&a.keys[p0]), cid1 = __ldg(&a.keys[p1]);
      w[it] = 0u;
      if (cid0 < g.n_cells) {
        const uint32_t row0 = cid0 / (uint32_t)g.dim[0];
        const int c0a = (int)(cid0 - row0 * (uint32_t)g.dim[0]);
        const int c0b = (cid1 < g.n_cells && cid1 / (uint32_t)g.dim[0] == row0) ? (int)(cid1 - row0 * (uint32_t)g.dim[0]) : g.dim[0] - 1;
        const int c1 = (int)(row0 % (uint32_t)g.dim[1]);
        int c2 = (int)(row0 / (uint32_t)g.dim[1]);
        const int zoff = pcp_unfold_z(g, c2);
        const int xlo = max(c0a - R, 0), xhi = min(c0b + R, g.dim[0] - 1);
        uint32_t total = 0;
        for (int e2 = max(c2 - R, 0); e2 <= min(c2 + R, g.dim[2] - 1); ++e2)
          for (int e1 = max(c1 - R, 0); e1 <= min(c1 + R, g.dim[1] - 1); ++e1) {
            const uint32_t base = ((uint32_t)(zoff + e2) * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
            total += __ldg(&a.E[base + xhi + 1]) - __ldg(&a.E[base + xlo]);
          }
        w[it] = total;
      }
    }
  }
  Sort(tmp).Sort(w, id);  // tiles past the end sort last: the first (n_tiles - window start) ranks are real tiles
#pragma unroll
  for (int it = 0; it < T6_WIN_ITEMS; ++it) {
    const int64_t t = (int64_t)blockIdx.x * T6_WIN + threadIdx.x * T6_WIN_ITEMS + it;
    if (t < n_tiles) perm[t] = id[it];
  }
}

// LEAN: records only (query-centred; the k-th distance may be asked for): see knn_finish.cuh
template <int CAP, bool LEAN, int BIN>
__global__ void __launch_bounds__(T6_WARPS * 32, CAP == 32 ? (T6_MINB * 4) / T6_WARPS : 16 / T6_WARPS)
    knn_tile6_kernel(KArgs a, int R0, int max_groups, uint32_t* __restrict__ work_count, uint32_t* __restrict__ work_list,
                     const uint32_t* __restrict__ perm, int64_t n_tiles, int refine_min) {
  constexpr int NB = CAP - 1;
  constexpr int CAPM = CAP;  // members a served lane may have
  constexpr bool RAD = BIN == T6_BIN_RADIUS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T6Smem<CAP>& sm = reinterpret_cast<T6Smem<CAP>*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int64_t wid = (int64_t)blockIdx.x * T6_WARPS + (threadIdx.x >> 5);
  const bool live_warp = wid < n_tiles;  // a tail warp stays (inactive) for the CTA barriers below
  const int64_t pos0 = (live_warp ? (perm ? (int64_t)__ldg(&perm[wid]) : wid) : n_tiles) * 32;
  const int64_t p = pos0 + lane;
  const GridDesc& g = a.g;
  const SearchParams& sp = a.sp;
  const float4* __restrict__ sorted = a.sorted;
  const int k = sp.k;
  const bool in_range = live_warp && p < a.n;
  const int my_slot = in_range ? __ldg(&a.slot[p]) : -1;  // -1: tail lane or halo point (neighbour only)
  const float4 Q = in_range ? __ldg(&sorted[p]) : make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t cid = in_range ? __ldg(&a.keys[p]) : g.n_cells;
  const bool active = my_slot >= 0;
  const bool finite = active && cid < g.n_cells;
  uint32_t* const hcol = &sm.u.plane[0][lane];
  const uint32_t hcol_a = smem_addr(hcol), cand_a = smem_addr(sm.cand), cpos_a = smem_addr(sm.cpos);
#if T6_TMA
  uint32_t parity = 0;  // bit b: the phase of chunk buffer b's mbarrier the next wait is for
  if (lane == 0) {
    mbar_init(smem_addr(&sm.bar[0]), 1);
    mbar_init(smem_addr(&sm.bar[1]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
#endif
  // LogBin::row(Y) + hbase, or (bits of 2^23 + bin) * 128 + hbase (modulo 2^32) = address of plane[bin][lane]
  const uint32_t hbase = BIN == T6_BIN_LOG ? hcol_a - LogBin<CAP>::C0 * 128u : hcol_a - T6_MAGIC_BITS * 128u;
  uint32_t wp = hcol_a;      // next free entry of this lane's member column
  const uint32_t wend = hcol_a + (uint32_t)CAP * 128u;
  const uint32_t my_row = cid / (uint32_t)g.dim[0];
  const int c0 = (int)(cid - my_row * (uint32_t)g.dim[0]);
  bool pending = finite;     // not yet served by a row group
  bool ok = false;           // the lane's k nearest are in its column
  bool far = false;          // the k-th neighbour lies beyond the block of half-width 1 (told to the work-list kernel)
  float t_safe = 0.f;        // the k-th distance must lie below this for the selection to be proven

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
    const int ldr = __ffs(lead) - 1;
    const uint32_t row0 = __shfl_sync(PCP_FULL, my_row, ldr);
    const bool in_grp = pending && my_row == row0;
    pending = pending && !in_grp;
    const int c1 = (int)(row0 % (uint32_t)g.dim[1]);
    int c2 = (int)(row0 / (uint32_t)g.dim[1]);
    const int zoff = pcp_unfold_z(g, c2);  // frame batches: the frame's first z row; c2 is now frame-local
    const int c0min = __reduce_min_sync(PCP_FULL, in_grp ? c0 : 0x7fffffff);
    const int c0max = __reduce_max_sync(PCP_FULL, in_grp ? c0 : -1);
    const int R = R0;
    int nrows = 0, bstar = -1, cum_star = 0;
    bool whole = false;
    float cover_cells;
    const uint32_t total = tile_setup(a, sm.run_beg, sm.run_off, lane, R, c0min, c0max, c1, c2, zoff, c0, Q, nrows, whole, cover_cells);
    // local frame: centred between the group's first and last query (the queries of a row are ordered along the row, so
    // no query is farther from the centre than half the group's extent; any float vector would do, the error bound below
    // only wants |q'| small)
    const int lst = 31 - __clz(__ballot_sync(PCP_FULL, in_grp));
    const float cx = 0.5f * (__shfl_sync(PCP_FULL, Q.x, ldr) + __shfl_sync(PCP_FULL, Q.x, lst));
    const float cy = 0.5f * (__shfl_sync(PCP_FULL, Q.y, ldr) + __shfl_sync(PCP_FULL, Q.y, lst));
    const float cz = 0.5f * (__shfl_sync(PCP_FULL, Q.z, ldr) + __shfl_sync(PCP_FULL, Q.z, lst));
    const float qx = __fsub_rn(Q.x, cx), qy = __fsub_rn(Q.y, cy), qz = __fsub_rn(Q.z, cz);
    const float qq = __fmaf_rn(qz, qz, __fmaf_rn(qy, qy, __fmul_rn(qx, qx)));
    // cover radius in metres; a tile whose block reaches the grid boundary on every side has no finite cover
    // radius: every candidate lies within the block's diagonal
    float cover_m;
    {
      const float ex = (float)(min(c0max + R, g.dim[0] - 1) - max(c0min - R, 0) + 1), ey = (float)(min(c1 + R, g.dim[1] - 1) - max(c1 - R, 0) + 1);
      const float ez = (float)(min(c2 + R, g.dim[2] - 1) - max(c2 - R, 0) + 1);
      const float diag_cells = sqrtf(ex * ex + ey * ey + ez * ez) * 1.0001f;
      cover_m = fminf(cover_cells, diag_cells) * g.cell;
    }
    if (RAD) cover_m = sqrtf(sp.r2);  // the stencil covers the radius by construction (stencil_for_radius); Y = d2 / r2
    float S = RAD ? 1.f / sp.r2 : 1.f / (cover_m * cover_m * (1.f - 1e-6f));  // Y = S d2 + off reaches 1 just inside the cover radius
    float off = 0.f;
    float A0 = -2.f * S * qx, A1 = -2.f * S * qy, A2 = -2.f * S * qz, K = __fmaf_rn(S, qq, off);
    // ---- pass 1: histogram of the bin coordinate over the staged candidates; a lane whose threshold bin holds
    // too many candidates for its column (d_k << cover: dense scan lines next to a sparse surrounding, e.g. the
    // near field of a VLP-16 frame) gets ONE refinement pass with the window narrowed to that bin (candidates
    // below the window saturate to bin 0 and are all collected, those above it to bin NB)
    if constexpr (!RAD) {
    bool hist_on = in_grp, refined = false;
#pragma unroll 1
    for (;;) {
      if (hist_on) {
#pragma unroll
        for (int b = 0; b < CAP; ++b) hcol[b * 32] = 0;
      }
      const uint32_t inc = hist_on ? 1u : 0u;
      __syncwarp();
#if T6_TMA
      if (total > 0) chunk_issue(sorted, sm, nrows, 0u, (int)min((uint32_t)T6_CH, total), lane, 0);
#endif
#pragma unroll 1
      for (uint32_t cb = 0; cb < total; cb += T6_CH) {
        const int n_ch = (int)min((uint32_t)T6_CH, total - cb);
#if T6_TMA
        const int buf = (int)((cb / T6_CH) & 1u);
        if (cb + T6_CH < total) chunk_issue(sorted, sm, nrows, cb + T6_CH, (int)min((uint32_t)T6_CH, total - cb - T6_CH), lane, buf ^ 1);
        chunk_land<CAP, true, false>(sm, nrows, cb, n_ch, lane, buf, parity, cx, cy, cz);
        uint32_t ca = cand_a + (uint32_t)buf * (uint32_t)(T6_CH * 16);
#else
        stage6(sorted, sm, nrows, cb, n_ch, lane, cx, cy, cz);
        uint32_t ca = cand_a;
#endif
        // four candidates per step: broadcast loads, five FFMA each, and the histogram update as one IMAD (the
        // float's bits times the row stride, the constant part folded into hbase) + one shared reduction
        int j = 0;
#pragma unroll 2
        for (; j + 4 <= n_ch; j += 4, ca += 64) {
          const float4 P0 = lds_f4<0>(ca), P1 = lds_f4<16>(ca), P2 = lds_f4<32>(ca), P3 = lds_f4<48>(ca);
          if constexpr (BIN == T6_BIN_LOG) {
            const float y0 = bin_y(P0, S, K, A0, A1, A2), y1 = bin_y(P1, S, K, A0, A1, A2);
            const float y2 = bin_y(P2, S, K, A0, A1, A2), y3 = bin_y(P3, S, K, A0, A1, A2);
            red_add_u32(LogBin<CAP>::row(y0) + hbase, inc);
            red_add_u32(LogBin<CAP>::row(y1) + hbase, inc);
            red_add_u32(LogBin<CAP>::row(y2) + hbase, inc);
            red_add_u32(LogBin<CAP>::row(y3) + hbase, inc);
          } else {
            const float z0 = bin_z<NB>(P0, S, K, A0, A1, A2), z1 = bin_z<NB>(P1, S, K, A0, A1, A2);
            const float z2 = bin_z<NB>(P2, S, K, A0, A1, A2), z3 = bin_z<NB>(P3, S, K, A0, A1, A2);
            red_add_u32(__float_as_uint(z0) * 128u + hbase, inc);
            red_add_u32(__float_as_uint(z1) * 128u + hbase, inc);
            red_add_u32(__float_as_uint(z2) * 128u + hbase, inc);
            red_add_u32(__float_as_uint(z3) * 128u + hbase, inc);
          }
        }
        for (; j < n_ch; ++j, ca += 16) {
          const float4 P = lds_f4<0>(ca);
          if constexpr (BIN == T6_BIN_LOG) red_add_u32(LogBin<CAP>::row(bin_y(P, S, K, A0, A1, A2)) + hbase, inc);
          else red_add_u32(__float_as_uint(bin_z<NB>(P, S, K, A0, A1, A2)) * 128u + hbase, inc);
        }
        __syncwarp();
      }
      // the bin holding the k-th smallest, and how many candidates lie at or below it
      if (hist_on) {
        int cum = 0;
        bstar = -1;
#pragma unroll
        for (int b = 0; b < NB; ++b) {
          cum += (int)hcol[b * 32];
          if (bstar < 0 && cum >= k) {
            bstar = b;
            cum_star = cum;
          }
        }
      }
      if (refined) break;
      const bool refine = hist_on && bstar >= 0 && cum_star > CAPM;
      // a refinement pass walks every candidate again for the whole tile: worth it only when several lanes want it;
      // one or two go to the work list instead (their threshold bin stays too full: not served)
      if (__popc(__ballot_sync(PCP_FULL, refine)) < refine_min) break;
      T6STAT(9, refine ? 1 : 0);
      refined = true;
      hist_on = refine;
      if (refine) {
        // New Y' = (Y - w0) / (w1 - w0) with the window [w0, w1) a little wider than the old bin b*: the k-th stays below
        // the new overflow bin; everything below the window saturates to bin 0 (fewer than k candidates: all collected),
        // everything above it to bin NB.
        float w0, w1;
        if constexpr (BIN == T6_BIN_LOG) {
          // old bin: Y in [edge(b* - 1), edge(b*)); the members (Y < edge(b*)) stay below Y' = 1.  The new bins are geometric
          // in Y' (the top one takes the upper 16 % or 8 % of the window): coarse at the low end, fine where the k-th usually lies
          w1 = LogBin<CAP>::edge(bstar) * 1.001f;
          w0 = bstar > 0 ? LogBin<CAP>::edge(bstar - 1) * 0.999f : -0.04f * w1;
        } else {
          // old bin: NB sqrt(Y) in [b* - 1/2, b* + 1/2); window b* - 0.52 .. b* + 0.60: the k-th stays inside (0, 0.92)
          const float r0 = ((float)bstar - 0.52f) / (float)NB, r1 = ((float)bstar + 0.60f) / (float)NB;
          w0 = bstar > 0 ? r0 * r0 : -r1 * r1 * 0.04f;
          w1 = r1 * r1;
        }
        const float gain = 1.f / (w1 - w0);
        off = -w0 * gain;  // (the first pass has off = 0)
        S *= gain;
        A0 = -2.f * S * qx, A1 = -2.f * S * qy, A2 = -2.f * S * qz, K = __fmaf_rn(S, qq, off);
      }
    }
    }  // !RAD
    T6STAT(2, lane == 0 ? total : 0);
    T6STAT(5, (in_grp && bstar < 0) ? 1 : 0);                     // deferred: k-th beyond the cover radius
    T6STAT(6, (in_grp && bstar >= 0 && cum_star > CAPM) ? 1 : 0);  // deferred: threshold bin too full
    const bool grp_ok = RAD ? in_grp : (in_grp && bstar >= 0 && cum_star <= CAPM);
    if (!RAD && in_grp && bstar < 0 && R >= 1) far = true;  // d_k > the lane's cover radius >= one cell
    float t_reach = 0.f;  // largest exact d2 a member of this lane can have
    // Collect test of pass 2, one float compare against thr (other lanes collect nothing):
    //   SQRT    the pass-1 value itself: 2^23 + round(NB sqrt~(Y)) < 2^23 + b* + 1;
    //   LOG     bin(Y) <= b*  <=>  Y < edge(b*), exact: the bins are read off Y's own bits.
    float thr = BIN == T6_BIN_SQRT ? 0.f : -1.f;
    if (grp_ok) {
      const float reach = 2.f * sqrtf(qq) + 2.f * cover_m;
      const float eps_y = 6.03e-8f * (12.f * S * reach * reach + 12.f * fabsf(off) + 4.f);
      if constexpr (RAD) {
        // superset of the neighbours: Y_exact < 1 <=> d2 < r2 in real arithmetic; the reference compares ITS rounded d2
        // (relative error below 2^-21), the computed Y is within eps_y of the exact one
        thr = 1.f + eps_y + 1e-6f;
        t_safe = 0.f;
        t_reach = ((thr + eps_y) / S) * (1.f + 4e-6f);
      } else if constexpr (BIN == T6_BIN_SQRT) {
        thr = T6_MAGIC + (float)(bstar + 1);
        // a candidate that was NOT collected has NB sqrt~(Y) >= b* + 1/2, so Y >= edge^2 (1 - 2^-20) with the MUFU's error;
        // a member has Y <= edge^2 (1 + 2^-20)
        const float edge = ((float)bstar + 0.5f) / (float)NB;
        const float e2 = edge * edge;
        t_safe = ((e2 * (1.f - 2e-6f) - eps_y - off) / S) * (1.f - 4e-6f);
        t_reach = ((e2 * (1.f + 2e-6f) + eps_y - off) / S) * (1.f + 4e-6f);
      } else {
        thr = LogBin<CAP>::edge(bstar);
        // a candidate that was NOT collected has Y_computed >= thr, a member Y_computed < thr; Y_exact is within eps_y
        t_safe = ((thr - eps_y - off) / S) * (1.f - 4e-6f);
        t_reach = ((thr + eps_y - off) / S) * (1.f + 4e-6f);
      }
    }
    __syncwarp();
    // ---- pass 2: collect the sorted positions of the candidates at or below the threshold bin
    if (__any_sync(PCP_FULL, grp_ok)) {
      const bool resident = !RAD && total <= (uint32_t)T6_CH;  // one chunk: still staged from pass 1 (the radius mode has no pass 1)
      // A candidate farther from the bounding box of the served queries than the largest member distance is
      // nobody's member: each chunk is compacted in place (one candidate per lane, ballot + rank) before the
      // per-lane loop, which then runs over about half of it.  Local coordinates throughout.
      const float big = 3.0e38f;
      float bx0 = grp_ok ? qx : big, bx1 = grp_ok ? qx : -big, by0 = grp_ok ? qy : big, by1 = grp_ok ? qy : -big;
      float bz0 = grp_ok ? qz : big, bz1 = grp_ok ? qz : -big;
      float tmax = t_reach;
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
      tmax = tmax * 1.001f + 1e-5f * g.cell * g.cell;  // slack for the rounding of the local coordinates and of the box distance
#if T6_TMA
      if (!resident) chunk_issue(sorted, sm, nrows, 0u, (int)min((uint32_t)T6_CH, total), lane, 0);
#endif
#pragma unroll 1
      for (uint32_t cb = 0; cb < total; cb += T6_CH) {
        int n_ch = (int)min((uint32_t)T6_CH, total - cb);
#if T6_TMA
        const int buf = (int)((cb / T6_CH) & 1u);
        if (cb + T6_CH < total) chunk_issue(sorted, sm, nrows, cb + T6_CH, (int)min((uint32_t)T6_CH, total - cb - T6_CH), lane, buf ^ 1);
        if (resident) chunk_land<CAP, false, true>(sm, nrows, cb, n_ch, lane, buf, parity, cx, cy, cz);
        else chunk_land<CAP, true, true>(sm, nrows, cb, n_ch, lane, buf, parity, cx, cy, cz);
        float4* const cand = sm.cand[buf];
#else
        if (!resident) stage6(sorted, sm, nrows, cb, n_ch, lane, cx, cy, cz);
        float4* const cand = sm.cand[0];
#endif
        {
          int kept = 0;
#pragma unroll 1
          for (int i0 = 0; i0 < n_ch; i0 += 32) {
            const int i = i0 + lane;
            float4 P = make_float4(0.f, 0.f, 0.f, 0.f);
            uint32_t ps = 0;
            bool keep = false;
            if (i < n_ch) {
              P = cand[i];
              ps = sm.cpos[i];
              const float dx = fmaxf(fmaxf(bx0 - P.x, P.x - bx1), 0.f), dy = fmaxf(fmaxf(by0 - P.y, P.y - by1), 0.f);
              const float dz = fmaxf(fmaxf(bz0 - P.z, P.z - bz1), 0.f);
              keep = dx * dx + dy * dy + dz * dz <= tmax;
            }
            const unsigned km = __ballot_sync(PCP_FULL, keep);
            __syncwarp();  // every lane has read its candidate before any slot is overwritten
            if (keep) {
              const int d = kept + __popc(km & ((1u << lane) - 1u));
              cand[d] = P;
              sm.cpos[d] = ps;
            }
            kept += __popc(km);
          }
          __syncwarp();
          T6STAT(8, lane == 0 ? kept : 0);
          n_ch = kept;
        }
#if T6_TMA
        uint32_t ca = cand_a + (uint32_t)buf * (uint32_t)(T6_CH * 16), pa = cpos_a;
#else
        uint32_t ca = cand_a, pa = cpos_a;
#endif
        int j = 0;
#pragma unroll 2
        for (; j + 4 <= n_ch; j += 4, ca += 64, pa += 16) {
          const float4 P0 = lds_f4<0>(ca), P1 = lds_f4<16>(ca), P2 = lds_f4<32>(ca), P3 = lds_f4<48>(ca);
          float z0, z1, z2, z3;
          if constexpr (BIN == T6_BIN_SQRT) {
            z0 = bin_z<NB>(P0, S, K, A0, A1, A2), z1 = bin_z<NB>(P1, S, K, A0, A1, A2);
            z2 = bin_z<NB>(P2, S, K, A0, A1, A2), z3 = bin_z<NB>(P3, S, K, A0, A1, A2);
          } else {
            z0 = bin_y<!RAD>(P0, S, K, A0, A1, A2), z1 = bin_y<!RAD>(P1, S, K, A0, A1, A2);
            z2 = bin_y<!RAD>(P2, S, K, A0, A1, A2), z3 = bin_y<!RAD>(P3, S, K, A0, A1, A2);
          }
          // (radius: a neighbourhood may hold more hits than the column has rows; the pointer keeps counting, the
          // stores stop at the column's end and the lane is handed to the work list)
          if (z0 < thr) { if (!RAD || wp < wend) sts_u32(wp, lds_u32<0>(pa)); wp += 128u; }
          if (z1 < thr) { if (!RAD || wp < wend) sts_u32(wp, lds_u32<4>(pa)); wp += 128u; }
          if (z2 < thr) { if (!RAD || wp < wend) sts_u32(wp, lds_u32<8>(pa)); wp += 128u; }
          if (z3 < thr) { if (!RAD || wp < wend) sts_u32(wp, lds_u32<12>(pa)); wp += 128u; }
        }
        for (; j < n_ch; ++j, ca += 16, pa += 4) {
          const float4 P = lds_f4<0>(ca);
          const float z = BIN == T6_BIN_SQRT ? bin_z<NB>(P, S, K, A0, A1, A2) : bin_y<!RAD>(P, S, K, A0, A1, A2);
          if (z < thr) { if (!RAD || wp < wend) sts_u32(wp, lds_u32<0>(pa)); wp += 128u; }
        }
        __syncwarp();
      }
    }
    ok = ok || grp_ok;
  }
  int cnt = ok ? (int)((wp - hcol_a) >> 7) : 0;  // members collected into this lane's column
  if (RAD && cnt > CAP) {  // more candidates inside the radius than the column holds: the work list's 512-entry buffer takes it
    ok = false;
    cnt = 0;
  }
  T6STAT(0, finite ? 1 : 0);
  T6STAT(3, cnt);
  T6STAT(4, pending ? 1 : 0);  // deferred: more row groups in the tile than the kernel serves
  // The resolve step, the sorting network and the finish stage are straight-line code walked once per tile: the warps
  // of a CTA re-align here so that they fetch it together (one miss serves four warps) instead of evicting each
  // other's lines.
#if T6_SYNC
  __syncthreads();
#endif
  // ---- resolve + sort: every member's reference distance (FLANN L2_Simple arithmetic on the ORIGINAL coordinates) as a
  // 32-bit key prefix(d2) | slot on registers; a d2 >= 0 orders like its bit pattern, the slot bits only make the
  // keys distinct.  After the network, slot j of the column is rewritten with the position of the j-th key's member.
  uint32_t kth_key = 0xFFFFFFFFu;
  int n_in = 0;  // radius: members that pass the exact test (they sort in front of the dropped ones)
  {
    uint32_t keys[CAP];
    const int maxc = __reduce_max_sync(PCP_FULL, cnt);  // blocks of eight slots nobody uses are skipped (k = 10: half of them)
#pragma unroll
    for (int j0 = 0; j0 < CAP; j0 += 8) {
#pragma unroll
      for (int j = j0; j < j0 + 8; ++j) keys[j] = 0xFFFFFFFFu;
      if (j0 < maxc) {
#pragma unroll
        for (int j = j0; j < j0 + 8; ++j) {
          if (j < cnt) {
            const float4 P = __ldg(&sorted[hcol[j * 32]]);
            const float d2 = dist2_nofma(Q.x, Q.y, Q.z, P);
            if (!RAD || d2 < sp.r2) {  // radius: strict d2 < r2 with the reference's own arithmetic decides membership
              keys[j] = (__float_as_uint(d2) & ~(uint32_t)(CAP - 1)) | (uint32_t)j;
              ++n_in;
            }
          }
        }
      }
    }
    if (CAP == 32) sort32_regs(reinterpret_cast<uint32_t(&)[32]>(keys));
    else sort64_regs(reinterpret_cast<uint32_t(&)[64]>(keys));
#pragma unroll
    for (int j = 0; j < CAP; ++j) {
      if (j == k - 1) kth_key = keys[j];
      keys[j] = j < (RAD ? n_in : cnt) ? hcol[(keys[j] & (uint32_t)(CAP - 1)) * 32] : 0u;
    }
#pragma unroll
    for (int j = 0; j < CAP; ++j)
      if (j < (RAD ? n_in : cnt)) hcol[j * 32] = keys[j];
  }
  if (RAD) cnt = n_in;
  // ---- proof of the selection: every point that was not collected is at least t_safe away; the k-th member's
  // distance is at most its key with the slot bits set
  if (!RAD && ok && !(__uint_as_float(kth_key | (uint32_t)(CAP - 1)) < t_safe)) {
    ok = false;
    far = false;
    T6STAT(10, 1);
  }
  T6STAT(1, (finite && !ok) ? 1 : 0);
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
  const int n_nb = ok ? (RAD ? cnt : k) : 0;
  const bool emit = active && (ok || !finite);  // non-finite queries get their empty record here
  __syncwarp();
#if T6_SYNC
  __syncthreads();
#endif
  if (!live_warp) return;
  knn_finish_pos<LEAN>(a, hcol, cnt, n_nb, ~(uint32_t)(CAP - 1), Q, emit ? my_slot : -1, sm.u.rec, lane, RAD ? sp.min_pts : 1);
}

}  // namespace

#ifdef PCP_STATS
extern "C" __attribute__((visibility("default"))) int pcp_debug_tile6_stats(unsigned long long* out) {
  unsigned long long z[16] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_tile6_stats, sizeof(z));
  cudaMemcpyToSymbol(g_tile6_stats, z, sizeof(z));
  return 0;
}
#endif

// Launches the k-NN tile kernel over the whole cloud; deferred queries are left in ctx->work_list /
// ctx->work_count (device) for the work-list kernel.  32-member lanes for k <= 24, 64-member lanes for 24 < k <= 32.
// sp.mode == 1: fixed-radius search through the same kernel (T6_BIN_RADIUS), lanes of `radius_cap` (32 or 64) hits.
int pcp_knn_tile6_launch(pcp_ctx* ctx, const SearchParams& sp, const KArgs& a, int radius_cap) {
  int rc;
  if ((rc = pcp_reserve(ctx, ctx->work_list, (size_t)ctx->n * sizeof(uint32_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->work_count, 2 * sizeof(uint32_t)))) return rc;  // [0] list length, [1] ticket counter of the work-list kernel
  PCP_CUDA(cudaMemsetAsync(ctx->work_count.p, 0, 2 * sizeof(uint32_t), ctx->stream));
  const int64_t tiles = (ctx->n + 31) / 32;
  const int blocks = (int)((tiles + T6_WARPS - 1) / T6_WARPS);
  // Stencil rows served per tile.  One on big clouds (a second row group costs the urban cloud 5-10 %); up to four
  // when the rows are short — frame batches: 26 k points over a few hundred rows, where 8-30 % of the lanes sit in
  // another row than their tile's majority.
  const double pts_per_row = (double)ctx->n / ((double)ctx->grid.n_seg * (double)ctx->grid.dim[1] * (double)ctx->grid.dim[2]);
  int max_groups = pts_per_row < 256.0 ? 4 : 1;
  if (ctx->knobs.knn_groups > 0) max_groups = ctx->knobs.knn_groups;
  uint32_t* wc = (uint32_t*)ctx->work_count.p;
  uint32_t* wl = (uint32_t*)ctx->work_list.p;
  const uint32_t* perm = nullptr;
#if T6_ORDER
  if (tiles > 2 * T6_WARPS && !ctx->knobs.t6_no_order) {
    if ((rc = pcp_reserve(ctx, ctx->tile_perm, (size_t)tiles * sizeof(uint32_t)))) return rc;
    tile_order_kernel<<<(unsigned)((tiles + T6_WIN - 1) / T6_WIN), T6_WIN_THREADS, 0, ctx->stream>>>(a, sp.R0, tiles, (uint32_t*)ctx->tile_perm.p);
    PCP_LAUNCH_CHECK("tile_order_kernel");
    perm = (const uint32_t*)ctx->tile_perm.p;
  }
#endif
  const bool lean = sp.out != nullptr && sp.center_mode == PCP_CENTER_QUERY && sp.normals == nullptr && sp.nn_idx == nullptr && sp.valid == nullptr;
#define T6_LAUNCH(CAPV, LEANV, BINV)                                                                                            \
  do {                                                                                                                         \
    const size_t smem = sizeof(T6Smem<CAPV>) * T6_WARPS;                                                                       \
    PCP_CUDA(cudaFuncSetAttribute(knn_tile6_kernel<CAPV, LEANV, BINV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    PCP_CUDA(pcp_launch_search(ctx, knn_tile6_kernel<CAPV, LEANV, BINV>, dim3(blocks), dim3(T6_WARPS * 32), smem, a, sp.R0, max_groups, wc, wl, perm, tiles, refine_min)); \
  } while (0)
#define T6_LAUNCH_BIN(CAPV, LEANV)                                         \
  do {                                                                     \
    if (bin == T6_BIN_LOG) T6_LAUNCH(CAPV, LEANV, T6_BIN_LOG);             \
    else T6_LAUNCH(CAPV, LEANV, T6_BIN_SQRT);                              \
  } while (0)
  // bin mode (see T6_BIN_*): PCP_T6_BIN=sqrt|log forces one
  // Measured on one B200 (1 M-point urban cloud, search stage, sqrt / geometric bins): k = 10 1.11 / 1.05 ms, k = 15 1.145 /
  // 1.094, k = 20 1.231 / 1.187, k = 24 1.38 / 1.45 (four bins per octave are too coarse once 1.2-1.3 k members approach
  // the 32-row column), k = 30 (64-row columns, eight per octave) 1.68 / 1.61; 500 k-point playground, k = 20: 0.764 / 0.725.
  // A refinement pass runs when at least 6 lanes of a tile want one (always: k = 20 1.26 instead of 1.19 ms).
  const int refine_min = ctx->knobs.t6_refine_min > 0 ? ctx->knobs.t6_refine_min : 6;
  int bin = (sp.k >= 22 && sp.k <= PCP_T6_K32) ? T6_BIN_SQRT : T6_BIN_LOG;
  if (ctx->knobs.t6_bin >= 0) bin = ctx->knobs.t6_bin;
  if (sp.mode == 1) {
    if (radius_cap <= 32) {
      if (lean) T6_LAUNCH(32, true, T6_BIN_RADIUS); else T6_LAUNCH(32, false, T6_BIN_RADIUS);
    } else {
      if (lean) T6_LAUNCH(64, true, T6_BIN_RADIUS); else T6_LAUNCH(64, false, T6_BIN_RADIUS);
    }
  } else if (sp.k <= PCP_T6_K32) {
    if (lean) T6_LAUNCH_BIN(32, true); else T6_LAUNCH_BIN(32, false);
  } else {
    if (lean) T6_LAUNCH_BIN(64, true); else T6_LAUNCH_BIN(64, false);
  }
#undef T6_LAUNCH_BIN
#undef T6_LAUNCH
  PCP_LAUNCH_CHECK("knn_tile6_kernel");
  return PCP_OK;
}
