// search_pca.cu — the fused hot kernel: per-point neighbour search over the cell-sorted grid,
// (d2, index)-ordered neighbour list, second moment in the reference's accumulation order,
// 3×3 eigen-decomposition and structure features, without the neighbour list ever
// leaving the SM.
//
// Replaces, per point, the body of the loop at src/pcaAnalysis.cpp:18-40:
//   kdTreeFla.radiusSearch(i, radius, idx, d2)                     (:25)
//   pcl::computeCovarianceMatrixNormalized(cloud, idx, pt, C)      (:55)   SURVEY A.2
//   Eigen::SelfAdjointEigenSolver<Matrix3f>(C)                     (:57)   eigen3.cuh
//   lineStruc / planeStruc / scatterStruc                          (:29-31)
// and the k-nearest flavour reached through pcl::NormalEstimation (src/smoothing.cpp:239-246).
//
// Work decomposition: one warp owns a tile of 32 consecutive cell-sorted points.  For each
// query of the tile the 32 lanes scan the stencil rows (each row = 2R+1 x-adjacent cells =
// ONE contiguous run of the sorted float4 array, read coalesced), keep the running best-k in
// registers (one key per lane, bitonic sort/merge through shuffles) or, for radius search,
// append hits to a warp-private shared buffer; the ordered moment sums run as six
// independent chains on lanes 0-5.  After the 32 queries every lane solves the eigenproblem
// of ONE query, so the fp64 Jacobi work is fully lane-parallel.
#include <cfloat>
#include <cmath>
#include <cstdlib>

#include "eigen3.cuh"
#include "knn_finish.cuh"
#include "pcp_internal.cuh"
#include "search_common.cuh"

#ifdef PCP_STATS
__device__ unsigned long long g_stats[8];
#define STAT(i, v) do { if (lane == 0) atomicAdd(&g_stats[i], (unsigned long long)(v)); } while (0)
#else
#define STAT(i, v) do { } while (0)
#endif

namespace {

#ifndef PCP_RADIUS_T6_DEFAULT
#define PCP_RADIUS_T6_DEFAULT 1   // measured: 4 M points at configs[3] density (mean 31.6 hits) 5.12 -> 4.52 ms, urban 1 M (mean 8.7) 0.735 -> 0.706 ms
#endif
#ifndef SP_WARPS
#define SP_WARPS 8
#endif
#ifndef SP_MINB
#define SP_MINB 3     // CTAs per SM the kernels are compiled for
#endif
constexpr int WARPS = SP_WARPS;     // warps per CTA
constexpr int RADIUS_CAP = 512;     // radius mode: neighbours kept per query in shared memory
constexpr int REC_WORDS = 17;       // 68-byte record
constexpr int KNN_LW = 16;          // k-NN: queries per tile of the warp-per-query kernel (columns of the list planes)

// shared memory of one warp; the record staging area reuses the search buffers (disjoint phases)
template <int MODE>
struct WarpSmem {
  union U {
    struct S {
      u64 buf[MODE == 0 ? 64 : RADIUS_CAP];
      float prod[6][32];
      uint2 runs[32];  // (first sorted position, virtual offset) of the non-empty runs of a row chunk
    } s;
    float rec[32][REC_WORDS];
  } u;
  float mom[MODE == 0 ? 1 : 32][6];    // radius mode: query-centred moment (normalised)
  float momM[MODE == 0 ? 1 : 32][6];   // radius mode: mean-centred moment (normalised), only when needed
  int cnt[32];
  int gidq[32];
  // k-NN: the ordered neighbour lists of the tile's queries, query q in column q (KNN_LW columns, 32 rows of
  // ord and of bits(d2)), so that the moment / eigen / record stage runs once per tile with one query per
  // lane (knn_finish) instead of once per query with the warp idling behind six summing lanes
  uint32_t list_lo[MODE == 0 ? 32 * KNN_LW : 1];
  uint32_t list_hi[MODE == 0 ? 32 * KNN_LW : 1];
};

// Ordered second moment about (cx,cy,cz) of the m neighbours whose keys sit in list[0..m)
// (registers for m <= 32 via `regkey`, shared memory otherwise).  Lanes 0..5 return
// c00,c01,c02,c11,c12,c22 (already divided by m); accumulation order is PCL's.
template <bool FROM_SMEM>
__device__ __noinline__ float ordered_moment(const float4* __restrict__ sorted, const uint32_t* __restrict__ inv, const u64* list,
                                             u64 regkey, int m, float cx, float cy, float cz, float (*prod)[32], int lane) {
  float acc = 0.f;
  for (int base = 0; base < m; base += 32) {
    int j = base + lane;
    if (j < m) {
      u64 key = FROM_SMEM ? list[j] : regkey;
      float4 p = __ldg(&sorted[__ldg(&inv[(uint32_t)key])]);
      float px = __fsub_rn(p.x, cx), py = __fsub_rn(p.y, cy), pz = __fsub_rn(p.z, cz);
      prod[0][lane] = __fmul_rn(px, px);  // c00
      prod[1][lane] = __fmul_rn(py, px);  // c01
      prod[2][lane] = __fmul_rn(pz, px);  // c02
      prod[3][lane] = __fmul_rn(py, py);  // c11
      prod[4][lane] = __fmul_rn(py, pz);  // c12
      prod[5][lane] = __fmul_rn(pz, pz);  // c22
    }
    __syncwarp();
    if (lane < 6) {
      int cnt = min(32, m - base);
#pragma unroll 4
      for (int t = 0; t < cnt; ++t) acc = __fadd_rn(acc, prod[lane][t]);
    }
    __syncwarp();
  }
  if (m > 0) acc = __fdiv_rn(acc, (float)m);
  return acc;
}

// fp32 running-sum centroid in list order (pcl::compute3DCentroid); result on every lane
template <bool FROM_SMEM>
__device__ __noinline__ void ordered_centroid(const float4* __restrict__ sorted, const uint32_t* __restrict__ inv, const u64* list,
                                              u64 regkey, int m, float (*prod)[32], int lane, float& cx, float& cy, float& cz) {
  float acc = 0.f;
  for (int base = 0; base < m; base += 32) {
    int j = base + lane;
    if (j < m) {
      u64 key = FROM_SMEM ? list[j] : regkey;
      float4 p = __ldg(&sorted[__ldg(&inv[(uint32_t)key])]);
      prod[0][lane] = p.x;
      prod[1][lane] = p.y;
      prod[2][lane] = p.z;
    }
    __syncwarp();
    if (lane < 3) {
      int cnt = min(32, m - base);
      for (int t = 0; t < cnt; ++t) acc = __fadd_rn(acc, prod[lane][t]);
    }
    __syncwarp();
  }
  acc = __fdiv_rn(acc, (float)m);
  cx = __shfl_sync(PCP_FULL, acc, 0);
  cy = __shfl_sync(PCP_FULL, acc, 1);
  cz = __shfl_sync(PCP_FULL, acc, 2);
}

// Per-query search state of a warp.
struct Search {
  u64 best;         // k-NN: ascending best list, one key per lane
  u64 thr;          // k-NN: key of the current k-th best (INF until k candidates were merged)
  int cnt;          // entries waiting in the shared buffer (k-NN) / hits so far (radius)
};

// Scan every cell of the block of half-width R around (c0,c1,c2) that is NOT in the block of
// half-width Rp (Rp < 0: nothing scanned yet).  `bound2` prunes rows and clips runs by their
// distance to the query (radius: r2; k-NN: the current k-th distance or +inf).
template <int MODE>
__device__ __forceinline__ void scan_shell(const KArgs& a, WarpSmem<MODE>* ws, Search& st, const float4 Q, int c0, int c1, int c2,
                                           int zoff, float f0, float f1, float f2, int Rp, int R, float bound2, int lane, uint32_t dirs) {
  const GridDesc& g = a.g;
  const float4* __restrict__ sorted = a.sorted;
  const unsigned lt_mask = (1u << lane) - 1u, le_mask = (2u << lane) - 1u;
  const int lo1 = max(c1 - R, 0), hi1 = min(c1 + R, g.dim[1] - 1), n1 = hi1 - lo1 + 1;
  const int lo2 = max(c2 - R, 0), hi2 = min(c2 + R, g.dim[2] - 1), n2 = hi2 - lo2 + 1;
  const int nrows = n1 * n2;
  const int lo0 = max(c0 - R, 0), hi0 = min(c0 + R, g.dim[0] - 1);
  const bool prune = bound2 < 3.0e38f;
  const float inv_c2 = prune ? (float)(g.inv_cell * g.inv_cell) * bound2 * 1.0002f : 0.f;  // bound in cells^2, widened
  for (int rb = 0; rb < nrows; rb += 32) {
    const int r = rb + lane;
    // up to two runs per row: [bA, eA) and [bB, eB) in sorted positions
    uint32_t bA = 0, eA = 0, bB = 0, eB = 0;
    if (r < nrows) {
      const int o2 = r / n1, o1 = r - o2 * n1;
      int e1 = c1 + o1, e2 = c2 + o2;  // start at the query's own row, wrap to the low side
      if (e1 > hi1) e1 -= n1;
      if (e2 > hi2) e2 -= n2;
      const int d1 = e1 - c1, d2 = e2 - c2;
      int xlo = lo0, xhi = hi0;
      bool keep = true;
      if (prune) {
        float g1 = cell_gap(d1, f1), g2 = cell_gap(d2, f2);
        float rest = inv_c2 - (g1 * g1 + g2 * g2) * 0.9998f;
        keep = rest >= 0.f;
        if (keep) {
          float hx = sqrtf(rest) + 4e-3f;  // cells; offsets taken relative to c0 to stay exact
          xlo = max(xlo, c0 + (int)floorf(f0 - hx));
          xhi = min(xhi, c0 + (int)floorf(f0 + hx));
        }
      }
      if (keep && xlo <= xhi) {
        const uint32_t base = ((uint32_t)(zoff + e2) * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
        const bool inner = Rp >= 0 && abs(d1) <= Rp && abs(d2) <= Rp;
        if (!inner) {
          bA = __ldg(&a.E[base + xlo]);
          eA = __ldg(&a.E[base + xhi + 1]);
        } else {
          const int ahi = min(xhi, c0 - Rp - 1), blo = max(xlo, c0 + Rp + 1);
          if (xlo <= ahi) {
            bA = __ldg(&a.E[base + xlo]);
            eA = __ldg(&a.E[base + ahi + 1]);
          }
          if (blo <= xhi) {
            bB = __ldg(&a.E[base + blo]);
            eB = __ldg(&a.E[base + xhi + 1]);
          }
        }
      }
    }
    // Flatten the non-empty runs of this row chunk into ONE virtual sequence so that every
    // 32-candidate step uses all lanes (stencil rows are often short): run descriptors are
    // compacted by rank into shared memory, and each step maps lane -> (run, offset) with one
    // warp OR-reduce of the run-start bits and a popc.
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const uint32_t rbeg = half == 0 ? bA : bB;
      const uint32_t len = half == 0 ? eA - bA : eB - bB;
      const unsigned ne = __ballot_sync(PCP_FULL, len > 0);
      if (ne == 0) continue;
      uint32_t inc = len;  // inclusive prefix of the run lengths
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(PCP_FULL, inc, o);
        if (lane >= o) inc += t;
      }
      const uint32_t total = __shfl_sync(PCP_FULL, inc, 31);
      const uint32_t myoff = inc - len;
      if (len > 0) ws->u.s.runs[__popc(ne & lt_mask)] = make_uint2(rbeg, myoff);
      __syncwarp();
      STAT(2, __popc(ne));
      uint32_t starts_before = 0;
      for (uint32_t t0 = 0; t0 < total; t0 += 32) {
        STAT(3, 1);
        STAT(4, min(32u, total - t0));
        const uint32_t rel = myoff - t0;  // wraps for runs that started earlier
        const unsigned smask = __reduce_or_sync(PCP_FULL, (len > 0 && rel < 32u) ? (1u << rel) : 0u);
        const uint32_t v = t0 + lane;
        const bool live = v < total;
        const int ridx = (int)starts_before + __popc(smask & le_mask) - 1;
        starts_before += __popc(smask);
        bool pass = false;
        u64 key = 0;
        if (live) {
          const uint2 rd = ws->u.s.runs[ridx];
          const float4 P = __ldg(&sorted[rd.x + (v - rd.y)]);
          const float d2v = dist2_nofma(Q.x, Q.y, Q.z, P);
          key = ((u64)__float_as_uint(d2v) << 32) | (u64)__float_as_uint(P.w);
          pass = MODE == 0 ? key < st.thr : d2v < bound2;  // radius mode: bound2 = r2 (or the k-by-cover radius of this pass)
        }
        const unsigned m = __ballot_sync(PCP_FULL, pass);
        if (m == 0) continue;
        STAT(5, __popc(m));
        const int off = st.cnt + __popc(m & lt_mask);
        if (MODE == 0) {
          if (pass) ws->u.s.buf[off] = key;
          st.cnt += __popc(m);
          __syncwarp();
          if (st.cnt >= 32) {
            STAT(6, 1);
            u64 v32 = warp_sort32(ws->u.s.buf[lane], dirs);
            st.best = warp_merge32(st.best, v32, lane, dirs);
            st.thr = shfl_u64(st.best, a.sp.k - 1);
            u64 rem = (lane + 32 < st.cnt) ? ws->u.s.buf[lane + 32] : 0ull;
            __syncwarp();
            if (lane + 32 < st.cnt) ws->u.s.buf[lane] = rem;
            st.cnt -= 32;
            __syncwarp();
          }
        } else {
          if (pass && off < RADIUS_CAP) ws->u.s.buf[off] = key;
          st.cnt += __popc(m);
        }
      }
      __syncwarp();  // runs[] is rewritten by the next half / row chunk
    }
  }
}

// One tile: lane l holds the sorted position `my_pos` of query l (valid for l < nq); the warp solves the
// Radius neighbourhoods larger than the warp buffer (RADIUS_CAP hits; e.g. r = 3 m on a dense scan,
// src/lidarOdometry.cpp:76): no list can be kept, so the hits are accumulated where they are found — one
// walk over the stencil per call, every lane summing its own candidates in DOUBLE, then a warp reduction:
// count, coordinate sums and the six second moments about (cx, cy, cz).  The reference sums the same
// terms in fp32 in (d2, index) order; the double sums differ from that by rounding only (well inside the
// 1e-4 parity bar, and closer to the exact moment).  Used for records / normals, never for neighbour lists.
__device__ __noinline__ void heavy_moment(const KArgs& a, const float4 Q, int c0, int c1, int c2, int zoff, int R, float r2, double cx,
                                          double cy, double cz, int lane, double* sums /*3*/, double* mom /*6*/) {
  const GridDesc& g = a.g;
  const int lo0 = max(c0 - R, 0), hi0 = min(c0 + R, g.dim[0] - 1);
  const int lo1 = max(c1 - R, 0), hi1 = min(c1 + R, g.dim[1] - 1);
  const int lo2 = max(c2 - R, 0), hi2 = min(c2 + R, g.dim[2] - 1);
  double s[3] = {0, 0, 0}, m[6] = {0, 0, 0, 0, 0, 0};
  for (int e2 = lo2; e2 <= hi2; ++e2)
    for (int e1 = lo1; e1 <= hi1; ++e1) {
      const uint32_t base = ((uint32_t)(zoff + e2) * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
      const uint32_t b = __ldg(&a.E[base + lo0]), e = __ldg(&a.E[base + hi0 + 1]);
      for (uint32_t j = b + lane; j < e; j += 32) {
        const float4 P = __ldg(&a.sorted[j]);
        if (dist2_nofma(Q.x, Q.y, Q.z, P) < r2) {
          const double px = (double)P.x - cx, py = (double)P.y - cy, pz = (double)P.z - cz;
          s[0] += (double)P.x; s[1] += (double)P.y; s[2] += (double)P.z;
          m[0] += px * px; m[1] += px * py; m[2] += px * pz; m[3] += py * py; m[4] += py * pz; m[5] += pz * pz;
        }
      }
    }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int t = 0; t < 3; ++t) s[t] += __shfl_xor_sync(PCP_FULL, s[t], o);
#pragma unroll
    for (int t = 0; t < 6; ++t) m[t] += __shfl_xor_sync(PCP_FULL, m[t], o);
  }
#pragma unroll
  for (int t = 0; t < 3; ++t) sums[t] = s[t];
#pragma unroll
  for (int t = 0; t < 6; ++t) mom[t] = m[t];
}

// cell coordinates of the query at P with cell id `cid` (c2 frame-local, zoff = its frame's first z row) and its fractional
// position inside the cell, clamped to [0, 1] (same double arithmetic as the key kernel; selects, not indexed arrays)
struct QueryCell {
  int c0, c1, c2, zoff;
  float f0, f1, f2;
};
__device__ __forceinline__ QueryCell query_cell(const GridDesc& g, const float4 P, const uint32_t cid) {
  QueryCell qc;
  qc.c0 = (int)(cid % (uint32_t)g.dim[0]);
  const uint32_t tq = cid / (uint32_t)g.dim[0];
  qc.c1 = (int)(tq % (uint32_t)g.dim[1]);
  qc.c2 = (int)(tq / (uint32_t)g.dim[1]);
  qc.zoff = pcp_unfold_z(g, qc.c2);
  const double qx = g.ax[0] == 0 ? P.x : (g.ax[0] == 1 ? P.y : P.z), qy = g.ax[1] == 0 ? P.x : (g.ax[1] == 1 ? P.y : P.z);
  const double qz = g.ax[2] == 0 ? P.x : (g.ax[2] == 1 ? P.y : P.z);
  const double ox = g.ax[0] == 0 ? g.origin[0] : (g.ax[0] == 1 ? g.origin[1] : g.origin[2]);
  const double oy = g.ax[1] == 0 ? g.origin[0] : (g.ax[1] == 1 ? g.origin[1] : g.origin[2]);
  const double oz = g.ax[2] == 0 ? g.origin[0] : (g.ax[2] == 1 ? g.origin[1] : g.origin[2]);
  qc.f0 = fminf(fmaxf((float)((qx - ox) * g.inv_cell - (double)qc.c0), 0.f), 1.f);
  qc.f1 = fminf(fmaxf((float)((qy - oy) * g.inv_cell - (double)qc.c1), 0.f), 1.f);
  qc.f2 = fminf(fmaxf((float)((qz - oz) * g.inv_cell - (double)qc.c2), 0.f), 1.f);
  return qc;
}

// nq queries one after the other and writes their outputs.
template <int MODE>
// `my_r0`: first stencil half-width of this lane's query (k-NN; the tile kernel flags the entries whose k-th
// neighbour it found beyond the half-width-1 block, so their scan of that block is not repeated for nothing)
__device__ __forceinline__ void search_tile(const KArgs& a, WarpSmem<MODE>* ws, const int lane, const int64_t my_pos, const int nq,
                                            const int my_r0) {
  const float4* __restrict__ sorted = a.sorted;
  const GridDesc& g = a.g;
  const SearchParams& sp = a.sp;
  const bool need_q = sp.out != nullptr && sp.center_mode == PCP_CENTER_QUERY;
  const bool need_m = sp.normals != nullptr || (sp.out != nullptr && sp.center_mode == PCP_CENTER_MEAN);
  const int my_slot = lane < nq ? a.slot[my_pos] : -1;
  const uint32_t dirs = bitonic_dirs(lane);
  const int max_dim = max(g.dim[0], max(g.dim[1], g.dim[2]));
  // lane j remembers query j of the tile and, once solved, its k-th distance: later queries seed
  // their threshold from them (triangle inequality: d_k(q) <= d_k(q_j) + |q − q_j|)
  const float4 myP = lane < nq ? __ldg(&sorted[my_pos]) : make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t my_cid = lane < nq ? __ldg(&a.keys[my_pos]) : g.n_cells;
  const uint32_t my_seg = (g.n_seg > 1 && lane < nq) ? my_cid / g.seg_cells : 0u;  // a seed must come from the same frame
  float my_dk = 3.4e38f;
  // k-NN: lane j works out the cell of ITS query once and parks it in warp-private shared memory (the product rows and
  // the id row, which only the radius mode uses during the search): the query loop then costs neither the two integer
  // divisions and the double arithmetic per query nor registers across the loop
  if (MODE == 0) {
    QueryCell mine = {0, 0, 0, 0, 0.f, 0.f, 0.f};
    if (my_cid < g.n_cells) mine = query_cell(g, myP, my_cid);
    __syncwarp();
    ws->u.s.prod[0][lane] = __int_as_float(mine.c0);
    ws->u.s.prod[1][lane] = __int_as_float(mine.c1);
    ws->u.s.prod[2][lane] = __int_as_float(mine.c2);
    ws->u.s.prod[3][lane] = mine.f0;
    ws->u.s.prod[4][lane] = mine.f1;
    ws->u.s.prod[5][lane] = mine.f2;
    ws->gidq[lane] = mine.zoff;
    __syncwarp();
  }

#pragma unroll 1
  for (int q = 0; q < nq; ++q) {
    const int slot_q = __shfl_sync(PCP_FULL, my_slot, q);
    if (slot_q < 0) continue;  // halo point: neighbour only
    STAT(0, 1);
    // the query's point and cell were loaded by lane q with the tile's other queries (one round trip to L2 per tile,
    // not two dependent ones per query)
    const float4 Q = make_float4(__shfl_sync(PCP_FULL, myP.x, q), __shfl_sync(PCP_FULL, myP.y, q), __shfl_sync(PCP_FULL, myP.z, q),
                                 __shfl_sync(PCP_FULL, myP.w, q));
    const uint32_t cid = __shfl_sync(PCP_FULL, my_cid, q);
    int n_nb = 0;
    bool overflow = false, heavy = false;
    Search st;
    st.best = PCP_KEY_INF;
    st.thr = PCP_KEY_INF;
    st.cnt = 0;
    if (cid < g.n_cells) {
      // cell coordinates, frame row offset and fractional position inside the own cell: lane q worked them out for
      // its query before the loop
      QueryCell qc;
      if (MODE == 0) {
        qc.c0 = __float_as_int(ws->u.s.prod[0][q]);
        qc.c1 = __float_as_int(ws->u.s.prod[1][q]);
        qc.c2 = __float_as_int(ws->u.s.prod[2][q]);
        qc.zoff = ws->gidq[q];
        qc.f0 = ws->u.s.prod[3][q];
        qc.f1 = ws->u.s.prod[4][q];
        qc.f2 = ws->u.s.prod[5][q];
      } else {
        qc = query_cell(g, Q, cid);
      }
      const int c0 = qc.c0, c1 = qc.c1, c2 = qc.c2, zoff = qc.zoff;
      const float f0 = qc.f0, f1 = qc.f1, f2 = qc.f2;
      if (MODE == 1 && sp.k > 0) {
        // k nearest for 32 < k <= 64 (the best list above holds one key per lane): collect EVERY point inside
        // a radius the stencil is known to cover, and keep the first k of the sorted hits.  Fewer than k hits:
        // grow the stencil; more than the buffer holds: bisect the radius between a value known to hold < k
        // points and the one that overflowed.  Exact, ties included (all points at the k-th distance lie
        // inside the radius and are ordered by index).
        int R = sp.R0;
        float lo2 = 0.f, hi2 = 0.f;
        const float diag2 = g.cell * g.cell * 1.01f *
                            ((float)g.dim[0] * (float)g.dim[0] + (float)g.dim[1] * (float)g.dim[1] + (float)g.dim[2] * (float)g.dim[2]) + 1e-30f;
        overflow = true;
#pragma unroll 1
        for (int it = 0; it < 96; ++it) {
          const bool whole = (c0 - R <= 0) && (c0 + R >= g.dim[0] - 1) && (c1 - R <= 0) && (c1 + R >= g.dim[1] - 1) &&
                             (c2 - R <= 0) && (c2 + R >= g.dim[2] - 1);
          // the whole grid in reach: every point lies within the box diagonal (a finite radius the bisection can halve)
          const float cov = whole ? diag2 : cover2_of(R, g.cell);
          const bool bisect = hi2 > 0.f && hi2 <= cov;
          const float r2q = bisect ? 0.5f * lo2 + 0.5f * hi2 : cov;
          if (bisect && !(r2q > lo2 && r2q < hi2)) break;  // more than RADIUS_CAP points at one distance: give up
          st.cnt = 0;
          __syncwarp();
          scan_shell<MODE>(a, ws, st, Q, c0, c1, c2, zoff, f0, f1, f2, -1, R, r2q, lane, dirs);
          __syncwarp();
          if (st.cnt > RADIUS_CAP) {
            hi2 = r2q;
            continue;
          }
          if (st.cnt >= sp.k || (whole && !bisect)) {
            overflow = false;
            break;
          }
          lo2 = r2q;
          if (!bisect) {  // fewer than k inside the covered radius: widen the stencil (planar density guess)
            const double want = sqrt((double)cov * 1.6 * (double)sp.k / (double)max(st.cnt, 1)) * g.inv_cell;
            R = min(max(R + 1, (int)fmin(ceil(want), 1.0e6)), max_dim);
          }
        }
        n_nb = overflow ? 0 : st.cnt;
      } else if (MODE == 1) {
        scan_shell<MODE>(a, ws, st, Q, c0, c1, c2, zoff, f0, f1, f2, -1, sp.R0, sp.r2, lane, dirs);
        __syncwarp();
        n_nb = st.cnt;
        overflow = st.cnt > RADIUS_CAP;
        if (overflow && sp.nn_idx == nullptr) {
          // more hits than the buffer holds: unordered double-precision accumulation instead of a list
          overflow = false;
          heavy = true;
          if (n_nb >= sp.min_pts && (need_q || need_m)) {
            double sums[3], mom[6];
            const double inv_n = 1.0 / (double)n_nb;
            if (need_q) {
              heavy_moment(a, Q, c0, c1, c2, zoff, sp.R0, sp.r2, (double)Q.x, (double)Q.y, (double)Q.z, lane, sums, mom);
              if (lane < 6) ws->mom[q][lane] = (float)(mom[lane] * inv_n);
            }
            if (need_m) {
              if (!need_q) heavy_moment(a, Q, c0, c1, c2, zoff, sp.R0, sp.r2, 0.0, 0.0, 0.0, lane, sums, mom);
              const double mx = sums[0] * inv_n, my = sums[1] * inv_n, mz = sums[2] * inv_n;
              heavy_moment(a, Q, c0, c1, c2, zoff, sp.R0, sp.r2, mx, my, mz, lane, sums, mom);
              if (lane < 6) ws->momM[q][lane] = (float)(mom[lane] * inv_n);
            }
          }
        }
      } else {
        int Rp = -1, R = min(__shfl_sync(PCP_FULL, my_r0, q), max_dim);
        // seed: the tightest upper bound on this query's k-th distance from the solved queries of
        // the tile; with it the first pass already filters and prunes like a radius search
        float bound = 3.4e38f;
        {
          float bj = 3.4e38f;
          if (lane < q && my_dk < 3.0e38f && my_seg == (g.n_seg > 1 ? cid / g.seg_cells : 0u)) {
            float dx = Q.x - myP.x, dy = Q.y - myP.y, dz = Q.z - myP.z;
            bj = sqrtf(my_dk) + sqrtf(dx * dx + dy * dy + dz * dz);
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) bj = fminf(bj, __shfl_xor_sync(PCP_FULL, bj, o));
          // a seed is only worth taking from a nearby query (tiles cross row ends): ignore it when
          // it would widen the first stencil beyond 2 cells
          const float rs = bj * 1.00001f * (float)g.inv_cell + 1e-5f;
          if (rs <= 2.0f) {
            bound = bj * bj * 1.00002f;
            st.thr = ((u64)__float_as_uint(bound) << 32) | 0xFFFFFFFFull;
            R = min(max(R, (int)ceilf(rs)), max_dim);
          }
        }
#pragma unroll 1
        for (;;) {
          STAT(1, 1);
          scan_shell<MODE>(a, ws, st, Q, c0, c1, c2, zoff, f0, f1, f2, Rp, R, bound, lane, dirs);
          if (st.cnt > 0) {
            STAT(7, 1);
            u64 v = warp_sort32(lane < st.cnt ? ws->u.s.buf[lane] : PCP_KEY_INF, dirs);
            st.best = warp_merge32(st.best, v, lane, dirs);
            st.cnt = 0;
            __syncwarp();
          }
          const u64 kth = shfl_u64(st.best, sp.k - 1);
          if (kth < st.thr) st.thr = kth;
          const bool found = kth != PCP_KEY_INF;
          if (found) bound = __uint_as_float((uint32_t)(kth >> 32));
          const bool whole = (c0 - R <= 0) && (c0 + R >= g.dim[0] - 1) && (c1 - R <= 0) && (c1 + R >= g.dim[1] - 1) &&
                             (c2 - R <= 0) && (c2 + R >= g.dim[2] - 1);
          if (whole) break;
          int Rn;
          if (found) {
            const float dk = __uint_as_float((uint32_t)(kth >> 32));
            if (dk <= cover2_of(R, g.cell)) break;
            Rn = (int)fmin(ceil(sqrt((double)dk) * (1.0 + 2e-6) * g.inv_cell + 2e-6), 1.0e9);
            Rn = max(R + 1, Rn);
          } else {
            Rn = 2 * R;
          }
          Rp = R;
          R = min(Rn, max_dim);
        }
        n_nb = __popc(__ballot_sync(PCP_FULL, lane < sp.k && st.best != PCP_KEY_INF));
        const u64 kfin = shfl_u64(st.best, sp.k - 1);
        if (lane == q && kfin != PCP_KEY_INF) my_dk = __uint_as_float((uint32_t)(kfin >> 32));
        st.thr = kfin;
      }
    }

    if (MODE == 0) {
      // the ordered list (one key per lane) goes to column q; everything after the search runs per tile below
      if (lane < sp.k) {
        ws->list_lo[lane * KNN_LW + q] = (uint32_t)st.best;
        ws->list_hi[lane * KNN_LW + q] = (uint32_t)(st.best >> 32);
      }
      if (lane == 0) ws->cnt[q] = n_nb;
      continue;
    }
    // ---- ordered neighbour list -> outputs (radius mode)
    const int ord_q = __float_as_int(Q.w);
    const int gid_q = a.gid_of_ord ? __ldg(&a.gid_of_ord[ord_q]) : ord_q;
    const bool valid = !overflow && (MODE == 0 ? n_nb > 0 : n_nb >= sp.min_pts);
    u64* buf = ws->u.s.buf;
    if (MODE == 1 && !overflow && !heavy && n_nb > 0 && (valid || sp.nn_idx != nullptr)) {
      // sort the hits by (d2, index)
      if (n_nb <= 32) {
        st.best = warp_sort32(lane < n_nb ? buf[lane] : PCP_KEY_INF, dirs);
        __syncwarp();
        if (lane < n_nb) buf[lane] = st.best;
        __syncwarp();
      } else {
        int m2 = 64;
        while (m2 < n_nb) m2 <<= 1;
        for (int t = n_nb + lane; t < m2; t += 32) buf[t] = PCP_KEY_INF;
        __syncwarp();
        smem_sort_pow2(buf, m2, lane);
      }
    }
    if (MODE == 1 && sp.k > 0 && n_nb > sp.k) n_nb = sp.k;  // k by cover radius: the first k of the sorted hits
    if (MODE == 1 && overflow && sp.overflow_flag && lane == 0) atomicExch(sp.overflow_flag, 1);
    if (MODE == 1 && sp.counts && lane == 0) sp.counts[slot_q] = n_nb;
    if (MODE == 0 && sp.kth_d2 && lane == 0)
      sp.kth_d2[slot_q] = n_nb >= sp.k ? __uint_as_float((uint32_t)(st.thr >> 32)) : INFINITY;
    if (MODE == 1 && sp.k > 0 && sp.kth_d2 && lane == 0)
      sp.kth_d2[slot_q] = n_nb >= sp.k ? __uint_as_float((uint32_t)(buf[sp.k - 1] >> 32)) : INFINITY;
    if (sp.nn_idx != nullptr && !overflow) {
      if (MODE == 0) {
        if (lane < sp.k) {
          const int64_t o = (int64_t)slot_q * sp.k + lane;
          const bool has = lane < n_nb;
          const int ord = (int)(uint32_t)st.best;
          sp.nn_idx[o] = has ? (a.gid_of_ord ? __ldg(&a.gid_of_ord[ord]) : ord) : -1;
          if (sp.nn_d2) sp.nn_d2[o] = has ? __uint_as_float((uint32_t)(st.best >> 32)) : INFINITY;
        }
      } else if (sp.k > 0) {
        for (int t = lane; t < sp.k; t += 32) {
          const int64_t o = (int64_t)slot_q * sp.k + t;
          const bool has = t < n_nb;
          const u64 key = has ? buf[t] : 0ull;
          const int ord = (int)(uint32_t)key;
          sp.nn_idx[o] = has ? (a.gid_of_ord ? __ldg(&a.gid_of_ord[ord]) : ord) : -1;
          if (sp.nn_d2) sp.nn_d2[o] = has ? __uint_as_float((uint32_t)(key >> 32)) : INFINITY;
        }
      } else if (sp.offsets) {
        const int64_t o = sp.offsets[slot_q];
        for (int t = lane; t < n_nb; t += 32) {
          const u64 key = buf[t];
          const int ord = (int)(uint32_t)key;
          sp.nn_idx[o + t] = a.gid_of_ord ? __ldg(&a.gid_of_ord[ord]) : ord;
          if (sp.nn_d2) sp.nn_d2[o + t] = __uint_as_float((uint32_t)(key >> 32));
        }
      }
    }
    if (valid && !heavy && (need_q || need_m)) {
      if (need_q) {
        float c = ordered_moment<MODE == 1>(sorted, a.inv, buf, st.best, n_nb, Q.x, Q.y, Q.z, ws->u.s.prod, lane);
        if (lane < 6) ws->mom[q][lane] = c;
      }
      if (need_m) {
        float mx, my, mz;
        ordered_centroid<MODE == 1>(sorted, a.inv, buf, st.best, n_nb, ws->u.s.prod, lane, mx, my, mz);
        float c = ordered_moment<MODE == 1>(sorted, a.inv, buf, st.best, n_nb, mx, my, mz, ws->u.s.prod, lane);
        if (lane < 6) ws->momM[q][lane] = c;
      }
    }
    if (lane == 0) {
      ws->cnt[q] = valid ? n_nb : -n_nb - 1;  // negative = no PCA for this record
      ws->gidq[q] = gid_q;
    }
    __syncwarp();
  }
  __syncwarp();

  if (MODE == 0) {
    const int n_mine = my_slot >= 0 ? ws->cnt[lane] : 0;
    knn_finish(a, ws->list_lo + lane, ws->list_hi + lane, KNN_LW, n_mine, 1, myP, my_slot, ws->u.rec, lane);
    return;
  }
  // ---- lane-parallel eigen + features: lane l finishes query l of the tile
  if (my_slot >= 0) {
    const int c = ws->cnt[lane];
    const bool valid = c >= 0;
    const int n_nb = valid ? c : -c - 1;
    float* rec = ws->u.rec[lane];
#pragma unroll
    for (int w = 0; w < REC_WORDS; ++w) rec[w] = 0.f;
    rec[12] = __int_as_float(ws->gidq[lane]);
    rec[13] = __int_as_float(n_nb);
    Sym3Eig eq;
    bool have_q = false;
    if (valid && sp.out != nullptr) {
      sym3_eigen(sp.center_mode == PCP_CENTER_QUERY ? ws->mom[lane] : ws->momM[lane], eq);
      have_q = true;
      rec[0] = eq.val[0]; rec[1] = eq.val[1]; rec[2] = eq.val[2];
#pragma unroll
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int r = 0; r < 3; ++r) rec[3 + 3 * k + r] = eq.vec[k][r];
      // src/pcaAnalysis.cpp:29-31 in fp32 on the ascending eigenvalues
      float s1 = __fsqrt_rn(eq.val[0]), s2 = __fsqrt_rn(eq.val[1]), s3 = __fsqrt_rn(eq.val[2]);
      rec[14] = __fdiv_rn(__fsub_rn(s1, s2), s1);
      rec[15] = __fdiv_rn(__fsub_rn(s2, s3), s1);
      rec[16] = __fdiv_rn(s3, s1);
    }
    if (sp.normals != nullptr) {
      pcp_normal16 nrm;
      if (!valid || n_nb < 3) {
        nrm.nx = nrm.ny = nrm.nz = nrm.curvature = __int_as_float(0x7FC00000);
      } else {
        Sym3Eig em;
        if (have_q && sp.center_mode == PCP_CENTER_MEAN) em = eq;
        else sym3_eigen(ws->momM[lane], em);
        const float* M = ws->momM[lane];
        float tr = M[0] + M[3] + M[5];
        float4 P = myP;
        float nx = em.vec[0][0], ny = em.vec[0][1], nz = em.vec[0][2];
        // flipNormalTowardsViewpoint with viewpoint (0,0,0)
        float cs = (0.f - P.x) * nx + (0.f - P.y) * ny + (0.f - P.z) * nz;
        if (cs < 0.f) { nx = -nx; ny = -ny; nz = -nz; }
        nrm.nx = nx; nrm.ny = ny; nrm.nz = nz;
        nrm.curvature = tr != 0.f ? fabsf(em.val[0] / tr) : 0.f;
      }
      sp.normals[my_slot] = nrm;
    }
    if (sp.valid) sp.valid[my_slot] = valid ? 1 : 0;
  }
  __syncwarp();
  // ---- 68-byte records: lanes 0..16 write one record per step (contiguous 68 B)
  if (sp.out != nullptr) {
    float* outw = reinterpret_cast<float*>(sp.out);
    for (int r = 0; r < nq; ++r) {
      int s = __shfl_sync(PCP_FULL, my_slot, r);
      if (s >= 0 && lane < REC_WORDS) outw[(int64_t)s * REC_WORDS + lane] = ws->u.rec[r][lane];
    }
  }
}

// every point of the cloud: tile t = sorted positions [32 t, 32 t + 32)
template <int MODE>
__global__ void __launch_bounds__(WARPS * 32, SP_MINB) search_pca_kernel(KArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpSmem<MODE>* ws = reinterpret_cast<WarpSmem<MODE>*>(smem_raw) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int64_t pos0 = ((int64_t)blockIdx.x * WARPS + (threadIdx.x >> 5)) * 32;
  if (pos0 >= a.n) return;
  if (MODE == 0) {  // k-NN tiles hold KNN_LW queries
    for (int h = 0; h < 32; h += KNN_LW) {
      if (pos0 + h >= a.n) break;
      search_tile<MODE>(a, ws, lane, pos0 + h + lane, (int)min((int64_t)KNN_LW, a.n - pos0 - h), a.sp.R0);
      __syncwarp();
    }
  } else {
    search_tile<MODE>(a, ws, lane, pos0 + lane, (int)min((int64_t)32, a.n - pos0), a.sp.R0);
  }
}

// the queries a faster kernel deferred: work_list[0 .. work_count[0]) are sorted positions.  Warps draw tiles of
// `tile` list entries from a ticket counter (work_count[1]) until the list is empty — the count is only known on
// the device, and drawing tickets keeps every resident warp busy to the end whatever the list length.
template <int MODE>
__global__ void __launch_bounds__(WARPS * 32, SP_MINB) search_pca_worklist_kernel(KArgs a, uint32_t* work_count, const uint32_t* __restrict__ work_list,
                                                                         int tile) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpSmem<MODE>* ws = reinterpret_cast<WarpSmem<MODE>*>(smem_raw) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int64_t n_items = (int64_t)__ldg(work_count);
#pragma unroll 1
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) t = atomicAdd(work_count + 1, 1u);
    t = __shfl_sync(PCP_FULL, t, 0);
    const int64_t i0 = (int64_t)t * tile;
    if (i0 >= n_items) break;
    const int nq = (int)min((int64_t)tile, n_items - i0);
    const uint32_t entry = lane < nq ? work_list[i0 + lane] : 0u;  // bit 31: "not covered by the half-width-1 block"
    const int64_t my_pos = (int64_t)(entry & 0x7FFFFFFFu);
    search_tile<MODE>(a, ws, lane, my_pos, nq, (MODE == 0 && (entry >> 31)) ? max(2, a.sp.R0) : a.sp.R0);
    __syncwarp();
  }
}

}  // namespace

#ifdef PCP_STATS
extern "C" __attribute__((visibility("default"))) int pcp_debug_stats(unsigned long long* out) {
  unsigned long long z[8] = {0};
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, g_stats, sizeof(z));
  cudaMemcpyToSymbol(g_stats, z, sizeof(z));
  return 0;
}
#endif

// the kernel arguments every search kernel takes: the grid of the context's current cloud + the call's parameters
void pcp_make_kargs(pcp_ctx* ctx, const SearchParams& sp, KArgs& a) {
  a.sorted = (const float4*)ctx->sorted_pts.p;
  a.keys = (const uint32_t*)ctx->keys_sorted.p;
  a.slot = (const int32_t*)ctx->sorted_slot.p;
  a.E = (const uint32_t*)ctx->cell_end.p;
  a.inv = (const uint32_t*)ctx->inv_pos.p;
  a.gid_of_ord = (ctx->d_gid && !ctx->gid_direct) ? (const int32_t*)ctx->gid_sorted.p : nullptr;
  a.g = ctx->grid;
  a.n = ctx->n;
  a.sp = sp;
}

int pcp_search_launch(pcp_ctx* ctx, const SearchParams& sp) {
  // PCP_KNN_IMPL (read once in pcp_create) selects the k-NN kernel for A/B timing: default = tile kernel of
  // knn_tile6.cu + this file's warp-per-query kernel for the queries it defers; "tile4" = the round-1 tile
  // kernel (knn_tile.cu) instead; "v3" = warp-per-query for every point.  All give identical results.
  const PcpKnobs& kn = ctx->knobs;
  const bool use_v3 = sp.mode == 0 && kn.knn_impl == 3;
  KArgs a;
  pcp_make_kargs(ctx, sp, a);
  const int64_t tiles = (ctx->n + 31) / 32;
  const int blocks = (int)((tiles + WARPS - 1) / WARPS);
  pcp_tic(ctx, PCP_T_SEARCH_PCA);
  // radius + PCA records (no neighbour lists requested) also starts in the tile kernel when the
  // neighbourhoods are expected to be small: lanes hold up to 32 neighbours, larger ones go to the
  // warp-per-query kernel.  Measured on B200: mean 8.7 neighbours -> 0.82 vs 0.99 ms per 1 M points, mean 32 ->
  // 3.2 vs 2.4 ms, so the switch is a surface-density guess n*pi*r^2 / (two largest extents) <= 40
  // (the urban clouds hold ~0.4 of that guess).  PCP_RADIUS_IMPL=tile|v3 forces either.
  bool radius_tile = false;
  int radius_cap = 32;
  if (sp.mode == 1 && sp.k == 0 && sp.out != nullptr && sp.counts == nullptr && sp.nn_idx == nullptr) {
    const double area = (double)ctx->grid.dim[2] * (double)ctx->grid.dim[0] * (double)ctx->grid.cell * (double)ctx->grid.cell;
    const double guess = (double)ctx->n / (double)ctx->grid.n_seg * 3.14159265 * (double)sp.r2 / (area > 0.0 ? area : 1.0);
    // lanes of 32 neighbours for sparse neighbourhoods, of 64 for dense ones (configs[3]: ~32 on average),
    // the warp-per-query kernel alone beyond that
    radius_tile = guess <= 200.0;
    radius_cap = guess <= 40.0 ? 32 : 64;
    if (kn.radius_impl == 1 || kn.radius_impl == 2) { radius_tile = true; radius_cap = kn.radius_impl == 2 ? 64 : 32; }
    if (kn.radius_impl == 3) radius_tile = false;
    // the tile kernel's run tables hold (2R+1)^2 <= 49 stencil rows: a cell smaller than r/3 (R0 > 3: automatic
    // k-NN cells, or a small cell chosen by the caller) goes to the warp-per-query kernel, which has no such bound
    if (sp.R0 > 3) radius_tile = false;
  }
  if ((sp.mode == 0 && !use_v3) || radius_tile) {
    // k-NN: knn_tile6.cu for single clouds; frame batches (segmented grid: short rows, up to four row groups per tile,
    // k = 10) stay on the round-1 tile kernel, which re-stages less per row group (measured, 128 VLP-16 frames: 3.7 vs 4.1 ms)
    const bool use_t6 = sp.mode == 0 && (kn.knn_impl == 6 || (kn.knn_impl == 0 && ctx->grid.n_seg == 1));  // PCP_KNN_IMPL=tile6 forces it
    // fixed-radius tiles: knn_tile6.cu's radius mode for single clouds (PCP_RADIUS_T6=0 keeps the round-1 kernel)
    const bool rad_t6 = sp.mode == 1 && (kn.radius_t6 == 1 || (kn.radius_t6 < 0 && PCP_RADIUS_T6_DEFAULT && ctx->grid.n_seg == 1));
    int rc = (use_t6 || rad_t6) ? pcp_knn_tile6_launch(ctx, sp, a, radius_cap) : pcp_knn_tile_launch(ctx, sp, a, radius_cap);
    if (rc) return rc;
    // one wave of resident CTAs (3 per SM at 80 registers); tiles of 16 list entries drawn by ticket
    // (tuning hooks: PCP_WL_BLOCKS, PCP_WL_TILE)
    int wl_blocks = blocks < 148 * SP_MINB ? blocks : 148 * SP_MINB;
    int wl_tile = 16;
    if (kn.wl_blocks > 0) wl_blocks = kn.wl_blocks;
    if (kn.wl_tile > 0) wl_tile = kn.wl_tile;
    if (sp.mode == 0 && wl_tile > KNN_LW) wl_tile = KNN_LW;
    if (sp.mode == 0) {
      size_t smem = sizeof(WarpSmem<0>) * WARPS;
      PCP_CUDA(cudaFuncSetAttribute(search_pca_worklist_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      PCP_CUDA(pcp_launch_search(ctx, search_pca_worklist_kernel<0>, dim3(wl_blocks), dim3(WARPS * 32), smem, a, (uint32_t*)ctx->work_count.p,
                                 (const uint32_t*)ctx->work_list.p, wl_tile));
    } else {
      size_t smem = sizeof(WarpSmem<1>) * WARPS;
      PCP_CUDA(cudaFuncSetAttribute(search_pca_worklist_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      PCP_CUDA(pcp_launch_search(ctx, search_pca_worklist_kernel<1>, dim3(wl_blocks), dim3(WARPS * 32), smem, a, (uint32_t*)ctx->work_count.p,
                                 (const uint32_t*)ctx->work_list.p, wl_tile));
    }
  } else if (sp.mode == 0) {
    size_t smem = sizeof(WarpSmem<0>) * WARPS;
    PCP_CUDA(cudaFuncSetAttribute(search_pca_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PCP_CUDA(pcp_launch_search(ctx, search_pca_kernel<0>, dim3(blocks), dim3(WARPS * 32), smem, a));
  } else {
    size_t smem = sizeof(WarpSmem<1>) * WARPS;
    PCP_CUDA(cudaFuncSetAttribute(search_pca_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    PCP_CUDA(pcp_launch_search(ctx, search_pca_kernel<1>, dim3(blocks), dim3(WARPS * 32), smem, a));
  }
  PCP_LAUNCH_CHECK("search_pca_kernel");
  pcp_toc(ctx, PCP_T_SEARCH_PCA);
  return PCP_OK;
}
