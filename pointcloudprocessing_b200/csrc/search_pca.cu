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

#include "eigen3.cuh"
#include "pcp_internal.cuh"
#include "search_common.cuh"

namespace {

constexpr int WARPS = 8;            // warps per CTA
constexpr int RADIUS_CAP = 512;     // radius mode: neighbours kept per query in shared memory
constexpr int REC_WORDS = 17;       // 68-byte record

struct KArgs {
  const float4* sorted;
  const uint32_t* keys;
  const int32_t* slot;
  const uint32_t* E;
  GridDesc g;
  int64_t n;
  SearchParams sp;
};

// shared memory of one warp
template <int MODE>
struct WarpSmem {
  u64 buf[MODE == 0 ? 64 : RADIUS_CAP];
  float prod[6][32];
  float mom[32][6];    // query-centred moment (normalised)
  float momM[32][6];   // mean-centred moment (normalised), only when needed
  float rec[32][REC_WORDS];
  int cnt[32];
  int gidq[32];
};

__device__ __forceinline__ float cover2_of(int R, float cell) {
  double rs = ((double)R - 1e-6) * (double)cell;
  return (float)(rs * rs * (1.0 - 1e-6));
}

// Ordered second moment about (cx,cy,cz) of the m neighbours whose keys sit in list[0..m)
// (registers for m <= 32 via `regkey`, shared memory otherwise).  Lanes 0..5 return
// c00,c01,c02,c11,c12,c22 (already divided by m); accumulation order is PCL's.
template <bool FROM_SMEM>
__device__ __forceinline__ float ordered_moment(const float4* __restrict__ sorted, const u64* list, u64 regkey, int m,
                                                float cx, float cy, float cz, float (*prod)[32], int lane) {
  float acc = 0.f;
  for (int base = 0; base < m; base += 32) {
    int j = base + lane;
    if (j < m) {
      u64 key = FROM_SMEM ? list[j] : regkey;
      float4 p = __ldg(&sorted[(uint32_t)key]);
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
      for (int t = 0; t < cnt; ++t) acc = __fadd_rn(acc, prod[lane][t]);
    }
    __syncwarp();
  }
  if (m > 0) acc = __fdiv_rn(acc, (float)m);
  return acc;
}

// fp32 running-sum centroid in list order (pcl::compute3DCentroid); result on every lane
template <bool FROM_SMEM>
__device__ __forceinline__ void ordered_centroid(const float4* __restrict__ sorted, const u64* list, u64 regkey, int m,
                                                 float (*prod)[32], int lane, float& cx, float& cy, float& cz) {
  float acc = 0.f;
  for (int base = 0; base < m; base += 32) {
    int j = base + lane;
    if (j < m) {
      u64 key = FROM_SMEM ? list[j] : regkey;
      float4 p = __ldg(&sorted[(uint32_t)key]);
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

template <int MODE>
__global__ void __launch_bounds__(WARPS * 32) search_pca_kernel(KArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpSmem<MODE>* ws = reinterpret_cast<WarpSmem<MODE>*>(smem_raw) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int64_t tile = (int64_t)blockIdx.x * WARPS + (threadIdx.x >> 5);
  const int64_t pos0 = tile * 32;
  if (pos0 >= a.n) return;
  const float4* __restrict__ sorted = a.sorted;
  const GridDesc& g = a.g;
  const SearchParams& sp = a.sp;
  const bool need_q = sp.out != nullptr && sp.center_mode == PCP_CENTER_QUERY;
  const bool need_m = sp.normals != nullptr || (sp.out != nullptr && sp.center_mode == PCP_CENTER_MEAN);
  const int nq = (int)min((int64_t)32, a.n - pos0);
  const int my_slot = lane < nq ? a.slot[pos0 + lane] : -1;
  const unsigned lt_mask = (1u << lane) - 1u;

  for (int q = 0; q < nq; ++q) {
    const int slot_q = __shfl_sync(PCP_FULL, my_slot, q);
    if (slot_q < 0) continue;  // halo point: neighbour only
    const int64_t pq = pos0 + q;
    const float4 Q = __ldg(&sorted[pq]);
    const uint32_t cid = __ldg(&a.keys[pq]);
    int n_nb = 0;          // neighbour count of this query
    u64 best = PCP_KEY_INF;  // k-NN: sorted best list, one per lane
    bool overflow = false;
    if (cid < g.n_cells) {
      const int c0 = (int)(cid % (uint32_t)g.dim[0]);
      const uint32_t tq = cid / (uint32_t)g.dim[0];
      const int c1 = (int)(tq % (uint32_t)g.dim[1]);
      const int c2 = (int)(tq / (uint32_t)g.dim[1]);
      int R = sp.R0;
      for (;;) {
        best = PCP_KEY_INF;
        u64 thr = PCP_KEY_INF;
        uint32_t thr_hi = 0xFFFFFFFFu;
        int cnt = 0;  // entries in ws->buf
        // stencil rows (axes 1 and 2), clipped to the grid; enumeration starts at the query's own row
        const int lo1 = max(c1 - R, 0), hi1 = min(c1 + R, g.dim[1] - 1), n1 = hi1 - lo1 + 1;
        const int lo2 = max(c2 - R, 0), hi2 = min(c2 + R, g.dim[2] - 1), n2 = hi2 - lo2 + 1;
        const int nrows = n1 * n2;
        const int lo0 = max(c0 - R, 0), hi0 = min(c0 + R, g.dim[0] - 1);
        for (int rb = 0; rb < nrows; rb += 32) {
          const int r = rb + lane;
          uint32_t rbeg = 0, rend = 0;
          if (r < nrows) {
            int o1 = r % n1, o2 = r / n1;
            int e1 = lo1 + (c1 - lo1 + o1) % n1, e2 = lo2 + (c2 - lo2 + o2) % n2;
            uint32_t base = ((uint32_t)e2 * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
            rbeg = __ldg(&a.E[base + lo0]);
            rend = __ldg(&a.E[base + hi0 + 1]);
          }
          const int nr = min(32, nrows - rb);
          for (int j = 0; j < nr; ++j) {
            const uint32_t bj = __shfl_sync(PCP_FULL, rbeg, j), ej = __shfl_sync(PCP_FULL, rend, j);
            for (uint32_t cb = bj; cb < ej; cb += 32) {
              const uint32_t cp = cb + lane;
              bool pass = false;
              u64 key = 0;
              if (cp < ej) {
                float4 P = __ldg(&sorted[cp]);
                float d2 = dist2_nofma(Q.x, Q.y, Q.z, P);
                uint32_t d2b = __float_as_uint(d2);
                key = ((u64)d2b << 32) | cp;
                if (MODE == 0) pass = d2b < thr_hi || (d2b == thr_hi && key_less(key, thr, sorted));
                else pass = d2 < sp.r2;
              }
              const unsigned m = __ballot_sync(PCP_FULL, pass);
              if (m == 0) continue;
              const int off = cnt + __popc(m & lt_mask);
              if (MODE == 0) {
                if (pass) ws->buf[off] = key;
                cnt += __popc(m);
                __syncwarp();
                if (cnt >= 32) {
                  u64 v = warp_sort32(ws->buf[lane], lane, sorted);
                  best = warp_merge32(best, v, lane, sorted);
                  thr = shfl_u64(best, sp.k - 1);
                  thr_hi = (uint32_t)(thr >> 32);
                  u64 rem = (lane + 32 < cnt) ? ws->buf[lane + 32] : 0ull;
                  __syncwarp();
                  if (lane + 32 < cnt) ws->buf[lane] = rem;
                  cnt -= 32;
                  __syncwarp();
                }
              } else {
                if (pass && off < RADIUS_CAP) ws->buf[off] = key;
                cnt += __popc(m);
              }
            }
          }
        }
        if (MODE == 1) {
          __syncwarp();
          n_nb = cnt;
          overflow = cnt > RADIUS_CAP;
          break;
        }
        if (cnt > 0) {
          u64 v = warp_sort32(lane < cnt ? ws->buf[lane] : PCP_KEY_INF, lane, sorted);
          best = warp_merge32(best, v, lane, sorted);
          __syncwarp();
        }
        const u64 kth = shfl_u64(best, sp.k - 1);
        const bool found = kth != PCP_KEY_INF;
        const bool whole = (c0 - R <= 0) && (c0 + R >= g.dim[0] - 1) && (c1 - R <= 0) && (c1 + R >= g.dim[1] - 1) &&
                           (c2 - R <= 0) && (c2 + R >= g.dim[2] - 1);
        if (whole) break;
        if (found) {
          const float dk = __uint_as_float((uint32_t)(kth >> 32));
          if (dk <= cover2_of(R, g.cell)) break;
          int Rn = (int)fmin(ceil(sqrt((double)dk) * (1.0 + 2e-6) * g.inv_cell + 2e-6), 1.0e9);
          R = max(R + 1, Rn);
        } else {
          R = 2 * R;
        }
        R = min(R, max(g.dim[0], max(g.dim[1], g.dim[2])));
      }
      if (MODE == 0) n_nb = __popc(__ballot_sync(PCP_FULL, lane < sp.k && best != PCP_KEY_INF));
    }

    // ---- ordered neighbour list -> outputs
    const int gid_q = __float_as_int(Q.w);
    bool valid = !overflow && (MODE == 0 ? n_nb > 0 : n_nb >= sp.min_pts);
    if (MODE == 1 && !overflow && n_nb > 0 && (valid || sp.nn_idx != nullptr)) {
      // sort the hits by (d2, index)
      if (n_nb <= 32) {
        best = warp_sort32(lane < n_nb ? ws->buf[lane] : PCP_KEY_INF, lane, sorted);
        __syncwarp();
        if (lane < n_nb) ws->buf[lane] = best;
        __syncwarp();
      } else {
        int m2 = 64;
        while (m2 < n_nb) m2 <<= 1;
        for (int t = n_nb + lane; t < m2; t += 32) ws->buf[t] = PCP_KEY_INF;
        __syncwarp();
        smem_sort_pow2(ws->buf, m2, lane, sorted);
      }
    }
    if (MODE == 1 && overflow && sp.overflow_flag && lane == 0) atomicExch(sp.overflow_flag, 1);
    if (MODE == 1 && sp.counts && lane == 0) sp.counts[slot_q] = n_nb;
    if (sp.nn_idx != nullptr && !overflow) {
      if (MODE == 0) {
        if (lane < sp.k) {
          int64_t o = (int64_t)slot_q * sp.k + lane;
          bool has = lane < n_nb;
          sp.nn_idx[o] = has ? __float_as_int(__ldg(&sorted[(uint32_t)best].w)) : -1;
          if (sp.nn_d2) sp.nn_d2[o] = has ? __uint_as_float((uint32_t)(best >> 32)) : INFINITY;
        }
      } else if (sp.offsets) {
        int64_t o = sp.offsets[slot_q];
        for (int t = lane; t < n_nb; t += 32) {
          u64 key = ws->buf[t];
          sp.nn_idx[o + t] = __float_as_int(__ldg(&sorted[(uint32_t)key].w));
          if (sp.nn_d2) sp.nn_d2[o + t] = __uint_as_float((uint32_t)(key >> 32));
        }
      }
    }
    if (valid && (need_q || need_m)) {
      if (need_q) {
        float c = ordered_moment<MODE == 1>(sorted, ws->buf, best, n_nb, Q.x, Q.y, Q.z, ws->prod, lane);
        if (lane < 6) ws->mom[q][lane] = c;
      }
      if (need_m) {
        float mx, my, mz;
        ordered_centroid<MODE == 1>(sorted, ws->buf, best, n_nb, ws->prod, lane, mx, my, mz);
        float c = ordered_moment<MODE == 1>(sorted, ws->buf, best, n_nb, mx, my, mz, ws->prod, lane);
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

  // ---- lane-parallel eigen + features: lane l finishes query l of the tile
  if (my_slot >= 0) {
    const int c = ws->cnt[lane];
    const bool valid = c >= 0;
    const int n_nb = valid ? c : -c - 1;
    float* rec = ws->rec[lane];
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
        float4 P = __ldg(&sorted[pos0 + lane]);
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
      if (s >= 0 && lane < REC_WORDS) outw[(int64_t)s * REC_WORDS + lane] = ws->rec[r][lane];
    }
  }
}

}  // namespace

int pcp_search_launch(pcp_ctx* ctx, const SearchParams& sp) {
  KArgs a;
  a.sorted = (const float4*)ctx->sorted_pts.p;
  a.keys = (const uint32_t*)ctx->keys_sorted.p;
  a.slot = (const int32_t*)ctx->sorted_slot.p;
  a.E = (const uint32_t*)ctx->cell_end.p;
  a.g = ctx->grid;
  a.n = ctx->n;
  a.sp = sp;
  const int64_t tiles = (ctx->n + 31) / 32;
  const int blocks = (int)((tiles + WARPS - 1) / WARPS);
  pcp_tic(ctx, PCP_T_SEARCH_PCA);
  if (sp.mode == 0) {
    size_t smem = sizeof(WarpSmem<0>) * WARPS;
    PCP_CUDA(cudaFuncSetAttribute(search_pca_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    search_pca_kernel<0><<<blocks, WARPS * 32, smem, ctx->stream>>>(a);
  } else {
    size_t smem = sizeof(WarpSmem<1>) * WARPS;
    PCP_CUDA(cudaFuncSetAttribute(search_pca_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    search_pca_kernel<1><<<blocks, WARPS * 32, smem, ctx->stream>>>(a);
  }
  PCP_LAUNCH_CHECK("search_pca_kernel");
  pcp_toc(ctx, PCP_T_SEARCH_PCA);
  return PCP_OK;
}
