// knn_finish.cuh — what every k-NN kernel does once a lane holds its query's ordered neighbour list:
// optional list / k-th distance outputs, the second moment in the reference's accumulation order
// (pcl::computeCovarianceMatrixNormalized), the eigen-solve (eigen3.cuh), the structure
// features (src/pcaAnalysis.cpp:29-31), the PCL-twin normal + curvature (src/smoothing.cpp:236-248),
// and the coalesced 68-byte record write.  Must be called by all 32 lanes of the warp.
#pragma once
#include "eigen3.cuh"
#include "search_common.cuh"

// Second half of every finish stage: eigen-solve of the moment(s), structure features, PCL-twin normal, record staging
// and the coalesced 68-byte record write.  mq / mm = query- / mean-centred moment {c00, c01, c02, c11, c12, c22}
// (already divided by n); `valid` = the PCA is to be computed.  Must be called by all 32 lanes of the warp.
// LEAN = the caller guarantees: records wanted, query-centred, no normals, no validity bytes, no lists (the k-NN hot
// path): the code of the optional outputs is not generated (the tile kernel's straight-line part is bound by instruction fetch).
template <bool LEAN = false>
__device__ __forceinline__ void knn_emit(const KArgs& a, const float (&mq)[6], const float (&mm)[6], const bool valid, const int n_nb,
                                         const float4 Q, const int my_slot, const int gid_q, float (*rec_all)[17], const int lane) {
  const SearchParams& sp = a.sp;
  // ---- eigen + features, staged as 68-byte records (the search columns are dead now)
  __syncwarp();
  if (my_slot >= 0) {
    float* rec = rec_all[lane];
#pragma unroll
    for (int w = 0; w < 17; ++w) rec[w] = 0.f;
    rec[12] = __int_as_float(gid_q);
    rec[13] = __int_as_float(n_nb);
    // one call site of the (noinline) solver on the hot path: the matrix the records need, which is also
    // the normals' matrix unless query-centred records AND normals were requested
    const bool want_out = valid && (LEAN || sp.out != nullptr);
    const bool want_nrm = !LEAN && sp.normals != nullptr && valid && n_nb >= 3;
    const bool first_is_mean = !LEAN && (!want_out || sp.center_mode == PCP_CENTER_MEAN);
    Sym3Eig e1;
    if (want_out || want_nrm) sym3_eigen_call(first_is_mean ? mm : mq, e1);
    if (want_out) {
      rec[0] = e1.val[0]; rec[1] = e1.val[1]; rec[2] = e1.val[2];
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int r = 0; r < 3; ++r) rec[3 + 3 * c + r] = e1.vec[c][r];
      // src/pcaAnalysis.cpp:29-31 in fp32 on the ascending eigenvalues
      const float s1 = __fsqrt_rn(e1.val[0]), s2 = __fsqrt_rn(e1.val[1]), s3 = __fsqrt_rn(e1.val[2]);
      rec[14] = __fdiv_rn(__fsub_rn(s1, s2), s1);
      rec[15] = __fdiv_rn(__fsub_rn(s2, s3), s1);
      rec[16] = __fdiv_rn(s3, s1);
    }
    if (!LEAN && sp.normals != nullptr) {
      pcp_normal16 nrm;
      if (!want_nrm) {
        nrm.nx = nrm.ny = nrm.nz = nrm.curvature = __int_as_float(0x7FC00000);
      } else {
        if (!first_is_mean) sym3_eigen_call(mm, e1);  // records were query-centred: solve the mean-centred moment too
        const float tr = mm[0] + mm[3] + mm[5];
        float nx = e1.vec[0][0], ny = e1.vec[0][1], nz = e1.vec[0][2];
        // flipNormalTowardsViewpoint with viewpoint (0,0,0)
        const float cs = (0.f - Q.x) * nx + (0.f - Q.y) * ny + (0.f - Q.z) * nz;
        if (cs < 0.f) { nx = -nx; ny = -ny; nz = -nz; }
        nrm.nx = nx; nrm.ny = ny; nrm.nz = nz;
        nrm.curvature = tr != 0.f ? fabsf(e1.val[0] / tr) : 0.f;
      }
      sp.normals[my_slot] = nrm;
    }
    if (!LEAN && sp.valid) sp.valid[my_slot] = valid ? 1 : 0;
  }
  __syncwarp();
  // ---- lanes 0..16 write one contiguous 68-byte record per step
  if (LEAN || sp.out != nullptr) {
    float* outw = reinterpret_cast<float*>(sp.out);
#pragma unroll 1
    for (int r = 0; r < 32; ++r) {
      const int s = __shfl_sync(PCP_FULL, my_slot, r);
      if (s >= 0 && lane < 17) outw[(int64_t)s * 17 + lane] = rec_all[r][lane];
    }
  }
}

// lo[j * stride] / hi[j * stride] (j < n_nb) = the two words of neighbour j's key bits(d2) << 32 | ord, ascending
// by (d2, index), in the lane's OWN shared-memory column (every loop over the list is rolled: code size,
// not instruction count, is what this stage costs the kernels that inline it); the column is overwritten.
// my_slot < 0: the lane has no query; min_valid = fewest neighbours for which the PCA is computed (1 for
// k-NN, the reference's 13 for radius search).  rec_all = 32 x 17 floats of warp-private shared memory;
// it may overlap the columns: the records are staged after the last read of the list.
__device__ __forceinline__ void knn_finish(const KArgs& a, uint32_t* lo, const uint32_t* hi, const int stride, const int n_nb,
                                           const int min_valid, const float4 Q, const int my_slot, float (*rec_all)[17],
                                           const int lane) {
  const SearchParams& sp = a.sp;
  const int k = sp.k;
  // ---- neighbour list -> outputs (each lane for its own query)
  const float4* __restrict__ sorted = a.sorted;
  const uint32_t* __restrict__ inv = a.inv;
  const int ord_q = __float_as_int(Q.w);
  const int gid_q = (my_slot >= 0 && a.gid_of_ord) ? __ldg(&a.gid_of_ord[ord_q]) : ord_q;
  const bool valid = my_slot >= 0 && n_nb >= min_valid;  // radius search: the 13-neighbour gate (src/pcaAnalysis.cpp:50)
  if (my_slot >= 0) {
    if (sp.kth_d2) sp.kth_d2[my_slot] = n_nb >= k ? __uint_as_float(hi[(k - 1) * stride]) : INFINITY;
    if (sp.nn_idx != nullptr) {
      const int64_t o = (int64_t)my_slot * k;
#pragma unroll 1
      for (int j = 0; j < k; ++j) {
        const bool has = j < n_nb;
        const int ord = has ? (int)lo[j * stride] : 0;
        sp.nn_idx[o + j] = has ? (a.gid_of_ord ? __ldg(&a.gid_of_ord[ord]) : ord) : -1;
        if (sp.nn_d2) sp.nn_d2[o + j] = has ? __uint_as_float(hi[j * stride]) : INFINITY;
      }
    }
  }
  const bool need_q = sp.out != nullptr && sp.center_mode == PCP_CENTER_QUERY;
  const bool need_m = sp.normals != nullptr || (sp.out != nullptr && sp.center_mode == PCP_CENTER_MEAN);
  float mq[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, mm[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (valid && (need_q || need_m)) {
    // ord -> sorted position, in place (the moment re-reads coordinates from them); eight loads in flight
#pragma unroll 8
    for (int j = 0; j < n_nb; ++j) lo[j * stride] = __ldg(&inv[lo[j * stride]]);
    const float fn = (float)n_nb;
    if (need_q) {
      // pcl::computeCovarianceMatrixNormalized about the query: six fp32 chains in list order
#pragma unroll 8
      for (int j = 0; j < n_nb; ++j) {
        const float4 P = __ldg(&sorted[lo[j * stride]]);
        const float px = __fsub_rn(P.x, Q.x), py = __fsub_rn(P.y, Q.y), pz = __fsub_rn(P.z, Q.z);
        mq[0] = __fadd_rn(mq[0], __fmul_rn(px, px));
        mq[1] = __fadd_rn(mq[1], __fmul_rn(py, px));
        mq[2] = __fadd_rn(mq[2], __fmul_rn(pz, px));
        mq[3] = __fadd_rn(mq[3], __fmul_rn(py, py));
        mq[4] = __fadd_rn(mq[4], __fmul_rn(py, pz));
        mq[5] = __fadd_rn(mq[5], __fmul_rn(pz, pz));
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) mq[c] = __fdiv_rn(mq[c], fn);
    }
    if (need_m) {
      // pcl::compute3DCentroid (fp32 running sums in list order), then the moment about it
      float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll 8
      for (int j = 0; j < n_nb; ++j) {
        const float4 P = __ldg(&sorted[lo[j * stride]]);
        sx = __fadd_rn(sx, P.x);
        sy = __fadd_rn(sy, P.y);
        sz = __fadd_rn(sz, P.z);
      }
      const float cx = __fdiv_rn(sx, fn), cy = __fdiv_rn(sy, fn), cz = __fdiv_rn(sz, fn);
#pragma unroll 8
      for (int j = 0; j < n_nb; ++j) {
        const float4 P = __ldg(&sorted[lo[j * stride]]);
        const float px = __fsub_rn(P.x, cx), py = __fsub_rn(P.y, cy), pz = __fsub_rn(P.z, cz);
        mm[0] = __fadd_rn(mm[0], __fmul_rn(px, px));
        mm[1] = __fadd_rn(mm[1], __fmul_rn(py, px));
        mm[2] = __fadd_rn(mm[2], __fmul_rn(pz, px));
        mm[3] = __fadd_rn(mm[3], __fmul_rn(py, py));
        mm[4] = __fadd_rn(mm[4], __fmul_rn(py, pz));
        mm[5] = __fadd_rn(mm[5], __fmul_rn(pz, pz));
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) mm[c] = __fdiv_rn(mm[c], fn);
    }
  }

  knn_emit(a, mq, mm, valid, n_nb, Q, my_slot, gid_q, rec_all, lane);
}

// ---- finish from a column of sorted POSITIONS (knn_tile6.cu) -----------------------------------------------------------
// col[j * 32], j < cnt = the members' positions in the cell-sorted array, ordered by the PREFIX of their squared
// distance (the float's top bits; the register sorting network of the tile kernel works on 32-bit keys
// prefix | slot).  The reference order is (d2, index) on the full float and the point's index, so the walk below
// recomputes every member's full key bits(d2) << 32 | ord from the point itself (FLANN L2_Simple arithmetic) and
// repairs the order where two members share a prefix (exact ties on lattices and range-quantised scans; otherwise two
// distances within 2^-18 of each other): an out-of-order member is moved back by insertion and the walk starts
// again.  The first n_nb entries in that order are the neighbours (n_nb = k, or 0 when the lane has no result);
// entries beyond them only matter while they share the prefix of the n_nb-th.  Outputs as knn_finish: neighbour
// lists, k-th distance, moments in list order, then knn_emit.  `kth_out` receives the exact k-th squared distance.
__device__ __forceinline__ u64 knn_full_key(const float4 Q, const float4 P) {
  return ((u64)__float_as_uint(dist2_nofma(Q.x, Q.y, Q.z, P)) << 32) | (u64)__float_as_uint(P.w);
}

// moves member j (position `pos`, full key `key`) back to its place among the members before it
static __device__ __noinline__ void knn_fix_order(const float4* __restrict__ sorted, uint32_t* col, const float4 Q, const int j, const uint32_t pos,
                                           const u64 key) {
  int i = j - 1;
  while (i >= 0) {
    const uint32_t pi = col[i * 32];
    if (knn_full_key(Q, __ldg(&sorted[pi])) > key) {
      col[(i + 1) * 32] = pi;
      --i;
    } else {
      break;
    }
  }
  col[(i + 1) * 32] = pos;
}

template <bool LEAN = false>
__device__ __forceinline__ void knn_finish_pos(const KArgs& a, uint32_t* col, const int cnt, const int n_nb, const uint32_t prefix_mask,
                                               const float4 Q, const int my_slot, float (*rec_all)[17], const int lane,
                                               const int min_valid = 1) {
  const SearchParams& sp = a.sp;
  const int k = sp.k;
  const float4* __restrict__ sorted = a.sorted;
  const int ord_q = __float_as_int(Q.w);
  const int gid_q = (my_slot >= 0 && a.gid_of_ord) ? __ldg(&a.gid_of_ord[ord_q]) : ord_q;
  const bool valid = my_slot >= 0 && n_nb >= max(min_valid, 1);  // radius search: the 13-neighbour gate (src/pcaAnalysis.cpp:50)
  const bool need_q = LEAN || (sp.out != nullptr && sp.center_mode == PCP_CENTER_QUERY);
  const bool need_m = !LEAN && (sp.normals != nullptr || (sp.out != nullptr && sp.center_mode == PCP_CENTER_MEAN));
  float mq[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, mm[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float sx = 0.f, sy = 0.f, sz = 0.f;
  if (valid) {
    const int64_t o = (int64_t)my_slot * k;
    bool redo;
    u64 prev;
#pragma unroll 1
    do {
      redo = false;
      prev = 0ull;
      sx = sy = sz = 0.f;
#pragma unroll
      for (int c = 0; c < 6; ++c) mq[c] = 0.f;
      // four members per step: positions, then the four points in flight, then the ordered accumulation
#pragma unroll 1
      for (int j0 = 0; j0 < n_nb && !redo; j0 += 4) {
        uint32_t ps[4];
        float4 P[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) ps[u] = col[min(j0 + u, cnt - 1) * 32];
#pragma unroll
        for (int u = 0; u < 4; ++u) P[u] = __ldg(&sorted[ps[u]]);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + u;
          if (j < n_nb && !redo) {
            const u64 key = knn_full_key(Q, P[u]);
            if (key < prev) {
              knn_fix_order(sorted, col, Q, j, ps[u], key);
              redo = true;
            } else {
              prev = key;
              if (!LEAN && sp.nn_idx != nullptr) {
                const int ord = (int)(uint32_t)key;
                sp.nn_idx[o + j] = a.gid_of_ord ? __ldg(&a.gid_of_ord[ord]) : ord;
                if (sp.nn_d2) sp.nn_d2[o + j] = __uint_as_float((uint32_t)(key >> 32));
              }
              // pcl::computeCovarianceMatrixNormalized about the query: six fp32 chains in list order;
              // pcl::compute3DCentroid: fp32 running sums in list order
              const float px = __fsub_rn(P[u].x, Q.x), py = __fsub_rn(P[u].y, Q.y), pz = __fsub_rn(P[u].z, Q.z);
              mq[0] = __fadd_rn(mq[0], __fmul_rn(px, px));
              mq[1] = __fadd_rn(mq[1], __fmul_rn(py, px));
              mq[2] = __fadd_rn(mq[2], __fmul_rn(pz, px));
              mq[3] = __fadd_rn(mq[3], __fmul_rn(py, py));
              mq[4] = __fadd_rn(mq[4], __fmul_rn(py, pz));
              mq[5] = __fadd_rn(mq[5], __fmul_rn(pz, pz));
              if (!LEAN) {
                sx = __fadd_rn(sx, P[u].x);
                sy = __fadd_rn(sy, P[u].y);
                sz = __fadd_rn(sz, P[u].z);
              }
            }
          }
        }
      }
      // members beyond the n_nb-th that share its prefix may still precede it in the full order
#pragma unroll 1
      for (int j = n_nb; j < cnt && !redo; ++j) {
        const uint32_t pj = col[j * 32];
        const u64 key = knn_full_key(Q, __ldg(&sorted[pj]));
        if ((((uint32_t)(key >> 32)) & prefix_mask) != (((uint32_t)(prev >> 32)) & prefix_mask)) break;
        if (key < prev) {
          knn_fix_order(sorted, col, Q, j, pj, key);
          redo = true;
        }
      }
    } while (redo);
    if (sp.kth_d2) sp.kth_d2[my_slot] = n_nb >= k ? __uint_as_float((uint32_t)(prev >> 32)) : INFINITY;
    const float fn = (float)n_nb;
#pragma unroll
    for (int c = 0; c < 6; ++c) mq[c] = __fdiv_rn(mq[c], fn);
    if (need_m) {
      const float cx = __fdiv_rn(sx, fn), cy = __fdiv_rn(sy, fn), cz = __fdiv_rn(sz, fn);
#pragma unroll 4
      for (int j = 0; j < n_nb; ++j) {
        const float4 P = __ldg(&sorted[col[j * 32]]);
        const float px = __fsub_rn(P.x, cx), py = __fsub_rn(P.y, cy), pz = __fsub_rn(P.z, cz);
        mm[0] = __fadd_rn(mm[0], __fmul_rn(px, px));
        mm[1] = __fadd_rn(mm[1], __fmul_rn(py, px));
        mm[2] = __fadd_rn(mm[2], __fmul_rn(pz, px));
        mm[3] = __fadd_rn(mm[3], __fmul_rn(py, py));
        mm[4] = __fadd_rn(mm[4], __fmul_rn(py, pz));
        mm[5] = __fadd_rn(mm[5], __fmul_rn(pz, pz));
      }
#pragma unroll
      for (int c = 0; c < 6; ++c) mm[c] = __fdiv_rn(mm[c], fn);
    }
    (void)need_q;
  } else if (my_slot >= 0) {
    // no result here (non-finite query, or handed to the work list which rewrites these): empty lists
    if (sp.kth_d2) sp.kth_d2[my_slot] = INFINITY;
    if (!LEAN && sp.nn_idx != nullptr) {
      const int64_t o = (int64_t)my_slot * k;
      for (int j = 0; j < k; ++j) {
        sp.nn_idx[o + j] = -1;
        if (sp.nn_d2) sp.nn_d2[o + j] = INFINITY;
      }
    }
  }
  knn_emit<LEAN>(a, mq, mm, valid, n_nb, Q, my_slot, gid_q, rec_all, lane);
}
