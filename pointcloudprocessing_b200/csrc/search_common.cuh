// search_common.cuh — warp-level building blocks of the fused search + PCA kernel.
//
// Candidate ordering everywhere is the reference's (d2, index) order (FLANN
// DistanceIndex::operator<).  A candidate is carried as one 64-bit key
//     key = (bits(d2) << 32) | ord
// where d2 >= 0 (its IEEE bit pattern orders like the value) and `ord` is the point's rank in
// index order (the local index itself for an unsharded cloud, the rank of its global id for a
// shard), so the reference order is a plain unsigned 64-bit compare: no tie path, no loads.
// `inv[ord]` gives the cell-sorted position, from which the moment stage re-reads coordinates.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pcp_internal.cuh"

typedef unsigned long long u64;
#define PCP_KEY_INF 0xFFFFFFFFFFFFFFFFull
#define PCP_FULL 0xffffffffu

// FLANN L2_Simple<float>: ((dx*dx) + dy*dy) + dz*dz with separate multiplies and adds.
__device__ __forceinline__ float dist2_nofma(float qx, float qy, float qz, float4 p) {
  float dx = __fsub_rn(qx, p.x), dy = __fsub_rn(qy, p.y), dz = __fsub_rn(qz, p.z);
  float r = __fmul_rn(dx, dx);
  r = __fadd_rn(r, __fmul_rn(dy, dy));
  r = __fadd_rn(r, __fmul_rn(dz, dz));
  return r;
}

__device__ __forceinline__ u64 shfl_u64(u64 v, int src) { return __shfl_sync(PCP_FULL, v, src); }
__device__ __forceinline__ u64 shfl_xor_u64(u64 v, int m) { return __shfl_xor_sync(PCP_FULL, v, m); }

// Direction bits of the 32-lane bitonic network, one bit per compare-exchange step in issue
// order (15 sort steps, then the 5 steps of the final merge network are the last 5 of those):
// bit s set = this lane keeps the SMALLER key in step s.
__device__ __forceinline__ uint32_t bitonic_dirs(int lane) {
  uint32_t d = 0;
  int s = 0;
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1, ++s) {
      bool lower = (lane & j) == 0;
      bool asc = (k == 32) ? true : ((lane & k) == 0);
      if (lower == asc) d |= 1u << s;
    }
  }
  return d;
}

// ascending bitonic sort of one key per lane
__device__ __forceinline__ u64 warp_sort32(u64 v, uint32_t dirs) {
  int s = 0;
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1, ++s) {
      u64 o = shfl_xor_u64(v, j);
      bool take_min = (dirs >> s) & 1u;
      v = ((o < v) == take_min) ? o : v;
    }
  }
  return v;
}

// best (ascending, one per lane) <- the 32 smallest of best ∪ v, v ascending
__device__ __forceinline__ u64 warp_merge32(u64 best, u64 v, int lane, uint32_t dirs) {
  u64 vr = shfl_u64(v, 31 - lane);
  u64 w = vr < best ? vr : best;  // bitonic sequence holding the 32 smallest
  int s = 10;                     // the k = 32 stage of the network: steps 10..14
#pragma unroll
  for (int j = 16; j > 0; j >>= 1, ++s) {
    u64 o = shfl_xor_u64(w, j);
    bool take_min = (dirs >> s) & 1u;
    w = ((o < w) == take_min) ? o : w;
  }
  return w;
}

// in-place ascending bitonic sort of m (power of two, >= 64) keys in warp-private shared memory
__device__ __forceinline__ void smem_sort_pow2(u64* buf, int m, int lane) {
  for (int k = 2; k <= m; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (m >> 1); t += 32) {
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));  // t-th pair: insert a zero bit at log2(j)
        int p = i | j;
        bool asc = (i & k) == 0;
        u64 a = buf[i], b = buf[p];
        if ((b < a) == asc) {
          buf[i] = b;
          buf[p] = a;
        }
      }
      __syncwarp();
    }
  }
}

// kernel arguments shared by the radius (search_pca.cu) and k-NN (knn_pca.cu) kernels
struct KArgs {
  const float4* sorted;       // cell-ordered points, .w = ord bits
  const uint32_t* keys;       // cell id per sorted position
  const int32_t* slot;        // output slot per sorted position (−1 = halo)
  const uint32_t* E;          // E[c] = #points with cell id < c
  const uint32_t* inv;        // ord -> sorted position
  const int32_t* gid_of_ord;  // ord -> global id (null = identity)
  GridDesc g;
  int64_t n;
  SearchParams sp;
};


// squared radius around any point of the query cell that the block of half-width R is guaranteed to cover
__device__ __forceinline__ float cover2_of(int R, float cell) {
  double rs = ((double)R - 1e-6) * (double)cell;
  return (float)(rs * rs * (1.0 - 1e-6));
}

// gap (in cells, >= 0) between the query at fractional offset f of its cell and the cell d cells
// away; gaps below 2e-3 cells count as 0 so fp rounding of f can never prune a real neighbour
__device__ __forceinline__ float cell_gap(int d, float f) {
  float gp = d > 0 ? (float)d - f : (d < 0 ? f - (float)d - 1.f : 0.f);
  return gp < 2e-3f ? 0.f : gp;
}


#include "sortnet32.cuh"
