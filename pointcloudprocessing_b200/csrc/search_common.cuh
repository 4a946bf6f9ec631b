// search_common.cuh — warp-level building blocks of the fused search + PCA kernel.
//
// Candidate ordering everywhere is the reference's (d2, index) order (FLANN
// DistanceIndex::operator<).  A candidate is carried as one 64-bit key
//     key = (bits(d2) << 32) | sorted_position
// d2 >= 0, so its IEEE bit pattern orders like the value.  Ordering two keys with different
// d2 is a plain 64-bit compare; only when the d2 bits are EQUAL (exact fp32 ties) the index
// stored in .w of the sorted point decides, which costs two L1-resident loads on a rare path.
// Carrying the sorted position (not the index) lets the moment stage re-read neighbour
// coordinates from the cell-ordered array, i.e. from lines the scan just touched.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

typedef unsigned long long u64;
#define PCP_KEY_INF 0xFFFFFFFFFFFFFFFFull
#define PCP_FULL 0xffffffffu

__device__ __forceinline__ bool key_less(u64 a, u64 b, const float4* __restrict__ sorted) {
  uint32_t ha = (uint32_t)(a >> 32), hb = (uint32_t)(b >> 32);
  if (ha != hb) return ha < hb;
  uint32_t pa = (uint32_t)a, pb = (uint32_t)b;
  if (pa == pb) return false;
  if (ha == 0xFFFFFFFFu) return pa < pb;  // padding entries
  int ga = __float_as_int(__ldg(&sorted[pa].w));
  int gb = __float_as_int(__ldg(&sorted[pb].w));
  return ga != gb ? ga < gb : pa < pb;
}

// FLANN L2_Simple<float>: ((dx*dx) + dy*dy) + dz*dz with separate multiplies and adds.
__device__ __forceinline__ float dist2_nofma(float qx, float qy, float qz, float4 p) {
  float dx = __fsub_rn(qx, p.x), dy = __fsub_rn(qy, p.y), dz = __fsub_rn(qz, p.z);
  float r = __fmul_rn(dx, dx);
  r = __fadd_rn(r, __fmul_rn(dy, dy));
  r = __fadd_rn(r, __fmul_rn(dz, dz));
  return r;
}

__device__ __forceinline__ u64 shfl_u64(u64 v, int src) {
  return __shfl_sync(PCP_FULL, v, src);
}
__device__ __forceinline__ u64 shfl_xor_u64(u64 v, int m) {
  return __shfl_xor_sync(PCP_FULL, v, m);
}

// ascending bitonic sort of one key per lane
__device__ __forceinline__ u64 warp_sort32(u64 v, int lane, const float4* __restrict__ sorted) {
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      u64 o = shfl_xor_u64(v, j);
      bool lower = (lane & j) == 0;
      bool asc = (k == 32) ? true : ((lane & k) == 0);
      bool take_min = (lower == asc);
      bool o_less = key_less(o, v, sorted);
      v = (take_min == o_less) ? o : v;
    }
  }
  return v;
}

// best (ascending, one per lane) <- the 32 smallest of best ∪ v, v ascending
__device__ __forceinline__ u64 warp_merge32(u64 best, u64 v, int lane, const float4* __restrict__ sorted) {
  u64 vr = shfl_u64(v, 31 - lane);
  u64 w = key_less(vr, best, sorted) ? vr : best;  // bitonic sequence of the 32 smallest
#pragma unroll
  for (int j = 16; j > 0; j >>= 1) {
    u64 o = shfl_xor_u64(w, j);
    bool lower = (lane & j) == 0;
    bool o_less = key_less(o, w, sorted);
    w = (lower == o_less) ? o : w;
  }
  return w;
}

// in-place ascending bitonic sort of m (power of two, >= 64) keys in warp-private shared memory
__device__ __forceinline__ void smem_sort_pow2(u64* buf, int m, int lane, const float4* __restrict__ sorted) {
  for (int k = 2; k <= m; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = lane; t < (m >> 1); t += 32) {
        // t-th pair of this stage: insert a zero bit at position log2(j)
        int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        int p = i | j;
        bool asc = (i & k) == 0;
        u64 a = buf[i], b = buf[p];
        bool b_less = key_less(b, a, sorted);
        if (b_less == asc) {
          buf[i] = b;
          buf[p] = a;
        }
      }
      __syncwarp();
    }
  }
}
