// tile_common.cuh — what the tile kernels (knn_tile.cu: radius search, knn_tile6.cu: k-NN) share: the stencil rows of a
// tile of 32 cell-sorted queries as contiguous runs of the sorted point array, and every lane's cover radius.
#pragma once
#include "pcp_internal.cuh"
#include "search_common.cuh"

constexpr int TA_MAXROWS = 64;     // stencil rows of a tile: (2R+1)^2 <= 49, R <= 3
constexpr int TA_MAXR = 3;

// fractional position of Q inside its cell c along the x/y/z axis `ax`, clamped to [0,1] (same double
// arithmetic as the key kernel; selects instead of dynamic indexing keep the arguments in the constant bank)
__device__ __forceinline__ float frac_of(const GridDesc& g, int ax, const float4 Q, int c) {
  const double v = ax == 0 ? (double)Q.x : (ax == 1 ? (double)Q.y : (double)Q.z);
  const double o = ax == 0 ? g.origin[0] : (ax == 1 ? g.origin[1] : g.origin[2]);
  return fminf(fmaxf((float)((v - o) * g.inv_cell - (double)c), 0.f), 1.f);
}

// One row group of a tile: the lanes whose query lies in stencil row `row0`.  Looks up the (2R+1)^2 runs
// of the union of their stencils, writes them to sm.run_*, and returns the number of candidates;
// cover_cells = radius (in cells) around THIS lane's query that the union is guaranteed to cover: its
// distance to the nearest face of the union block that still has cells behind it (a face on the grid
// boundary has no points beyond it), minus 2e-3 cells for the rounding of the fractional position.
__device__ __forceinline__ uint32_t tile_setup(const KArgs& a, uint32_t* __restrict__ run_beg, uint32_t* __restrict__ run_off, const int lane, const int R, const int c0min,
                                               const int c0max, const int c1, const int c2, const int zoff, const int c0, const float4 Q,
                                               int& nrows, bool& whole, float& cover_cells) {
  const GridDesc& g = a.g;
  const int xlo = max(c0min - R, 0), xhi = min(c0max + R, g.dim[0] - 1);
  const int lo1 = max(c1 - R, 0), n1 = min(c1 + R, g.dim[1] - 1) - lo1 + 1;
  const int lo2 = max(c2 - R, 0), n2 = min(c2 + R, g.dim[2] - 1) - lo2 + 1;
  nrows = n1 * n2;
  uint32_t beg0 = 0, len0 = 0, beg1 = 0, len1 = 0;
  if (lane < nrows) {
    const int o2 = lane / n1, o1 = lane - o2 * n1;
    const uint32_t base = ((uint32_t)(zoff + lo2 + o2) * (uint32_t)g.dim[1] + (uint32_t)(lo1 + o1)) * (uint32_t)g.dim[0];
    beg0 = __ldg(&a.E[base + xlo]);
    len0 = __ldg(&a.E[base + xhi + 1]) - beg0;
  }
  if (lane + 32 < nrows) {
    const int o2 = (lane + 32) / n1, o1 = (lane + 32) - o2 * n1;
    const uint32_t base = ((uint32_t)(zoff + lo2 + o2) * (uint32_t)g.dim[1] + (uint32_t)(lo1 + o1)) * (uint32_t)g.dim[0];
    beg1 = __ldg(&a.E[base + xlo]);
    len1 = __ldg(&a.E[base + xhi + 1]) - beg1;
  }
  uint32_t inc0 = len0, inc1 = len1;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t0 = __shfl_up_sync(PCP_FULL, inc0, o), t1 = __shfl_up_sync(PCP_FULL, inc1, o);
    if (lane >= o) {
      inc0 += t0;
      inc1 += t1;
    }
  }
  const uint32_t tot0 = __shfl_sync(PCP_FULL, inc0, 31);
  const uint32_t total = tot0 + __shfl_sync(PCP_FULL, inc1, 31);
  __syncwarp();
  run_beg[lane] = beg0;
  run_off[lane] = inc0 - len0;
  run_beg[lane + 32] = beg1;
  run_off[lane + 32] = tot0 + inc1 - len1;
  __syncwarp();
  whole = xlo == 0 && xhi == g.dim[0] - 1 && n1 == g.dim[1] && n2 == g.dim[2];
  const float f0 = frac_of(g, g.ax[0], Q, c0), f1 = frac_of(g, g.ax[1], Q, c1), f2 = frac_of(g, g.ax[2], Q, c2);
  const float inf = 3.0e18f;
  const float m0a = xlo > 0 ? (float)(c0 - xlo) + f0 : inf, m0b = xhi < g.dim[0] - 1 ? (float)(xhi + 1 - c0) - f0 : inf;
  const float m1a = c1 - R > 0 ? (float)R + f1 : inf, m1b = c1 + R < g.dim[1] - 1 ? (float)(R + 1) - f1 : inf;
  const float m2a = c2 - R > 0 ? (float)R + f2 : inf, m2b = c2 + R < g.dim[2] - 1 ? (float)(R + 1) - f2 : inf;
  cover_cells = fminf(fminf(fminf(m0a, m0b), fminf(m1a, m1b)), fminf(m2a, m2b)) - 2e-3f;
  return total;
}

