// odom.cu — the lidarOdometry consumers of the radius search and the range image.
//
// Replaces lidarOdometryClass::getRangeAndptIdImage (src/lidarOdometry.cpp:21-113) and
// lidarOdometryClass::calculateSmoothness (src/lidarOdometry.cpp:121-187).
//
//  * radius_count_kernel: the reference runs a full radiusSearch(i, 3.0) per point only to test
//    `size() < 6` (:76-78).  Here one thread per point walks the stencil rows of the grid (own row
//    first) and stops at `cap` hits, so a VLP-16 return costs a handful of candidates instead of the
//    thousands inside 3 m.  The hit test is FLANN's (fp32 L2_Simple, strict d2 < (float)(r*r)).
//  * angles: the reference's atan() calls resolve to atanf (float overload, "math.h" in C++), then
//    `* 180.0 / M_PI` in double.  Rows of a VLP-16 sit at (almost) exact multiples of 2 degrees above
//    the minimum, so int(angle / 2.0) is sensitive to the last bit of atanf: the device evaluates the
//    same fdlibm algorithm glibc's atanf uses (checked equal on all 2^32 inputs on the host side,
//    tests/), with separate fp32 multiplies and adds.
//  * last writer wins (:97-100) = atomicMax of the frame-local index, as in ring.cu.
//  * smoothness: one warp per (frame, row, block of 180 columns); the window sum runs in double in
//    column order (:141-147), the two thresholds smoothnessRow[size-3] / smoothnessRow[3] of the sorted
//    block (:160-172) are order statistics found by rank counting in shared memory.
#include <cfloat>
#include <cmath>

#include "search_common.cuh"

namespace {

// ---- fdlibm atanf (the algorithm of glibc's sysdeps/ieee754/flt-32/s_atanf.c up to 2.40), no FMA ----
__device__ __forceinline__ float atanf_fdlibm(float x) {
  const float hi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
  const float lo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
  const int32_t hx = __float_as_int(x), ix = hx & 0x7fffffff;
  if (ix >= 0x4c000000) {  // |x| >= 2^25
    if (ix > 0x7f800000) return __fadd_rn(x, x);
    const float r = __fadd_rn(hi[3], lo[3]);
    return hx > 0 ? r : -r;
  }
  int id;
  if (ix < 0x3ee00000) {  // |x| < 0.4375
    if (ix < 0x31000000) return x;
    id = -1;
  } else {
    x = fabsf(x);
    if (ix < 0x3f980000) {
      if (ix < 0x3f300000) {
        id = 0;
        x = __fdiv_rn(__fsub_rn(__fmul_rn(2.0f, x), 1.0f), __fadd_rn(2.0f, x));
      } else {
        id = 1;
        x = __fdiv_rn(__fsub_rn(x, 1.0f), __fadd_rn(x, 1.0f));
      }
    } else {
      if (ix < 0x401c0000) {
        id = 2;
        x = __fdiv_rn(__fsub_rn(x, 1.5f), __fadd_rn(1.0f, __fmul_rn(1.5f, x)));
      } else {
        id = 3;
        x = __fdiv_rn(-1.0f, x);
      }
    }
  }
  const float z = __fmul_rn(x, x), w = __fmul_rn(z, z);
  float a = __fmul_rn(w, 1.6285819933e-02f);
  a = __fmul_rn(w, __fadd_rn(4.9768779427e-02f, a));
  a = __fmul_rn(w, __fadd_rn(6.6610731184e-02f, a));
  a = __fmul_rn(w, __fadd_rn(9.0908870101e-02f, a));
  a = __fmul_rn(w, __fadd_rn(1.4285714924e-01f, a));
  const float s1 = __fmul_rn(z, __fadd_rn(3.3333334327e-01f, a));
  float b = __fmul_rn(w, -3.6531571299e-02f);
  b = __fmul_rn(w, __fadd_rn(-5.8335702866e-02f, b));
  b = __fmul_rn(w, __fadd_rn(-7.6918758452e-02f, b));
  b = __fmul_rn(w, __fadd_rn(-1.1111110449e-01f, b));
  const float s2 = __fmul_rn(w, __fadd_rn(-2.0000000298e-01f, b));
  const float xs = __fmul_rn(x, __fadd_rn(s1, s2));
  if (id < 0) return __fsub_rn(x, xs);
  const float hi_id = id == 0 ? hi[0] : (id == 1 ? hi[1] : (id == 2 ? hi[2] : hi[3]));
  const float lo_id = id == 0 ? lo[0] : (id == 1 ? lo[1] : (id == 2 ? lo[2] : lo[3]));
  const float zz = __fsub_rn(hi_id, __fsub_rn(__fsub_rn(xs, lo_id), x));
  return hx < 0 ? -zz : zz;
}

__device__ __forceinline__ double deg_of(float a) {
  const double kPi = 3.14159265358979323846;
  return __ddiv_rn(__dmul_rn((double)a, 180.0), kPi);
}
// atan(z / sqrt(x*x + y*y)) * 180.0 / M_PI  (src/lidarOdometry.cpp:46-47)
__device__ __forceinline__ double vert_angle(float4 p) {
  const float xy = __fsqrt_rn(__fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y)));
  return deg_of(atanf_fdlibm(__fdiv_rn(p.z, xy)));
}
// atan(y / x) * 180.0 / M_PI  (:49)
__device__ __forceinline__ double hori_angle(float4 p) { return deg_of(atanf_fdlibm(__fdiv_rn(p.y, p.x))); }

// ---- capped radius count: one thread per cell-sorted point ----
__global__ void __launch_bounds__(128) radius_count_kernel(KArgs a, int R, float r2, int cap, int32_t* __restrict__ counts) {
  const int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= a.n) return;
  const int slot = __ldg(&a.slot[p]);
  if (slot < 0) return;  // halo point of a shard: neighbour only
  const GridDesc& g = a.g;
  const uint32_t cid = __ldg(&a.keys[p]);
  if (cid >= g.n_cells) {  // non-finite point: no neighbours
    counts[slot] = 0;
    return;
  }
  const float4 Q = __ldg(&a.sorted[p]);
  const int c0 = (int)(cid % (uint32_t)g.dim[0]);
  const uint32_t t = cid / (uint32_t)g.dim[0];
  const int c1 = (int)(t % (uint32_t)g.dim[1]);
  int c2 = (int)(t / (uint32_t)g.dim[1]);
  const int zoff = pcp_unfold_z(g, c2);
  const int lo0 = max(c0 - R, 0), hi0 = min(c0 + R, g.dim[0] - 1);
  const int lo1 = max(c1 - R, 0), hi1 = min(c1 + R, g.dim[1] - 1);
  const int lo2 = max(c2 - R, 0), hi2 = min(c2 + R, g.dim[2] - 1);
  const int n1 = hi1 - lo1 + 1, rows = n1 * (hi2 - lo2 + 1);
  const int own = (c2 - lo2) * n1 + (c1 - lo1);
  const uint32_t limit = cap > 0 ? (uint32_t)cap : 0xFFFFFFFFu;
  uint32_t cnt = 0;
#pragma unroll 1
  for (int it = 0; it < rows && cnt < limit; ++it) {
    // the own row first (it almost always holds `cap` hits), then the rest in raster order
    const int r = it == 0 ? own : (it <= own ? it - 1 : it);
    const int e2 = lo2 + r / n1, e1 = lo1 + r % n1;
    const uint32_t base = ((uint32_t)(zoff + e2) * (uint32_t)g.dim[1] + (uint32_t)e1) * (uint32_t)g.dim[0];
    const uint32_t beg = __ldg(&a.E[base + lo0]), end = __ldg(&a.E[base + hi0 + 1]);
#pragma unroll 1
    for (uint32_t j = beg; j < end && cnt < limit; ++j) {
      const float d2 = dist2_nofma(Q.x, Q.y, Q.z, __ldg(&a.sorted[j]));
      cnt += d2 < r2 ? 1u : 0u;
    }
  }
  counts[slot] = (int32_t)cnt;
}

struct OdomGeom {
  int n_rows, n_cols;
  double vert_res, hori_res;
  int min_nb;
};

// one block per frame: {vert min, vert max, hori min, hori max} with the reference's initial values
// (max 0, min 1000, :40-41); NaN never wins a comparison, as upstream.
__global__ void __launch_bounds__(256) odom_minmax_kernel(const float4* __restrict__ pts, const int64_t* __restrict__ frame_offsets,
                                                          int64_t n_single, double* __restrict__ minmax) {
  const int f = blockIdx.x;
  const int64_t beg = frame_offsets ? frame_offsets[f] : 0;
  const int64_t end = frame_offsets ? frame_offsets[f + 1] : n_single;
  double vmin = 1000.0, vmax = 0.0, hmin = 1000.0, hmax = 0.0;
  for (int64_t i = beg + threadIdx.x; i < end; i += blockDim.x) {
    const float4 p = __ldg(pts + i);
    const double v = vert_angle(p), h = hori_angle(p);
    if (v > vmax) vmax = v;
    if (v < vmin) vmin = v;
    if (h > hmax) hmax = h;
    if (h < hmin) hmin = h;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    vmin = fmin(vmin, __shfl_xor_sync(PCP_FULL, vmin, o));
    vmax = fmax(vmax, __shfl_xor_sync(PCP_FULL, vmax, o));
    hmin = fmin(hmin, __shfl_xor_sync(PCP_FULL, hmin, o));
    hmax = fmax(hmax, __shfl_xor_sync(PCP_FULL, hmax, o));
  }
  __shared__ double s[8][4];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s[w][0] = vmin;
    s[w][1] = vmax;
    s[w][2] = hmin;
    s[w][3] = hmax;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 8; ++k) {
      vmin = fmin(vmin, s[k][0]);
      vmax = fmax(vmax, s[k][1]);
      hmin = fmin(hmin, s[k][2]);
      hmax = fmax(hmax, s[k][3]);
    }
    minmax[4 * f + 0] = vmin;
    minmax[4 * f + 1] = vmax;
    minmax[4 * f + 2] = hmin;
    minmax[4 * f + 3] = hmax;
  }
}

__global__ void __launch_bounds__(256) odom_init_kernel(int32_t* __restrict__ pt_index, int32_t* __restrict__ stats, int64_t cells,
                                                        int64_t n_frames) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < cells) pt_index[i] = -1;
  if (i < 2 * n_frames) stats[i] = 0;
}

// grid = (ceil(max_frame_len / 256), n_frames)
__global__ void __launch_bounds__(256) odom_scatter_kernel(const float4* __restrict__ pts, const int64_t* __restrict__ frame_offsets,
                                                           int64_t n_single, OdomGeom g, const int32_t* __restrict__ nb_counts,
                                                           const double* __restrict__ minmax, int32_t* __restrict__ pt_index,
                                                           int32_t* __restrict__ stats) {
  const int f = blockIdx.y;
  const int64_t beg = frame_offsets ? frame_offsets[f] : 0;
  const int64_t end = frame_offsets ? frame_offsets[f + 1] : n_single;
  const int64_t i = beg + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool accepted = false;
  if (i < end && nb_counts[i] >= g.min_nb) {  // radiusSearch(i, 3.0).size() >= 6 (:76-78)
    const float4 p = __ldg(pts + i);
    const double v = __dsub_rn(vert_angle(p), minmax[4 * f + 0]);  // (:80-83)
    const double h = __dsub_rn(hori_angle(p), minmax[4 * f + 2]);
    const double rd = __ddiv_rn(v, g.vert_res), cd = __ddiv_rn(h, g.hori_res);  // (:85-86)
    if (rd > -1.0 && rd < 2.0e9 && cd > -1.0 && cd < 2.0e9) {
      const int row = (int)rd, col = (int)cd;  // truncation toward zero
      if (row >= 0 && row < g.n_rows && col >= 0 && col < g.n_cols) {
        atomicMax(pt_index + ((int64_t)f * g.n_rows + row) * g.n_cols + col, (int32_t)(i - beg));
        accepted = true;
      }
    }
  }
  const unsigned m = __ballot_sync(PCP_FULL, accepted);
  if (m && (threadIdx.x & 31) == 0) atomicAdd(stats + 2 * f, __popc(m));  // writes; the fill turns it into overwrites
}

__global__ void __launch_bounds__(256) odom_fill_kernel(const float4* __restrict__ pts, const int64_t* __restrict__ frame_offsets,
                                                        OdomGeom g, const int32_t* __restrict__ pt_index, float* __restrict__ range_img,
                                                        float* __restrict__ ptid_img, int32_t* __restrict__ stats, int64_t cells) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool occupied = false;
  int f = 0;
  if (c < cells) {
    const int32_t w = pt_index[c];
    f = (int)(c / ((int64_t)g.n_rows * g.n_cols));
    float range = 0.f, id = 0.f;  // MatrixXf::Zero (:31-32)
    if (w >= 0) {
      const float4 p = __ldg(pts + (frame_offsets ? frame_offsets[f] : 0) + w);
      range = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y)), __fmul_rn(p.z, p.z)));  // (:97-99)
      id = (float)w;                                                                                             // (:100)
      occupied = range > 0.f;                                                                                    // (:108)
    }
    range_img[c] = range;
    if (ptid_img) ptid_img[c] = id;
  }
  // a warp may straddle two frames: count per frame of lane 0 and of the others
  const int f0 = __shfl_sync(PCP_FULL, f, 0);
  const unsigned same = __ballot_sync(PCP_FULL, occupied && f == f0);
  const unsigned other = __ballot_sync(PCP_FULL, occupied && f != f0);
  if ((threadIdx.x & 31) == 0 && same) atomicAdd(stats + 2 * f0 + 1, __popc(same));
  if (other && occupied && f != f0 && (threadIdx.x & 31) == (__ffs(other) - 1)) atomicAdd(stats + 2 * f + 1, __popc(other));
}

// stats[2f] = writes -> overwrites ("losing points num", :95-96) = writes - occupied cells
__global__ void odom_stats_kernel(int32_t* __restrict__ stats, int64_t n_frames) {
  const int64_t f = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (f < n_frames) stats[2 * f] -= stats[2 * f + 1];
}

struct SmoothGeom {
  int n_rows, n_cols, n_blocks, half_window, min_valid, pick, min_block_valid;
  float near_range;
};

#define SM_WARPS 4
#define SM_MAXW 256  // columns of one block a warp can hold

// one warp per (frame, row, block)
__global__ void __launch_bounds__(SM_WARPS * 32) odom_smooth_kernel(const float* __restrict__ range_img, SmoothGeom g, int64_t units,
                                                                    float* __restrict__ smooth_img, float* __restrict__ type_img) {
  __shared__ double vals[SM_WARPS][SM_MAXW];
  __shared__ double thr[SM_WARPS][2];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t u = blockIdx.x * (int64_t)SM_WARPS + wid;
  if (u >= units) return;
  const int b = (int)(u % g.n_blocks);
  const int64_t fr = u / g.n_blocks;  // frame * n_rows + row
  const float* __restrict__ R = range_img + fr * g.n_cols;
  const int bw = g.n_cols / g.n_blocks;
  const int col0 = bw * b + g.half_window, ncol = bw - g.half_window;  // columns col0 .. bw*(b+1)-1 (:133)
  double* v = vals[wid];
  int m_lane = 0;
  for (int t = lane; t < ncol; t += 32) {
    const int col = col0 + t;
    const float rc = R[col];
    double s = -1.0;  // marks "no smoothness" (real values are >= 0)
    if (!(rc < g.near_range)) {  // (:136-137)
      int valid = 0;
      double sum = 0.0;
      for (int i = col - g.half_window; i < col + g.half_window; ++i) {  // (:141)
        const float ri = i < g.n_cols ? R[i] : 0.f;
        if (ri > g.near_range) {
          ++valid;
          sum = __dadd_rn(sum, (double)ri);
        }
      }
      if (valid >= g.min_valid) {
        const float prod = __fmul_rn(rc, (float)valid);  // float * int (:151-152)
        s = fabs(__ddiv_rn(__dsub_rn(sum, (double)prod), (double)prod));
        ++m_lane;
      }
    }
    v[t] = s;
  }
  int m = m_lane;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(PCP_FULL, m, o);
  __syncwarp();
  const bool pickable = m > g.min_block_valid;  // (:162)
  if (pickable) {
    // order statistics by rank counting: sorted[k] = the value with #less <= k < #less-or-equal
    const int k_lo = g.pick, k_hi = m - g.pick;
    for (int t = lane; t < ncol; t += 32) {
      const double s = v[t];
      if (s < 0.0) continue;
      int lt = 0, le = 0;
      for (int j = 0; j < ncol; ++j) {
        const double o = v[j];
        lt += (o >= 0.0 && o < s) ? 1 : 0;
        le += (o >= 0.0 && o <= s) ? 1 : 0;
      }
      if (lt <= k_lo && k_lo < le) thr[wid][0] = s;  // every writer holds the same value
      if (lt <= k_hi && k_hi < le) thr[wid][1] = s;
    }
  }
  __syncwarp();
  // the first half_window columns of the block are never touched upstream: they stay 0
  for (int t = lane; t < bw; t += 32) {
    const int col = bw * b + t;
    float sm = 0.f, ty = 0.f;
    if (t >= g.half_window) {
      const double s = v[t - g.half_window];
      sm = s >= 0.0 ? (float)s : 0.f;  // smoothnessImg is a float matrix (:154)
      if (pickable) {
        if ((double)sm >= thr[wid][1]) ty = 1.f;  // (:167-169)
        if ((double)sm < thr[wid][0]) ty = 2.f;   // (:170-172)
      }
    }
    if (smooth_img) smooth_img[fr * g.n_cols + col] = sm;
    type_img[fr * g.n_cols + col] = ty;
  }
  // columns past n_blocks * bw (none for 900 / 5) stay 0
  if (b == g.n_blocks - 1)
    for (int col = bw * g.n_blocks + lane; col < g.n_cols; col += 32) {
      if (smooth_img) smooth_img[fr * g.n_cols + col] = 0.f;
      type_img[fr * g.n_cols + col] = 0.f;
    }
}

}  // namespace

int pcp_radius_count_impl(pcp_ctx* ctx, float r2, int R, int cap, int32_t* d_counts) {
  KArgs a;
  a.sorted = (const float4*)ctx->sorted_pts.p;
  a.keys = (const uint32_t*)ctx->keys_sorted.p;
  a.slot = (const int32_t*)ctx->sorted_slot.p;
  a.E = (const uint32_t*)ctx->cell_end.p;
  a.inv = (const uint32_t*)ctx->inv_pos.p;
  a.gid_of_ord = nullptr;
  a.g = ctx->grid;
  a.n = ctx->n;
  memset(&a.sp, 0, sizeof(a.sp));
  const unsigned blocks = (unsigned)((ctx->n + 127) / 128);
  radius_count_kernel<<<blocks, 128, 0, ctx->stream>>>(a, R, r2, cap, d_counts);
  PCP_LAUNCH_CHECK("radius_count_kernel");
  return PCP_OK;
}

int pcp_odom_range_impl(pcp_ctx* ctx, const pcp_odom_params* prm, const int64_t* d_off, int64_t n_frames, int64_t max_len,
                        const int32_t* d_counts, float* d_range, float* d_ptid, int32_t* d_index, int32_t* d_stats, double* d_minmax) {
  OdomGeom g;
  g.n_rows = prm->n_rows;
  g.n_cols = prm->n_cols;
  g.vert_res = prm->vert_res_deg;
  g.hori_res = prm->hori_res_deg;
  g.min_nb = prm->min_neighbors;
  const int64_t cells = (int64_t)g.n_rows * g.n_cols * n_frames;
  odom_minmax_kernel<<<(unsigned)n_frames, 256, 0, ctx->stream>>>(ctx->d_cloud, d_off, ctx->n, d_minmax);
  PCP_LAUNCH_CHECK("odom_minmax_kernel");
  const int64_t init_n = std::max(cells, 2 * n_frames);
  odom_init_kernel<<<(unsigned)((init_n + 255) / 256), 256, 0, ctx->stream>>>(d_index, d_stats, cells, n_frames);
  PCP_LAUNCH_CHECK("odom_init_kernel");
  if (max_len > 0) {
    dim3 grid((unsigned)((max_len + 255) / 256), (unsigned)n_frames);
    odom_scatter_kernel<<<grid, 256, 0, ctx->stream>>>(ctx->d_cloud, d_off, ctx->n, g, d_counts, d_minmax, d_index, d_stats);
    PCP_LAUNCH_CHECK("odom_scatter_kernel");
  }
  odom_fill_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_cloud, d_off, g, d_index, d_range, d_ptid, d_stats,
                                                                            cells);
  PCP_LAUNCH_CHECK("odom_fill_kernel");
  odom_stats_kernel<<<(unsigned)((n_frames + 255) / 256), 256, 0, ctx->stream>>>(d_stats, n_frames);
  PCP_LAUNCH_CHECK("odom_stats_kernel");
  return PCP_OK;
}

int pcp_odom_smooth_impl(pcp_ctx* ctx, const pcp_odom_params* prm, const float* d_range, int64_t n_frames, float* d_smooth,
                         float* d_type) {
  SmoothGeom g;
  g.n_rows = prm->n_rows;
  g.n_cols = prm->n_cols;
  g.n_blocks = prm->n_blocks;
  g.half_window = prm->half_window;
  g.min_valid = prm->min_valid;
  g.pick = prm->pick;
  g.min_block_valid = prm->min_block_valid;
  g.near_range = prm->near_range;
  const int64_t units = n_frames * g.n_rows * g.n_blocks;
  odom_smooth_kernel<<<(unsigned)((units + SM_WARPS - 1) / SM_WARPS), SM_WARPS * 32, 0, ctx->stream>>>(d_range, g, units, d_smooth, d_type);
  PCP_LAUNCH_CHECK("odom_smooth_kernel");
  return PCP_OK;
}
