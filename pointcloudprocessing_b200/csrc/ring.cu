// ring.cu — 16-beam range-image binning on the device.
//
// Replaces RingProjector::projectPointCloud (src/ringProjection.cpp:24-81; constants
// include/ringProjection.h:16-20).  The reference loop is sequential and a later point
// overwrites an earlier one in the same cell; here every point does an
// atomicMax of its frame-local index on the cell, which yields the same winner (the highest
// input index), and a second pass fills the range of the winner.  Mixed precision follows
// the reference's C++ promotions; atan2f is evaluated as a double atan2 rounded to float so
// the binning matches a correctly rounded libm atan2f.
#include <cfloat>
#include <cmath>

#include "pcp_internal.cuh"

namespace {

struct RingGeom {
  int n_scan, horizon;
  float ang_res_x, ang_res_y, ang_bottom;
};

__device__ __forceinline__ float atan2f_cr(float y, float x) { return (float)atan2((double)y, (double)x); }

__device__ __forceinline__ bool ring_cell(const RingGeom& g, float4 p, int& cell, float& range) {
  const double kPi = 3.14159265358979323846;
  float xy2 = __fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y));
  float vertical = (float)((double)__fmul_rn(atan2f_cr(p.z, __fsqrt_rn(xy2)), 180.0f) / kPi);
  float rowf = __fdiv_rn(__fadd_rn(vertical, g.ang_bottom), g.ang_res_y);
  if (!(rowf > -9.0e18f && rowf < 9.0e18f)) return false;
  long long row = (long long)rowf;  // truncation toward zero, like the size_t conversion for small negatives
  if (row < 0 || row >= g.n_scan) return false;
  float hang = (float)((double)__fmul_rn(atan2f_cr(p.x, p.y), 180.0f) / kPi);
  double cold = -round(((double)hang - 90.0) / (double)g.ang_res_x) + (double)(g.horizon / 2);
  if (!(cold > -9.0e18 && cold < 9.0e18)) return false;
  long long col = (long long)cold;
  if (col >= g.horizon) col -= g.horizon;
  if (col < 0 || col >= g.horizon) return false;
  range = __fsqrt_rn(__fadd_rn(xy2, __fmul_rn(p.z, p.z)));
  cell = (int)(col + row * g.horizon);
  return true;
}

__global__ void __launch_bounds__(256) ring_init_kernel(float* __restrict__ range_img, int32_t* __restrict__ pt_index,
                                                        int32_t* __restrict__ long_range, int64_t total_cells, int64_t n_frames) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < total_cells) {
    range_img[i] = FLT_MAX;  // rangeMat init, include/ringProjection.h:142
    pt_index[i] = -1;
  }
  if (long_range && i < n_frames) long_range[i] = 0;
}

// grid = (ceil(max_frame_len / 256), n_frames)
__global__ void __launch_bounds__(256) ring_scatter_kernel(const float4* __restrict__ pts, const int64_t* __restrict__ frame_offsets,
                                                           int64_t n_single, RingGeom g, int32_t* __restrict__ pt_index,
                                                           int32_t* __restrict__ long_range) {
  const int f = blockIdx.y;
  const int64_t beg = frame_offsets ? frame_offsets[f] : 0;
  const int64_t end = frame_offsets ? frame_offsets[f + 1] : n_single;
  const int64_t i = beg + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  bool far = false;
  if (i < end) {
    float4 p = __ldg(pts + i);
    int cell;
    float range;
    if (ring_cell(g, p, cell, range)) {
      atomicMax(pt_index + (int64_t)f * g.n_scan * g.horizon + cell, (int32_t)(i - beg));
      far = range > 60.f;
    }
  }
  if (long_range) {
    unsigned m = __ballot_sync(0xffffffffu, far);
    if (m && (threadIdx.x & 31) == 0) atomicAdd(long_range + f, __popc(m));
  }
}

__global__ void __launch_bounds__(256) ring_fill_kernel(const float4* __restrict__ pts, const int64_t* __restrict__ frame_offsets,
                                                        RingGeom g, const int32_t* __restrict__ pt_index, float* __restrict__ range_img,
                                                        int64_t total_cells) {
  int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= total_cells) return;
  int32_t w = pt_index[c];
  if (w < 0) return;
  int64_t f = c / ((int64_t)g.n_scan * g.horizon);
  int64_t beg = frame_offsets ? frame_offsets[f] : 0;
  float4 p = __ldg(pts + beg + w);
  float xy2 = __fadd_rn(__fmul_rn(p.x, p.x), __fmul_rn(p.y, p.y));
  range_img[c] = __fsqrt_rn(__fadd_rn(xy2, __fmul_rn(p.z, p.z)));
}

}  // namespace

int pcp_ring_project_impl(pcp_ctx* ctx, const pcp_ring_calib* calib, const int64_t* h_offsets, int64_t n_frames,
                          float* d_range, int32_t* d_index, int32_t* d_long) {
  RingGeom g;
  g.n_scan = calib->n_scan;
  g.horizon = calib->horizon_scan;
  g.ang_res_x = calib->ang_res_x;
  g.ang_res_y = calib->ang_res_y;
  g.ang_bottom = calib->ang_bottom;
  const int64_t cells = (int64_t)g.n_scan * g.horizon * n_frames;
  int64_t max_len = ctx->n;
  const int64_t* d_off = nullptr;
  if (h_offsets) {
    max_len = 0;
    for (int64_t f = 0; f < n_frames; ++f) max_len = std::max(max_len, h_offsets[f + 1] - h_offsets[f]);
    int rc = pcp_reserve(ctx, ctx->ring_offsets, (n_frames + 1) * sizeof(int64_t));
    if (rc) return rc;
    PCP_CUDA(cudaMemcpyAsync(ctx->ring_offsets.p, h_offsets, (n_frames + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
    d_off = (const int64_t*)ctx->ring_offsets.p;
  }
  const int64_t init_n = std::max(cells, n_frames);
  ring_init_kernel<<<(unsigned)((init_n + 255) / 256), 256, 0, ctx->stream>>>(d_range, d_index, d_long, cells, n_frames);
  PCP_LAUNCH_CHECK("ring_init_kernel");
  if (max_len > 0) {
    dim3 grid((unsigned)((max_len + 255) / 256), (unsigned)n_frames);
    ring_scatter_kernel<<<grid, 256, 0, ctx->stream>>>(ctx->d_cloud, d_off, ctx->n, g, d_index, d_long);
    PCP_LAUNCH_CHECK("ring_scatter_kernel");
    ring_fill_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_cloud, d_off, g, d_index, d_range, cells);
    PCP_LAUNCH_CHECK("ring_fill_kernel");
  }
  return PCP_OK;
}
