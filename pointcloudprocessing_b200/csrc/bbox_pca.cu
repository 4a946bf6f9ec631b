// bbox_pca.cu — pcaAnalysist::getMiniBoundingBox (src/pcaAnalysis.cpp:123-148) as device reductions: centroid of the
// whole cloud (pcl::compute3DCentroid, :125-126), its moment about that centroid (pcl::computeCovarianceMatrixNormalized,
// :128-129), the eigen-decomposition with the third axis replaced by col0 x col1 (:131-133), the cloud's extents in
// that frame (:135-146) and meanDiag = (max + min) / 2 (:147).  The reference discards the result (and dereferences
// a null cloud on the way), so there is nothing downstream to be bit-compatible with: the sums run in
// double (the reference's sequential fp32 running sums carry ~1e-4 relative error on a million points), the
// eigen-solve and the rotation are the ones every record uses.
#include <cfloat>

#include "eigen3.cuh"
#include "pcp_internal.cuh"

namespace {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ bool finite3b(float4 p) { return isfinite(p.x) && isfinite(p.y) && isfinite(p.z); }

// acc[0..2] += sum of x, y, z; acc[3] += number of finite points
__global__ void __launch_bounds__(256) obb_sum_kernel(const float4* __restrict__ pts, int64_t n, double* __restrict__ acc) {
  double s[4] = {0, 0, 0, 0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 p = __ldg(pts + i);
    if (finite3b(p)) {
      s[0] += (double)p.x; s[1] += (double)p.y; s[2] += (double)p.z; s[3] += 1.0;
    }
  }
  __shared__ double sh[8][4];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const double v = warp_sum(s[c]);
    if (lane == 0) sh[w][c] = v;
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double v = 0;
    for (int k = 0; k < 8; ++k) v += sh[k][threadIdx.x];
    atomicAdd(acc + threadIdx.x, v);
  }
}

// acc[4..9] += the six second moments about the centroid acc[0..2] / acc[3] (cast to fp32 first: the reference
// hands a Vector4f centroid to the covariance routine)
__global__ void __launch_bounds__(256) obb_moment_kernel(const float4* __restrict__ pts, int64_t n, double* __restrict__ acc) {
  const double cnt = acc[3] > 0 ? acc[3] : 1.0;
  const float cx = (float)(acc[0] / cnt), cy = (float)(acc[1] / cnt), cz = (float)(acc[2] / cnt);
  double m[6] = {0, 0, 0, 0, 0, 0};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 p = __ldg(pts + i);
    if (finite3b(p)) {
      const double px = (double)(p.x - cx), py = (double)(p.y - cy), pz = (double)(p.z - cz);
      m[0] += px * px; m[1] += px * py; m[2] += px * pz; m[3] += py * py; m[4] += py * pz; m[5] += pz * pz;
    }
  }
  __shared__ double sh[8][6];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    const double v = warp_sum(m[c]);
    if (lane == 0) sh[w][c] = v;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double v = 0;
    for (int k = 0; k < 8; ++k) v += sh[k][threadIdx.x];
    atomicAdd(acc + 4 + threadIdx.x, v);
  }
}

__device__ __forceinline__ uint32_t obb_enc(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float obb_dec(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

// one thread: centroid, eigen-solve, right-handed frame; initialises the extent words
__global__ void obb_frame_kernel(const double* __restrict__ acc, pcp_obb* __restrict__ out, uint32_t* __restrict__ ext) {
  const double cnt = acc[3] > 0 ? acc[3] : 1.0;
  float C[6];
  for (int c = 0; c < 6; ++c) C[c] = (float)(acc[4 + c] / cnt);  // {c00, c01, c02, c11, c12, c22}, divisor n (not n - 1)
  Sym3Eig e;
  sym3_eigen(C, e);
  out->centroid[0] = (float)(acc[0] / cnt); out->centroid[1] = (float)(acc[1] / cnt); out->centroid[2] = (float)(acc[2] / cnt);
  out->n_points = (int64_t)acc[3];
  for (int k = 0; k < 3; ++k) out->eigval[k] = e.val[k];
  const float* a = e.vec[0];
  const float* b = e.vec[1];
  const float c3[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};  // col2 = col0 x col1 (:133)
  for (int r = 0; r < 3; ++r) {
    out->axes[0][r] = a[r];
    out->axes[1][r] = b[r];
    out->axes[2][r] = c3[r];
  }
  for (int k = 0; k < 3; ++k) {
    ext[k] = 0xFFFFFFFFu;
    ext[3 + k] = 0u;
  }
}

// extents of R (p - c) over the cloud: the rigid transform of :135-140 (transMatrix = [R | -R c])
__global__ void __launch_bounds__(256) obb_extent_kernel(const float4* __restrict__ pts, int64_t n, const pcp_obb* __restrict__ fr,
                                                         uint32_t* __restrict__ ext) {
  float R[3][3], t[3];
  for (int k = 0; k < 3; ++k) {
    for (int r = 0; r < 3; ++r) R[k][r] = fr->axes[k][r];
    t[k] = -(R[k][0] * fr->centroid[0] + R[k][1] * fr->centroid[1] + R[k][2] * fr->centroid[2]);
  }
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 p = __ldg(pts + i);
    if (!finite3b(p)) continue;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float v = R[k][0] * p.x + R[k][1] * p.y + R[k][2] * p.z + t[k];
      lo[k] = fminf(lo[k], v);
      hi[k] = fmaxf(hi[k], v);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
      hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      atomicMin(ext + k, obb_enc(lo[k]));
      atomicMax(ext + 3 + k, obb_enc(hi[k]));
    }
  }
}

__global__ void obb_finish_kernel(const uint32_t* __restrict__ ext, pcp_obb* __restrict__ out) {
  for (int k = 0; k < 3; ++k) {
    const float lo = obb_dec(ext[k]), hi = obb_dec(ext[3 + k]);
    out->min_pt[k] = lo;
    out->max_pt[k] = hi;
    out->mean_diag[k] = 0.5f * (hi + lo);
  }
}

}  // namespace

extern "C" PCP_API int pcp_bbox_pca(pcp_ctx* ctx, pcp_obb* out) {
  if (!ctx || !out) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_bbox_pca: no cloud set");
  PCP_CUDA(cudaSetDevice(ctx->device));
  int rc;
  const size_t bytes = 10 * sizeof(double) + 8 * sizeof(uint32_t) + sizeof(pcp_obb);
  if ((rc = pcp_reserve(ctx, ctx->obb, bytes))) return rc;
  double* acc = (double*)ctx->obb.p;
  uint32_t* ext = (uint32_t*)(acc + 10);
  pcp_obb* d_out = (pcp_obb*)(ext + 8);
  PCP_CUDA(cudaMemsetAsync(acc, 0, 10 * sizeof(double), ctx->stream));
  const int blocks = (int)std::min<int64_t>((ctx->n + 255) / 256, 148 * 8);
  obb_sum_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_cloud, ctx->n, acc);
  PCP_LAUNCH_CHECK("obb_sum_kernel");
  obb_moment_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_cloud, ctx->n, acc);
  PCP_LAUNCH_CHECK("obb_moment_kernel");
  obb_frame_kernel<<<1, 1, 0, ctx->stream>>>(acc, d_out, ext);
  PCP_LAUNCH_CHECK("obb_frame_kernel");
  obb_extent_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_cloud, ctx->n, d_out, ext);
  PCP_LAUNCH_CHECK("obb_extent_kernel");
  obb_finish_kernel<<<1, 1, 0, ctx->stream>>>(ext, d_out);
  PCP_LAUNCH_CHECK("obb_finish_kernel");
  PCP_CUDA(cudaMemcpyAsync(out, d_out, sizeof(pcp_obb), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}
