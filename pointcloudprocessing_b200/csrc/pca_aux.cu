// pca_aux.cu — the smaller kernels around the fused search: segment PCA and the stale-record
// carry-forward of the reference loop.
#include <cub/device/device_scan.cuh>

#include "eigen3.cuh"
#include "pcp_internal.cuh"

namespace {

// ---- carry-forward (src/pcaAnalysis.cpp:16,26,39,47-51) -------------------------
// `tempFea` lives outside the reference loop, so a record whose neighbourhood is below the
// gate repeats λ / vectors / features of the last valid lower-index record.  The sequential
// dependence becomes: last_valid[i] = max-scan of (valid[i] ? i : −1), then a gather.
struct MaxInt {
  __device__ __forceinline__ int operator()(int a, int b) const { return a > b ? a : b; }
};

__global__ void __launch_bounds__(256) valid_index_kernel(const uint8_t* __restrict__ valid, int* __restrict__ lv, int64_t n) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) lv[i] = valid[i] ? (int)i : -1;
}

// one warp per invalid record: lanes 0..16 copy words, keeping ptID (12) and ptNum (13)
// frame batches (frames != nullptr): every frame is its own reference run, so nothing is carried across
// a frame start
__global__ void __launch_bounds__(256) carry_gather_kernel(float* __restrict__ out, const uint8_t* __restrict__ valid,
                                                           const int* __restrict__ lv, int64_t n, const int64_t* __restrict__ frames,
                                                           int n_frames) {
  int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (i >= n || valid[i]) return;
  int src = lv[i];
  if (src < 0 || lane >= 17 || lane == 12 || lane == 13) return;  // leading records stay zero
  if (frames) {
    int lo = 0, hi = n_frames;  // last f with frames[f] <= i
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(frames + mid) <= i) lo = mid; else hi = mid;
    }
    if ((int64_t)src < __ldg(frames + lo)) return;
  }
  out[i * 17 + lane] = out[(int64_t)src * 17 + lane];
}

// ---- segment PCA (src/pcaAnalysis.cpp:89-121) ------------------------------------------------
// One warp per segment: products of 32 members at a time, the six sums as ordered chains on
// lanes 0-5 (PCL accumulation order); lane 0 then solves and writes the record.
__global__ void __launch_bounds__(128) segment_pca_kernel(const float4* __restrict__ cloud, const float* __restrict__ centres,
                                                          const int64_t* __restrict__ offs, const int32_t* __restrict__ idx,
                                                          int64_t n_seg, float* __restrict__ out) {
  __shared__ float prod[4][6][32];
  __shared__ float mom[4][6];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t s = blockIdx.x * 4 + w;
  if (s >= n_seg) return;
  const int64_t b = offs[s], e = offs[s + 1];
  const float cx = centres[3 * s], cy = centres[3 * s + 1], cz = centres[3 * s + 2];
  float acc = 0.f;
  for (int64_t base = b; base < e; base += 32) {
    int64_t j = base + lane;
    if (j < e) {
      float4 p = __ldg(cloud + idx[j]);
      float px = __fsub_rn(p.x, cx), py = __fsub_rn(p.y, cy), pz = __fsub_rn(p.z, cz);
      prod[w][0][lane] = __fmul_rn(px, px);
      prod[w][1][lane] = __fmul_rn(py, px);
      prod[w][2][lane] = __fmul_rn(pz, px);
      prod[w][3][lane] = __fmul_rn(py, py);
      prod[w][4][lane] = __fmul_rn(py, pz);
      prod[w][5][lane] = __fmul_rn(pz, pz);
    }
    __syncwarp();
    if (lane < 6) {
      int cnt = (int)min((int64_t)32, e - base);
      for (int t = 0; t < cnt; ++t) acc = __fadd_rn(acc, prod[w][lane][t]);
    }
    __syncwarp();
  }
  const int64_t m = e - b;
  if (m > 0) acc = __fdiv_rn(acc, (float)m);
  if (lane < 6) mom[w][lane] = acc;
  __syncwarp();
  if (lane == 0) {
    Sym3Eig eg;
    sym3_eigen(mom[w], eg);
    float* rec = out + s * 17;
    rec[0] = eg.val[0]; rec[1] = eg.val[1]; rec[2] = eg.val[2];
    for (int k = 0; k < 3; ++k)
      for (int r = 0; r < 3; ++r) rec[3 + 3 * k + r] = eg.vec[k][r];
    rec[12] = __int_as_float(0);
    rec[13] = __int_as_float((int)m);
    float s1 = __fsqrt_rn(eg.val[0]), s2 = __fsqrt_rn(eg.val[1]), s3 = __fsqrt_rn(eg.val[2]);
    rec[14] = __fdiv_rn(__fsub_rn(s1, s2), s1);
    rec[15] = __fdiv_rn(__fsub_rn(s2, s3), s1);
    rec[16] = __fdiv_rn(s3, s1);
  }
}

}  // namespace

int pcp_carry_forward_impl(pcp_ctx* ctx, pcp_feature68* d_out, const uint8_t* d_valid, int64_t n) {
  int rc = pcp_reserve(ctx, ctx->last_valid, (size_t)n * sizeof(int));
  if (rc) return rc;
  int* lv = (int*)ctx->last_valid.p;
  pcp_tic(ctx, PCP_T_CARRY);
  valid_index_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_valid, lv, n);
  PCP_LAUNCH_CHECK("valid_index_kernel");
  size_t tmp = 0;
  PCP_CUDA(cub::DeviceScan::InclusiveScan(nullptr, tmp, lv, lv, MaxInt(), (int)n, ctx->stream));
  if ((rc = pcp_reserve(ctx, ctx->scan_tmp, tmp))) return rc;
  PCP_CUDA(cub::DeviceScan::InclusiveScan(ctx->scan_tmp.p, tmp, lv, lv, MaxInt(), (int)n, ctx->stream));
  ctx->launches += 2;
  carry_gather_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, ctx->stream>>>((float*)d_out, d_valid, lv, n,
                                                                                    ctx->n_frames > 1 ? (const int64_t*)ctx->frames.p : nullptr,
                                                                                    (int)ctx->n_frames);
  PCP_LAUNCH_CHECK("carry_gather_kernel");
  pcp_toc(ctx, PCP_T_CARRY);
  return PCP_OK;
}

int pcp_segment_pca_impl(pcp_ctx* ctx, const float* d_centres, const int64_t* d_offs, const int32_t* d_idx, int64_t n_seg,
                         pcp_feature68* d_out) {
  if (n_seg <= 0) return PCP_OK;
  segment_pca_kernel<<<(unsigned)((n_seg + 3) / 4), 128, 0, ctx->stream>>>(ctx->d_cloud, d_centres, d_offs, d_idx, n_seg, (float*)d_out);
  PCP_LAUNCH_CHECK("segment_pca_kernel");
  return PCP_OK;
}
