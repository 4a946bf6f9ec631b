// ring_seg.cu — ground removal and range-image segmentation on the device, one thread block per frame.
//
// Replaces RingProjector::groundRemoval (src/ringProjection.cpp:83-136), cloudSegmentation (:138-197) and
// labelComponents (:199-289) — the steps after projectPointCloud in RingProjector::fullSegmentation
// (include/ringProjection.h:174-179).  The reference floods sequentially from every unlabelled pixel in
// raster order.  Its neighbour table is declared with uint8_t members (include/ringProjection.h:69) while it
// is filled with -1 entries (:119-122), so the flood only ever grows RIGHT (+1 column, wrapping to column
// 0), FAR RIGHT (+255 columns, or column 0 past the end) and DOWN (+1 row): a directed graph.  A pixel's
// label is therefore decided by the smallest raster index among the pixels that can reach it (that pixel
// is a seed, and no earlier seed reaches it) — a min-propagation to a fixed point, done here in shared
// memory with one thread sweeping each 29-pixel piece of a row; component sizes and row masks then give
// the reference's validity rule, and a block scan numbers the valid seeds in raster order.
#include <cfloat>
#include <cmath>

#include "pcp_internal.cuh"

namespace {

constexpr int SEG_THREADS = 1024;
constexpr int SEG_FAR = 255;  // (uint8_t)(-1)

struct SegParams {
  int n_scan, horizon, ground_scan_ind, valid_points, valid_lines;
  float mount_angle, theta, sin_x, cos_x, sin_y, cos_y;
};

__device__ __forceinline__ float atan2f_cr(float y, float x) { return (float)atan2((double)y, (double)x); }

// angle = atan2(d2*sin(alpha), d1 - d2*cos(alpha)) > theta   (src/ringProjection.cpp:234-248)
__device__ __forceinline__ bool grows(float ra, float rb, float s, float c, float theta) {
  const float d1 = fmaxf(ra, rb), d2 = fminf(ra, rb);
  return atan2f_cr(__fmul_rn(d2, s), __fsub_rn(d1, __fmul_rn(d2, c))) > theta;
}

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* warp_tot, uint32_t& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = warp_tot[lane], winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    warp_tot[lane] = winc - w;
    if (lane == 31) warp_tot[32] = winc;
  }
  __syncthreads();
  total = warp_tot[32];
  return warp_tot[warp] + inc - v;
}

__global__ void __launch_bounds__(SEG_THREADS, 1)
    ring_segment_kernel(const float4* __restrict__ pts, const int64_t* __restrict__ frame_offsets, SegParams sp,
                        const float* __restrict__ range_img, const int32_t* __restrict__ pt_index, uint32_t* __restrict__ stat,
                        int8_t* __restrict__ ground_out, int32_t* __restrict__ label_out, float* __restrict__ colmax_out,
                        int32_t* __restrict__ above_idx, int32_t* __restrict__ above_count, int32_t* __restrict__ n_labels) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int H = sp.horizon, NS = sp.n_scan, cells = H * NS;
  uint32_t* sm_m = reinterpret_cast<uint32_t*>(smem);                       // min raster index reaching the pixel
  float* sm_cmax = reinterpret_cast<float*>(sm_m + cells);                  // max ground range per column
  uint32_t* sm_scan = reinterpret_cast<uint32_t*>(sm_cmax + H);             // 33 words
  uint8_t* sm_e = reinterpret_cast<uint8_t*>(sm_scan + 40);                 // edge bits 1|2|4, free 0x80
  int8_t* sm_g = reinterpret_cast<int8_t*>(sm_e + cells);                   // groundMat
  const int tid = threadIdx.x, f = blockIdx.x;
  const int64_t beg = frame_offsets ? frame_offsets[f] : 0;
  const float* rng = range_img + (size_t)f * cells;
  const int32_t* pti = pt_index + (size_t)f * cells;
  stat += (size_t)f * cells;

  // ---- groundRemoval: one thread per column, pairs of adjacent rings bottom-up (:88-114), then the
  // column's largest ground range (:118-128)
  for (int c = tid; c < cells; c += SEG_THREADS) sm_g[c] = 0;
  __syncthreads();
  for (int j = tid; j < H; j += SEG_THREADS) {
    for (int i = 0; i < sp.ground_scan_ind && i + 1 < NS; ++i) {
      const int lower = j + i * H, upper = lower + H;
      const int32_t wl = pti[lower], wu = pti[upper];
      if (wl < 0 || wu < 0) {
        sm_g[lower] = -1;
        continue;
      }
      const float4 pl = __ldg(pts + beg + wl), pu = __ldg(pts + beg + wu);
      const float dx = __fsub_rn(pu.x, pl.x), dy = __fsub_rn(pu.y, pl.y), dz = __fsub_rn(pu.z, pl.z);
      const float hyp = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
      const float angle = (float)((double)__fmul_rn(atan2f_cr(dz, hyp), 180.0f) / 3.14159265358979323846);
      if (fabsf(__fsub_rn(angle, sp.mount_angle)) <= 10.0f) {
        sm_g[lower] = 1;
        sm_g[upper] = 1;
      }
    }
    float mx = 0.f;  // maxGroundrangeOfcol: uninitialised upstream, defined as 0
    for (int i = 0; i < NS; ++i) {
      const int c = j + i * H;
      if (sm_g[c] == 1) mx = fmaxf(mx, rng[c]);
    }
    sm_cmax[j] = mx;
    colmax_out[(size_t)f * H + j] = mx;
  }
  __syncthreads();
  // ---- free pixels (labelMat == 0 after groundRemoval), then the directed edges of the flood
  for (int c = tid; c < cells; c += SEG_THREADS) sm_e[c] = (sm_g[c] != 1 && rng[c] != FLT_MAX) ? 0x80 : 0;
  __syncthreads();
  for (int c = tid; c < cells; c += SEG_THREADS) {
    uint32_t m = 0xFFFFFFFFu;
    if (sm_e[c] & 0x80) {
      const int i = c / H, j = c - i * H;
      const float r = rng[c];
      uint8_t e = 0x80;
      const int tr = i * H + (j + 1 >= H ? 0 : j + 1);
      if ((sm_e[tr] & 0x80) && grows(r, rng[tr], sp.sin_x, sp.cos_x, sp.theta)) e |= 1;
      const int tf = i * H + (j + SEG_FAR >= H ? 0 : j + SEG_FAR);
      if ((sm_e[tf] & 0x80) && grows(r, rng[tf], sp.sin_x, sp.cos_x, sp.theta)) e |= 2;
      if (i + 1 < NS && (sm_e[c + H] & 0x80) && grows(r, rng[c + H], sp.sin_y, sp.cos_y, sp.theta)) e |= 4;
      sm_e[c] = e;  // other threads only test bit 0x80, which does not change
      m = (uint32_t)c;
    }
    sm_m[c] = m;
  }
  __syncthreads();
  // ---- min-propagation along the edges to a fixed point; thread t sweeps one piece of one row left to right
  const int cpr = SEG_THREADS / NS;              // pieces per row
  const int plen = (H + cpr - 1) / cpr;          // 29 for 1800 / 64: odd, so the sweeps are bank-conflict free
  const int prow = tid / cpr, pcol0 = (tid - prow * cpr) * plen;
  const bool has_piece = prow < NS && pcol0 < H;
  const int pend = has_piece ? min(pcol0 + plen, H) : 0;
  for (;;) {
    bool changed = false;
    if (has_piece) {
      for (int j = pcol0; j < pend; ++j) {
        const int c = prow * H + j;
        const uint8_t e = sm_e[c];
        if ((e & 7) == 0) continue;
        const uint32_t v = *reinterpret_cast<volatile uint32_t*>(sm_m + c);
        if (e & 1) changed |= atomicMin(&sm_m[prow * H + (j + 1 >= H ? 0 : j + 1)], v) > v;
        if (e & 2) changed |= atomicMin(&sm_m[prow * H + (j + SEG_FAR >= H ? 0 : j + SEG_FAR)], v) > v;
        if (e & 4) changed |= atomicMin(&sm_m[c + H], v) > v;
      }
    }
    if (!__syncthreads_or(changed ? 1 : 0)) break;
  }
  // ---- component statistics at the seed: size (low 16 bits), rows of the pushed pixels (high 16 bits;
  // the seed itself is never "pushed", :254)
  for (int c = tid; c < cells; c += SEG_THREADS) stat[c] = 0;
  __syncthreads();
  for (int c = tid; c < cells; c += SEG_THREADS) {
    if (sm_e[c] & 0x80) {
      const uint32_t s = sm_m[c];
      atomicAdd(&stat[s], 1u);
      if (s != (uint32_t)c) atomicOr(&stat[s], 1u << (16 + c / H));
    }
  }
  __syncthreads();
  // ---- validity (:264-276) and numbering of the valid seeds in raster order (labelCount, :278-279)
  uint32_t valid_bits = 0, n_valid = 0;
  if (has_piece) {
    for (int j = pcol0; j < pend; ++j) {
      const int c = prow * H + j;
      if ((sm_e[c] & 0x80) && sm_m[c] == (uint32_t)c) {
        const uint32_t st = __ldcg(stat + c);
        const int size = (int)(st & 0xFFFFu), lines = __popc(st >> 16);
        if (size >= 30 || (size >= sp.valid_points && lines >= sp.valid_lines)) {
          valid_bits |= 1u << (j - pcol0);
          ++n_valid;
        }
      }
    }
  }
  uint32_t total = 0;
  uint32_t next = block_exclusive_scan(n_valid, sm_scan, total) + 1;
  if (has_piece) {
    for (int j = pcol0; j < pend; ++j) {
      const int c = prow * H + j;
      if ((sm_e[c] & 0x80) && sm_m[c] == (uint32_t)c) __stcg(stat + c, (valid_bits >> (j - pcol0)) & 1u ? next++ : 999999u);
    }
  }
  if (tid == 0) n_labels[f] = (int32_t)total;
  __syncthreads();
  for (int c = tid; c < cells; c += SEG_THREADS) {
    label_out[(size_t)f * cells + c] = (sm_e[c] & 0x80) ? (int32_t)__ldcg(stat + sm_m[c]) : -1;
    ground_out[(size_t)f * cells + c] = sm_g[c];
  }
  // ---- cloudaboveGround (:186-189): not ground and nearer than the column's farthest ground return,
  // original point indices in raster order (frame-local)
  uint32_t n_above = 0;
  if (has_piece)
    for (int j = pcol0; j < pend; ++j) {
      const int c = prow * H + j;
      n_above += (sm_g[c] != 1 && rng[c] < sm_cmax[j]) ? 1u : 0u;
    }
  uint32_t at = block_exclusive_scan(n_above, sm_scan, total);
  if (has_piece)
    for (int j = pcol0; j < pend; ++j) {
      const int c = prow * H + j;
      if (sm_g[c] != 1 && rng[c] < sm_cmax[j]) above_idx[(size_t)f * cells + at++] = pti[c];
    }
  if (tid == 0) above_count[f] = (int32_t)total;
}

}  // namespace

int pcp_ring_segment_impl(pcp_ctx* ctx, const pcp_ring_calib* calib, const pcp_ring_seg_params* prm, const int64_t* d_offsets,
                          int64_t n_frames, const float* d_range, const int32_t* d_index, int8_t* d_ground, int32_t* d_label,
                          float* d_colmax, int32_t* d_above, int32_t* d_above_count, int32_t* d_n_labels) {
  SegParams sp;
  sp.n_scan = calib->n_scan;
  sp.horizon = calib->horizon_scan;
  sp.ground_scan_ind = prm->ground_scan_ind;
  sp.valid_points = prm->segment_valid_point_num;
  sp.valid_lines = prm->segment_valid_line_num;
  sp.mount_angle = prm->sensor_mount_angle;
  sp.theta = prm->segment_theta;
  // segmentAlphaX/Y = ang_res / 180.0 * M_PI narrowed to float (include/ringProjection.h:27-28); sin/cos
  // of a float pick the float overloads, evaluated here by the host libm like the reference does
  const float ax = (float)((double)calib->ang_res_x / 180.0 * M_PI), ay = (float)((double)calib->ang_res_y / 180.0 * M_PI);
  sp.sin_x = sinf(ax);
  sp.cos_x = cosf(ax);
  sp.sin_y = sinf(ay);
  sp.cos_y = cosf(ay);
  const size_t cells = (size_t)sp.n_scan * sp.horizon;
  const size_t smem = cells * 4 + (size_t)sp.horizon * 4 + 40 * 4 + cells * 2 + 16;
  if (sp.n_scan < 1 || sp.n_scan > 16 || cells > 65535 || smem > 227 * 1024)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_ring_segment: image must have at most 16 rings and fit one block's shared memory");
  int rc = pcp_reserve(ctx, ctx->ring_key, (size_t)n_frames * cells * sizeof(uint32_t));
  if (rc) return rc;
  PCP_CUDA(cudaMemsetAsync(d_above, 0xFF, (size_t)n_frames * cells * sizeof(int32_t), ctx->stream));
  PCP_CUDA(cudaFuncSetAttribute(ring_segment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ring_segment_kernel<<<(unsigned)n_frames, SEG_THREADS, smem, ctx->stream>>>(ctx->d_cloud, d_offsets, sp, d_range, d_index,
                                                                               (uint32_t*)ctx->ring_key.p, d_ground, d_label, d_colmax,
                                                                               d_above, d_above_count, d_n_labels);
  PCP_LAUNCH_CHECK("ring_segment_kernel");
  return PCP_OK;
}
