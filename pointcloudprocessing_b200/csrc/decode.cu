// decode.cu — Velodyne HDL / VLP-16 data packets -> points, on the device.
//
// Replaces what the reference does on the host with its prebuilt decoder: PacketDecoder::DecodePacket ->
// ProcessHDLPacket -> PushFiringData / SplitFrame (include/PacketDecoder.h:79-121, driven at
// src/pcapReader.cpp:79-102 with 16db.xml).  A frame then enters the GPU as 1206-byte packets (3 bytes per
// return) instead of 16-32 byte points.  Arithmetic as fitted to that binary and pinned by
// tests/golden/decoder_*.npz: d = raw * 0.002; xy = d * cos(v); x = xy * sin(a); y = xy * cos(a);
// z = d * sin(v) in double, a from the 36001-entry azimuth tables ((i / 100.0) * M_PI / 180.0), v the
// laser's vertical correction; zero distances are dropped; a block whose azimuth is below the previous
// block's starts a new frame.  The tables are made by the host libm (as the reference makes its own), so
// the device only multiplies and the doubles come out bit-identical.
//
// One warp per 100-byte block (32 returns = 32 lanes): count pass, one packed exclusive scan
// (returns kept | frame splits << 40), emit pass with ballot ranks — arrival order is preserved.
#include <cub/device/device_scan.cuh>

#include <cmath>

#include "pcp_internal.cuh"

namespace {

constexpr int PKT_BYTES = 1206, BLK_BYTES = 100, BLK_PER_PKT = 12;
constexpr unsigned long long CNT_MASK = (1ull << 40) - 1ull;

struct VertTables {
  double cv[64], sv[64];
};

__device__ __forceinline__ unsigned rd16(const uint8_t* p) { return (unsigned)p[0] | ((unsigned)p[1] << 8); }

__global__ void __launch_bounds__(256) decode_count_kernel(const uint8_t* __restrict__ pk, int64_t n_blocks,
                                                           unsigned long long* __restrict__ packed) {
  const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= n_blocks) return;
  const uint8_t* blk = pk + (b / BLK_PER_PKT) * PKT_BYTES + (b % BLK_PER_PKT) * BLK_BYTES;
  const unsigned raw = rd16(blk + 4 + 3 * lane);
  const unsigned kept = __popc(__ballot_sync(0xffffffffu, raw != 0));
  if (lane == 0) {
    unsigned prev = 0;  // _last_azimuth starts at 0: the first block never splits
    if (b > 0) {
      const int64_t pb = b - 1;
      prev = rd16(pk + (pb / BLK_PER_PKT) * PKT_BYTES + (pb % BLK_PER_PKT) * BLK_BYTES + 2);
    }
    const unsigned az = rd16(blk + 2);
    packed[b] = (unsigned long long)kept | ((az < prev ? 1ull : 0ull) << 40);
  }
}

__global__ void __launch_bounds__(256)
    decode_emit_kernel(const uint8_t* __restrict__ pk, int64_t n_blocks, const unsigned long long* __restrict__ packed,
                       const unsigned long long* __restrict__ excl, VertTables vt, const double* __restrict__ cos_t,
                       const double* __restrict__ sin_t, float4* __restrict__ cloud, double* __restrict__ ox, double* __restrict__ oy,
                       double* __restrict__ oz, double* __restrict__ odist, uint8_t* __restrict__ ointen, uint8_t* __restrict__ olaser,
                       uint16_t* __restrict__ oaz, uint32_t* __restrict__ oms, int64_t* __restrict__ frame_offsets, int64_t max_frames,
                       int64_t* __restrict__ info) {
  const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (b >= n_blocks) return;
  const uint8_t* pkt = pk + (b / BLK_PER_PKT) * PKT_BYTES;
  const uint8_t* blk = pkt + (b % BLK_PER_PKT) * BLK_BYTES;
  const unsigned long long mine = packed[b], before = excl[b];
  const int64_t off = (int64_t)(before & CNT_MASK);
  const int64_t frame = (int64_t)(before >> 40) + (int64_t)(mine >> 40);
  const unsigned flag = rd16(blk), az = rd16(blk + 2);
  const unsigned raw = rd16(blk + 4 + 3 * lane);
  const unsigned m = __ballot_sync(0xffffffffu, raw != 0);
  if (lane == 0) {
    if (b == 0 && max_frames > 0) frame_offsets[0] = 0;
    if ((mine >> 40) && frame < max_frames) frame_offsets[frame] = off;  // this block opens frame `frame`
    if (b == n_blocks - 1) {
      const int64_t total = off + (int64_t)(mine & CNT_MASK);
      if (frame + 1 < max_frames) frame_offsets[frame + 1] = total;
      info[0] = total;   // points
      info[1] = frame;   // complete frames (the frame being filled is frame `frame`)
    }
  }
  if (raw == 0) return;
  const int64_t o = off + __popc(m & ((1u << lane) - 1u));
  const int l = (flag == 0xDDFFu ? 32 : 0) + lane;
  const double d = __dmul_rn((double)raw, 0.002);
  const double xy = __dmul_rn(d, vt.cv[l]);
  const unsigned ai = az > 36000u ? 36000u : az;
  const double x = __dmul_rn(xy, __ldg(sin_t + ai)), y = __dmul_rn(xy, __ldg(cos_t + ai)), z = __dmul_rn(d, vt.sv[l]);
  const uint8_t inten = blk[4 + 3 * lane + 2];
  cloud[o] = make_float4((float)x, (float)y, (float)z, (float)inten);
  if (ox) ox[o] = x;
  if (oy) oy[o] = y;
  if (oz) oz[o] = z;
  if (odist) odist[o] = d;
  if (ointen) ointen[o] = inten;
  if (olaser) olaser[o] = (uint8_t)l;
  if (oaz) oaz[o] = (uint16_t)az;
  if (oms) oms[o] = (uint32_t)pkt[1200] | ((uint32_t)pkt[1201] << 8) | ((uint32_t)pkt[1202] << 16) | ((uint32_t)pkt[1203] << 24);
}

}  // namespace

extern "C" PCP_API int pcp_decode_packets(pcp_ctx* ctx, const pcp_ring_calib* calib, const void* packets, int64_t n_packets,
                                          int packets_on_device, const pcp_decoded* out, int out_is_device, int64_t* frame_offsets,
                                          int64_t max_frames, int64_t* n_frames_complete, int64_t* n_points) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!calib || !packets || n_packets < 1 || n_packets > (int64_t)5000000 || !frame_offsets || max_frames < 2 || !n_frames_complete || !n_points)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_decode_packets: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  int rc;
  const int64_t n_blocks = n_packets * BLK_PER_PKT, cap = n_blocks * 32;
  // azimuth tables: host libm, once per context
  if (!ctx->dec_tables.p) {
    if ((rc = pcp_reserve(ctx, ctx->dec_tables, 2 * 36001 * sizeof(double)))) return rc;
    std::vector<double> t(2 * 36001);
    for (int i = 0; i < 36001; ++i) {
      const double rad = ((double)i / 100.0) * M_PI / 180.0;
      t[i] = cos(rad);
      t[36001 + i] = sin(rad);
    }
    PCP_CUDA(cudaMemcpyAsync(ctx->dec_tables.p, t.data(), t.size() * sizeof(double), cudaMemcpyHostToDevice, st));
    PCP_CUDA(cudaStreamSynchronize(st));
  }
  VertTables vt;
  for (int l = 0; l < 64; ++l) {
    const double rad = ((double)calib->vert_deg[l]) * M_PI / 180.0;
    vt.cv[l] = cos(rad);
    vt.sv[l] = sin(rad);
  }
  const uint8_t* d_pk = (const uint8_t*)packets;
  if (!packets_on_device) {
    if ((rc = pcp_reserve(ctx, ctx->staging, (size_t)n_packets * PKT_BYTES))) return rc;
    pcp_tic(ctx, PCP_T_H2D);
    PCP_CUDA(cudaMemcpyAsync(ctx->staging.p, packets, (size_t)n_packets * PKT_BYTES, cudaMemcpyHostToDevice, st));
    pcp_toc(ctx, PCP_T_H2D);
    d_pk = (const uint8_t*)ctx->staging.p;
  }
  if ((rc = pcp_reserve(ctx, ctx->dec_packed, (size_t)n_blocks * 2 * sizeof(unsigned long long)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->own_cloud, (size_t)cap * sizeof(float4)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->frames, (size_t)(max_frames + 2) * sizeof(int64_t)))) return rc;
  unsigned long long* packed = (unsigned long long*)ctx->dec_packed.p;
  unsigned long long* excl = packed + n_blocks;
  int64_t* d_off = (int64_t*)ctx->frames.p;
  int64_t* d_info = d_off + max_frames;
  // per-return outputs: the caller's device arrays, or ctx-owned staging copied back afterwards
  double *dx = nullptr, *dy = nullptr, *dz = nullptr, *dd = nullptr;
  uint8_t *di = nullptr, *dl = nullptr;
  uint16_t* da = nullptr;
  uint32_t* dm = nullptr;
  if (out && out_is_device) {
    dx = out->x; dy = out->y; dz = out->z; dd = out->distance; di = out->intensity; dl = out->laser_id; da = out->azimuth; dm = out->ms;
  } else if (out) {
    const size_t per = 4 * 8 + 1 + 1 + 2 + 4;
    if ((rc = pcp_reserve(ctx, ctx->dec_out, (size_t)cap * per))) return rc;
    unsigned char* base = (unsigned char*)ctx->dec_out.p;
    dx = (double*)base; dy = dx + cap; dz = dy + cap; dd = dz + cap;
    dm = (uint32_t*)(dd + cap);
    da = (uint16_t*)(dm + cap);
    di = (uint8_t*)(da + cap);
    dl = di + cap;
    if (!out->x) dx = nullptr;
    if (!out->y) dy = nullptr;
    if (!out->z) dz = nullptr;
    if (!out->distance) dd = nullptr;
    if (!out->ms) dm = nullptr;
    if (!out->azimuth) da = nullptr;
    if (!out->intensity) di = nullptr;
    if (!out->laser_id) dl = nullptr;
  }
  const unsigned grid = (unsigned)((n_blocks * 32 + 255) / 256);
  decode_count_kernel<<<grid, 256, 0, st>>>(d_pk, n_blocks, packed);
  PCP_LAUNCH_CHECK("decode_count_kernel");
  size_t tmp = 0;
  PCP_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, packed, excl, (int)n_blocks, st));
  if ((rc = pcp_reserve(ctx, ctx->scan_tmp, tmp))) return rc;
  PCP_CUDA(cub::DeviceScan::ExclusiveSum(ctx->scan_tmp.p, tmp, packed, excl, (int)n_blocks, st));
  ctx->launches += 2;
  const double* tabs = (const double*)ctx->dec_tables.p;
  decode_emit_kernel<<<grid, 256, 0, st>>>(d_pk, n_blocks, packed, excl, vt, tabs, tabs + 36001, (float4*)ctx->own_cloud.p, dx, dy, dz, dd,
                                           di, dl, da, dm, d_off, max_frames, d_info);
  PCP_LAUNCH_CHECK("decode_emit_kernel");
  // sizes back to the host: the caller (and the grid) need them
  int64_t info[2];
  PCP_CUDA(cudaMemcpyAsync(info, d_info, sizeof(info), cudaMemcpyDeviceToHost, st));
  PCP_CUDA(cudaStreamSynchronize(st));
  const int64_t n = info[0], complete = info[1];
  if (complete + 2 > max_frames) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_decode_packets: more frames than max_frames");
  PCP_CUDA(cudaMemcpyAsync(frame_offsets, d_off, (size_t)(complete + 2) * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  if (out && !out_is_device && n > 0) {
    if (out->x) PCP_CUDA(cudaMemcpyAsync(out->x, dx, n * 8, cudaMemcpyDeviceToHost, st));
    if (out->y) PCP_CUDA(cudaMemcpyAsync(out->y, dy, n * 8, cudaMemcpyDeviceToHost, st));
    if (out->z) PCP_CUDA(cudaMemcpyAsync(out->z, dz, n * 8, cudaMemcpyDeviceToHost, st));
    if (out->distance) PCP_CUDA(cudaMemcpyAsync(out->distance, dd, n * 8, cudaMemcpyDeviceToHost, st));
    if (out->ms) PCP_CUDA(cudaMemcpyAsync(out->ms, dm, n * 4, cudaMemcpyDeviceToHost, st));
    if (out->azimuth) PCP_CUDA(cudaMemcpyAsync(out->azimuth, da, n * 2, cudaMemcpyDeviceToHost, st));
    if (out->intensity) PCP_CUDA(cudaMemcpyAsync(out->intensity, di, n, cudaMemcpyDeviceToHost, st));
    if (out->laser_id) PCP_CUDA(cudaMemcpyAsync(out->laser_id, dl, n, cudaMemcpyDeviceToHost, st));
  }
  PCP_CUDA(cudaStreamSynchronize(st));
  *n_points = n;
  *n_frames_complete = complete;
  // the decoded points become the current cloud, its frames the segments of the next grid
  ctx->grid_ok = false;
  ctx->bbox_enqueued = false;
  ctx->d_cloud = (const float4*)ctx->own_cloud.p;
  ctx->d_gid = nullptr;
  ctx->n = ctx->n_query = n;
  const int64_t n_fr = complete + (frame_offsets[complete + 1] > frame_offsets[complete] ? 1 : 0);
  ctx->n_frames = n_fr > 0 ? n_fr : 1;
  ctx->h_frames.assign(frame_offsets, frame_offsets + ctx->n_frames + 1);
  if (n_fr == 0) ctx->h_frames = {0, n};
  return PCP_OK;
}
