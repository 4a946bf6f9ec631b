// halo.cu — halo classification for slab-sharded clouds (BASELINE configs[3]).
//
// A rank owns the slab [lo_face, hi_face) along one axis of a big cloud and must hand its two slab
// neighbours the points within `h` of the shared faces.  One kernel classifies every own point and
// appends it (float4 + global index) to the fixed-capacity send buffer of the face(s) it is near; the
// unused tail of each buffer is filled with non-finite points, which the grid puts in its trash cell:
// they are nobody's neighbour.  Fixed capacity means no count has to reach the host before the
// exchange and the grid build can be enqueued — the step has no host synchronisation at all.
// The reference has no counterpart (single process, src/pcaAnalysis.cpp:11-42); this is the exchange
// step of the multi-GPU form of that loop.
#include <cfloat>
#include <cstring>

#include "pcp_internal.cuh"

namespace {

__global__ void halo_pack_kernel(const float4* __restrict__ pts, const int32_t* __restrict__ gid, int64_t n, int axis, float lo_edge,
                                 float hi_edge, int64_t cap, float4* __restrict__ lo_pts, int32_t* __restrict__ lo_gid,
                                 float4* __restrict__ hi_pts, int32_t* __restrict__ hi_gid, unsigned int* __restrict__ counts) {
  const int lane = threadIdx.x & 31;
  for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; i0 < n; i0 += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = i0 + lane;
    float4 p = make_float4(0.f, 0.f, 0.f, 0.f);
    bool to_lo = false, to_hi = false;
    if (i < n) {
      p = pts[i];
      const float c = axis == 0 ? p.x : (axis == 1 ? p.y : p.z);
      to_lo = c < lo_edge;    // within h of the low face  (lo_edge = lo_face + h)
      to_hi = c >= hi_edge;   // within h of the high face (hi_edge = hi_face - h)
    }
    const int g = i < n ? (gid ? gid[i] : (int32_t)i) : 0;
    // warp-aggregated append: one atomic per warp and face
    const unsigned ml = __ballot_sync(0xffffffffu, to_lo), mh = __ballot_sync(0xffffffffu, to_hi);
    unsigned bl = 0, bh = 0;
    if (lane == 0) {
      if (ml) bl = atomicAdd(&counts[0], (unsigned)__popc(ml));
      if (mh) bh = atomicAdd(&counts[1], (unsigned)__popc(mh));
    }
    bl = __shfl_sync(0xffffffffu, bl, 0);
    bh = __shfl_sync(0xffffffffu, bh, 0);
    if (to_lo) {
      const int64_t o = (int64_t)bl + __popc(ml & ((1u << lane) - 1u));
      if (o < cap) {
        lo_pts[o] = p;
        lo_gid[o] = g;
      }
    }
    if (to_hi) {
      const int64_t o = (int64_t)bh + __popc(mh & ((1u << lane) - 1u));
      if (o < cap) {
        hi_pts[o] = p;
        hi_gid[o] = g;
      }
    }
  }
}

// entries [count, cap) of both buffers <- non-finite padding
__global__ void halo_pad_kernel(int64_t cap, float4* __restrict__ lo_pts, int32_t* __restrict__ lo_gid, float4* __restrict__ hi_pts,
                                int32_t* __restrict__ hi_gid, const unsigned int* __restrict__ counts) {
  const float qnan = __int_as_float(0x7FC00000);
  const float4 pad = make_float4(qnan, qnan, qnan, 0.f);
  const int64_t c0 = min((int64_t)counts[0], cap), c1 = min((int64_t)counts[1], cap);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += (int64_t)gridDim.x * blockDim.x) {
    if (i >= c0) {
      lo_pts[i] = pad;
      lo_gid[i] = 0x7fffffff;
    }
    if (i >= c1) {
      hi_pts[i] = pad;
      hi_gid[i] = 0x7fffffff;
    }
  }
}

// Peer form of the padding kernel: the strips went straight into the neighbours' halo regions (peer
// stores over NVLink); the LAST block to finish fences system-wide and raises the neighbours' flags to
// `epoch` — the receive side of the exchange is a flag, not a message.
__global__ void halo_pad_signal_kernel(int64_t cap, float4* __restrict__ lo_pts, int32_t* __restrict__ lo_gid, float4* __restrict__ hi_pts,
                                       int32_t* __restrict__ hi_gid, const unsigned int* __restrict__ counts, unsigned int* __restrict__ done,
                                       unsigned int* lo_flag, unsigned int* hi_flag, unsigned int epoch) {
  const float qnan = __int_as_float(0x7FC00000);
  const float4 pad = make_float4(qnan, qnan, qnan, 0.f);
  const int64_t c0 = min((int64_t)counts[0], cap), c1 = min((int64_t)counts[1], cap);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += (int64_t)gridDim.x * blockDim.x) {
    if (i >= c0) {
      lo_pts[i] = pad;
      lo_gid[i] = 0x7fffffff;
    }
    if (i >= c1) {
      hi_pts[i] = pad;
      hi_gid[i] = 0x7fffffff;
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {
      *done = 0;  // ready for the next launch
      __threadfence_system();
      if (lo_flag) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(lo_flag), "r"(epoch) : "memory");
      if (hi_flag) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(hi_flag), "r"(epoch) : "memory");
    }
  }
}

// One thread polls up to two flags until both reached `epoch`; gives up after timeout_ns and leaves 1 in
// *err (a neighbour that died must not hang this GPU).
__global__ void halo_wait_kernel(const unsigned int* flag_a, const unsigned int* flag_b, unsigned int epoch,
                                 unsigned long long timeout_ns, unsigned int* err) {
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    unsigned int a = epoch, b = epoch;
    if (flag_a) asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(a) : "l"(flag_a) : "memory");
    if (flag_b) asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(b) : "l"(flag_b) : "memory");
    if ((int)(a - epoch) >= 0 && (int)(b - epoch) >= 0) return;
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) {
      atomicExch(err, 1u);
      return;
    }
    __nanosleep(200);
  }
}

}  // namespace

extern "C" PCP_API int pcp_halo_push(pcp_ctx* ctx, const void* d_xyzi, const int32_t* d_global_id, int64_t n, int axis, float lo_face,
                                     float hi_face, float h, int64_t cap, void* p_lo_pts, int32_t* p_lo_gid, void* p_hi_pts,
                                     int32_t* p_hi_gid, uint32_t* d_counts, uint32_t* p_lo_flag, uint32_t* p_hi_flag, uint32_t epoch) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!d_xyzi || n < 0 || axis < 0 || axis > 2 || cap <= 0 || !p_lo_pts || !p_lo_gid || !p_hi_pts || !p_hi_gid || !d_counts || !(h >= 0.f))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_halo_push: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  int rc;
  const bool fresh = ctx->halo_done.p == nullptr;
  if ((rc = pcp_reserve(ctx, ctx->halo_done, sizeof(uint32_t)))) return rc;
  if (fresh) PCP_CUDA(cudaMemsetAsync(ctx->halo_done.p, 0, sizeof(uint32_t), ctx->stream));
  PCP_CUDA(cudaMemsetAsync(d_counts, 0, 2 * sizeof(uint32_t), ctx->stream));
  if (n > 0) {
    const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
    halo_pack_kernel<<<blocks, 256, 0, ctx->stream>>>((const float4*)d_xyzi, d_global_id, n, axis, lo_face + h, hi_face - h, cap,
                                                      (float4*)p_lo_pts, p_lo_gid, (float4*)p_hi_pts, p_hi_gid, d_counts);
    PCP_LAUNCH_CHECK("halo_pack_kernel");
  }
  const int blocks = (int)std::min<int64_t>((cap + 255) / 256, 148 * 2);
  halo_pad_signal_kernel<<<blocks, 256, 0, ctx->stream>>>(cap, (float4*)p_lo_pts, p_lo_gid, (float4*)p_hi_pts, p_hi_gid, d_counts,
                                                          (unsigned int*)ctx->halo_done.p, p_lo_flag, p_hi_flag, epoch);
  PCP_LAUNCH_CHECK("halo_pad_signal_kernel");
  return PCP_OK;
}

extern "C" PCP_API int pcp_halo_wait(pcp_ctx* ctx, const uint32_t* d_flag_a, const uint32_t* d_flag_b, uint32_t epoch, int timeout_ms,
                                     uint32_t* d_err) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!d_err || timeout_ms < 0) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_halo_wait: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  if (!d_flag_a && !d_flag_b) return PCP_OK;
  halo_wait_kernel<<<1, 1, 0, ctx->stream>>>(d_flag_a, d_flag_b, epoch, (unsigned long long)timeout_ms * 1000000ull, d_err);
  PCP_LAUNCH_CHECK("halo_wait_kernel");
  return PCP_OK;
}

// Memory another process can map: plain cudaMalloc + CUDA IPC handle (64 bytes).
extern "C" PCP_API int pcp_peer_alloc(pcp_ctx* ctx, size_t bytes, void** d_ptr, void* handle64) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!d_ptr || !handle64 || bytes == 0) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_peer_alloc: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) return pcp_fail(ctx, PCP_ERR_OOM, "pcp_peer_alloc: cudaMalloc", e);
  cudaIpcMemHandle_t hd;
  e = cudaIpcGetMemHandle(&hd, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return pcp_fail(ctx, PCP_ERR_CUDA, "pcp_peer_alloc: cudaIpcGetMemHandle", e);
  }
  memcpy(handle64, &hd, 64);
  *d_ptr = p;
  return PCP_OK;
}
extern "C" PCP_API int pcp_peer_free(pcp_ctx* ctx, void* d_ptr) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  PCP_CUDA(cudaSetDevice(ctx->device));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  PCP_CUDA(cudaFree(d_ptr));
  return PCP_OK;
}
extern "C" PCP_API int pcp_peer_open(pcp_ctx* ctx, const void* handle64, void** d_ptr) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!d_ptr || !handle64) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_peer_open: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  cudaIpcMemHandle_t hd;
  memcpy(&hd, handle64, 64);
  PCP_CUDA(cudaIpcOpenMemHandle(d_ptr, hd, cudaIpcMemLazyEnablePeerAccess));
  return PCP_OK;
}
extern "C" PCP_API int pcp_peer_close(pcp_ctx* ctx, void* d_ptr) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  PCP_CUDA(cudaSetDevice(ctx->device));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  PCP_CUDA(cudaIpcCloseMemHandle(d_ptr));
  return PCP_OK;
}

extern "C" PCP_API int pcp_halo_pack(pcp_ctx* ctx, const void* d_xyzi, const int32_t* d_global_id, int64_t n, int axis, float lo_face,
                                     float hi_face, float h, int64_t cap, void* d_lo_pts, int32_t* d_lo_gid, void* d_hi_pts,
                                     int32_t* d_hi_gid, uint32_t* d_counts) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!d_xyzi || n < 0 || axis < 0 || axis > 2 || cap < 0 || !d_lo_pts || !d_lo_gid || !d_hi_pts || !d_hi_gid || !d_counts || !(h >= 0.f))
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_halo_pack: bad argument");
  PCP_CUDA(cudaSetDevice(ctx->device));
  PCP_CUDA(cudaMemsetAsync(d_counts, 0, 2 * sizeof(uint32_t), ctx->stream));
  if (n > 0) {
    const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
    halo_pack_kernel<<<blocks, 256, 0, ctx->stream>>>((const float4*)d_xyzi, d_global_id, n, axis, lo_face + h, hi_face - h, cap,
                                                      (float4*)d_lo_pts, d_lo_gid, (float4*)d_hi_pts, d_hi_gid, d_counts);
    PCP_LAUNCH_CHECK("halo_pack_kernel");
  }
  if (cap > 0) {
    const int blocks = (int)std::min<int64_t>((cap + 255) / 256, 148 * 4);
    halo_pad_kernel<<<blocks, 256, 0, ctx->stream>>>(cap, (float4*)d_lo_pts, d_lo_gid, (float4*)d_hi_pts, d_hi_gid, d_counts);
    PCP_LAUNCH_CHECK("halo_pad_kernel");
  }
  return PCP_OK;
}
