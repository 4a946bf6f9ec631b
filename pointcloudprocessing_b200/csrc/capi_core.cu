// capi_core.cu — lifetime, buffers, timing and cloud hand-over of the pcp_* C-ABI (include/pcp.h).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "pcp_internal.cuh"

int pcp_fail(pcp_ctx* ctx, int status, const char* what, cudaError_t ce) {
  if (ctx) {
    ctx->err = what ? what : "";
    if (ce != cudaSuccess) {
      ctx->err += ": ";
      ctx->err += cudaGetErrorString(ce);
    }
  }
  return status;
}

int pcp_reserve(pcp_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap && b.p) return PCP_OK;
  if (b.p) {
    cudaStreamSynchronize(ctx->stream);
    cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
  }
  size_t want = bytes < 256 ? 256 : bytes;
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    b.p = nullptr;
    return pcp_fail(ctx, PCP_ERR_OOM, "cudaMalloc", e);
  }
  b.cap = want;
  if (!b.listed) {
    b.listed = true;
    ctx->owned.push_back(&b);
  }
  return PCP_OK;
}

void pcp_tic(pcp_ctx* ctx, int stage) { cudaEventRecord(ctx->ev[2 * stage], ctx->stream); }
void pcp_toc(pcp_ctx* ctx, int stage) {
  cudaEventRecord(ctx->ev[2 * stage + 1], ctx->stream);
  ctx->ev_used[stage] = true;
}
void pcp_collect_timings(pcp_ctx* ctx) {
  for (int s = 0; s < PCP_T_COUNT; ++s) {
    if (!ctx->ev_used[s]) continue;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->ev[2 * s], ctx->ev[2 * s + 1]) == cudaSuccess) ctx->t_ms[s] = ms;
  }
}

static void free_buf(DevBuf& b) {
  if (b.p) cudaFree(b.p);
  b.p = nullptr;
  b.cap = 0;
}

// the environment is looked at here and nowhere else
static void read_knobs(PcpKnobs& kn) {
  if (const char* e = getenv("PCP_KNN_IMPL")) kn.knn_impl = !strcmp(e, "v3") ? 3 : (!strcmp(e, "tile4") ? 4 : (!strcmp(e, "tile6") ? 6 : 0));
  if (const char* e = getenv("PCP_KNN_GROUPS")) kn.knn_groups = atoi(e);
  if (const char* e = getenv("PCP_RADIUS_IMPL")) kn.radius_impl = !strcmp(e, "v3") ? 3 : (!strcmp(e, "tile64") ? 2 : (!strncmp(e, "tile", 4) ? 1 : 0));
  if (const char* e = getenv("PCP_WL_BLOCKS")) kn.wl_blocks = atoi(e) > 0 ? atoi(e) : 0;
  if (const char* e = getenv("PCP_WL_TILE")) kn.wl_tile = (atoi(e) >= 1 && atoi(e) <= 32) ? atoi(e) : 0;
  if (const char* e = getenv("PCP_GRID_SHIFT")) kn.grid_shift = atof(e) < 0.0 ? 0.0 : (atof(e) > 0.999 ? 0.999 : atof(e));
  if (const char* e = getenv("PCP_GRID_SHIFT_AXIS")) kn.grid_shift_axis = atoi(e);
  if (const char* e = getenv("PCP_RADIUS_T6")) kn.radius_t6 = atoi(e) != 0 ? 1 : 0;
  if (const char* e = getenv("PCP_T6_BIN")) kn.t6_bin = !strcmp(e, "sqrt") ? 0 : (!strcmp(e, "log") ? 1 : -1);
  if (const char* e = getenv("PCP_T6_REFINE_MIN")) kn.t6_refine_min = atoi(e) > 0 ? atoi(e) : 0;
  if (const char* e = getenv("PCP_T6_ORDER")) kn.t6_no_order = atoi(e) == 0 ? 1 : 0;
  if (const char* e = getenv("PCP_PRIO")) kn.flat_priority = !strcmp(e, "flat") ? 1 : (!strcmp(e, "two") ? 3 : 0);
  if (const char* e = getenv("PCP_SHARD_SORT")) kn.shard_full_sort = !strcmp(e, "full") ? 1 : (!strcmp(e, "rank") ? 2 : 0);
}

extern "C" {

PCP_API const char* pcp_version(void) { return "pcp 0.1 (sm_100a)"; }

PCP_API int pcp_create(pcp_ctx** out, int device) {
  if (!out) return PCP_ERR_INVALID_ARG;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return PCP_ERR_NO_DEVICE;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return PCP_ERR_NO_DEVICE;
  if (prop.major != 10) return PCP_ERR_NO_DEVICE;  // kernels are built for sm_100a only
  if (cudaSetDevice(device) != cudaSuccess) return PCP_ERR_CUDA;
  pcp_ctx* ctx = new (std::nothrow) pcp_ctx();
  if (!ctx) return PCP_ERR_OOM;
  ctx->device = device;
  read_knobs(ctx->knobs);
  int prio_least = 0, prio_greatest = 0;
  cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
  const int prio_stream = (ctx->knobs.flat_priority == 1 || ctx->knobs.flat_priority == 3) ? prio_greatest : 0;
  if (cudaStreamCreateWithPriority(&ctx->stream, cudaStreamNonBlocking, prio_stream) != cudaSuccess) {
    delete ctx;
    return PCP_ERR_CUDA;
  }
  ctx->prio_search = ctx->knobs.flat_priority == 3 ? prio_least : prio_stream;
  ctx->prio_high = prio_greatest;
  for (int i = 0; i < 2 * PCP_T_COUNT; ++i) cudaEventCreate(&ctx->ev[i]);
  *out = ctx;
  return PCP_OK;
}

PCP_API int pcp_destroy(pcp_ctx* ctx) {
  if (!ctx) return PCP_OK;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (DevBuf* b : ctx->owned) free_buf(*b);  // everything pcp_reserve ever allocated, whichever call asked for it
  if (ctx->h_occ) cudaFreeHost(ctx->h_occ);
  if (ctx->occ_ev) cudaEventDestroy(ctx->occ_ev);
  for (int i = 0; i < 2 * PCP_T_COUNT; ++i)
    if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
  return PCP_OK;
}

PCP_API const char* pcp_last_error(const pcp_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
PCP_API void* pcp_stream(pcp_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
PCP_API int64_t pcp_launch_count(const pcp_ctx* ctx) { return ctx ? ctx->launches : 0; }

PCP_API int pcp_synchronize(pcp_ctx* ctx) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCP_OK;
}

// page-locked host memory for the asynchronous host-pointer calls (plain cudaHostAlloc; no ctx needed)
PCP_API void* pcp_alloc_pinned(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
PCP_API void pcp_free_pinned(void* p) {
  if (p) cudaFreeHost(p);
}

PCP_API int pcp_set_host_async(pcp_ctx* ctx, int enable) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  ctx->host_async = enable != 0;
  return PCP_OK;
}

PCP_API int pcp_last_timings(pcp_ctx* ctx, float* ms, int n) {
  if (!ctx || !ms) return PCP_ERR_INVALID_ARG;
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  pcp_collect_timings(ctx);
  for (int i = 0; i < n && i < PCP_T_COUNT; ++i) ms[i] = ctx->t_ms[i];
  return PCP_OK;
}

// host records of `words` floats each -> packed float4 (the record's first four floats; 0 for the fourth when the record has three)
__global__ void pack_cloud_kernel(const float* __restrict__ src, int64_t n, int words, float4* __restrict__ dst) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float* r = src + i * words;
    const float w = words >= 4 ? r[3] : 0.f;  // the fourth float travels along; no kernel reads it
    dst[i] = make_float4(r[0], r[1], r[2], w);
  }
}

// strided host AoS -> packed float4 on the device
PCP_API int pcp_set_cloud_host(pcp_ctx* ctx, const void* pts, int64_t n, int stride_bytes) {
  if (!ctx || (!pts && n > 0) || n < 0 || stride_bytes < 12) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_cloud_host");
  PCP_CUDA(cudaSetDevice(ctx->device));
  ctx->grid_ok = false;
  ctx->bbox_enqueued = false;
  ctx->n = ctx->n_query = n;
  ctx->d_gid = nullptr;
  ctx->n_frames = 1;
  int rc = pcp_reserve(ctx, ctx->own_cloud, (size_t)std::max<int64_t>(n, 1) * sizeof(float4));
  if (rc) return rc;
  ctx->d_cloud = (const float4*)ctx->own_cloud.p;
  if (n == 0) return PCP_OK;
  pcp_tic(ctx, PCP_T_H2D);
  if (stride_bytes == 16) {
    // already packed: one linear copy
    PCP_CUDA(cudaMemcpyAsync(ctx->own_cloud.p, pts, (size_t)n * 16, cudaMemcpyHostToDevice, ctx->stream));
  } else if (stride_bytes % 4 == 0) {
    // strided AoS (32 B for pcl::PointXYZI, 12 B for packed xyz): upload the records as they are with ONE
    // linear copy and pack on the device — a 2-D copy of 16-byte rows runs at a quarter of the PCIe rate
    if ((rc = pcp_reserve(ctx, ctx->staging, (size_t)n * (size_t)stride_bytes))) return rc;
    PCP_CUDA(cudaMemcpyAsync(ctx->staging.p, pts, (size_t)n * (size_t)stride_bytes, cudaMemcpyHostToDevice, ctx->stream));
    const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
    pack_cloud_kernel<<<blocks, 256, 0, ctx->stream>>>((const float*)ctx->staging.p, n, stride_bytes / 4, (float4*)ctx->own_cloud.p);
    PCP_LAUNCH_CHECK("pack_cloud_kernel");
  } else if (stride_bytes > 16) {
    PCP_CUDA(cudaMemcpy2DAsync(ctx->own_cloud.p, 16, pts, (size_t)stride_bytes, 16, (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    // packed xyz at an odd stride: widen on the way
    PCP_CUDA(cudaMemsetAsync(ctx->own_cloud.p, 0, (size_t)n * 16, ctx->stream));
    PCP_CUDA(cudaMemcpy2DAsync(ctx->own_cloud.p, 16, pts, (size_t)stride_bytes, 12, (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
  }
  pcp_toc(ctx, PCP_T_H2D);
  if (!ctx->host_async) PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  else if ((rc = pcp_bbox_enqueue(ctx))) return rc;  // the box travels with the upload: pcp_build_grid will not have to wait for it
  return PCP_OK;
}

PCP_API int pcp_set_cloud_device_sharded(pcp_ctx* ctx, const void* d_xyzi, const int32_t* d_global_id, int64_t n, int64_t n_query) {
  if (!ctx || (!d_xyzi && n > 0) || n < 0 || n_query < 0 || n_query > n) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_cloud_device");
  PCP_CUDA(cudaSetDevice(ctx->device));
  ctx->grid_ok = false;
  ctx->bbox_enqueued = false;
  ctx->d_cloud = (const float4*)d_xyzi;
  ctx->d_gid = d_global_id;
  ctx->n = n;
  ctx->n_query = n_query;
  ctx->n_frames = 1;
  return PCP_OK;
}

// Frame batches: the current cloud is n_frames independent clouds laid end to end; the next pcp_build_grid
// builds one segmented grid over all of them (searches never cross a frame).
PCP_API int pcp_set_frames(pcp_ctx* ctx, const int64_t* frame_offsets, int64_t n_frames) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_set_frames: no cloud set");
  if (!frame_offsets || n_frames < 1 || n_frames > 65535) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_frames: bad argument");
  if (ctx->d_gid || ctx->n_query != ctx->n) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_frames: not available for sharded clouds");
  if (frame_offsets[0] != 0 || frame_offsets[n_frames] != ctx->n)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_frames: offsets must start at 0 and end at n");
  for (int64_t f = 0; f < n_frames; ++f)
    if (frame_offsets[f + 1] < frame_offsets[f]) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_frames: offsets must not decrease");
  PCP_CUDA(cudaSetDevice(ctx->device));
  ctx->grid_ok = false;
  ctx->n_frames = n_frames;
  ctx->h_frames.assign(frame_offsets, frame_offsets + n_frames + 1);
  int rc = pcp_reserve(ctx, ctx->frames, (size_t)(n_frames + 1) * sizeof(int64_t));
  if (rc) return rc;
  PCP_CUDA(cudaMemcpyAsync(ctx->frames.p, ctx->h_frames.data(), (size_t)(n_frames + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
  return PCP_OK;
}

PCP_API int pcp_set_cloud_device(pcp_ctx* ctx, const void* d_xyzi, int64_t n) {
  return pcp_set_cloud_device_sharded(ctx, d_xyzi, nullptr, n, n);
}

PCP_API int pcp_build_grid(pcp_ctx* ctx, float cell, int k_hint) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->d_cloud || ctx->n <= 0) return pcp_fail(ctx, PCP_ERR_NO_CLOUD, "pcp_build_grid: no cloud set");
  PCP_CUDA(cudaSetDevice(ctx->device));
  return pcp_grid_build_impl(ctx, cell, k_hint);
}

PCP_API int pcp_grid_info(pcp_ctx* ctx, int32_t dims[3], float* cell, int64_t* n_cells) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (!ctx->grid_ok) return pcp_fail(ctx, PCP_ERR_NO_GRID, "pcp_grid_info: no grid");
  if (dims)
    for (int a = 0; a < 3; ++a) dims[ctx->grid.ax[a]] = ctx->grid.dim[a];
  if (cell) *cell = ctx->grid.cell;
  if (n_cells) *n_cells = ctx->grid.n_cells;
  return PCP_OK;
}

}  // extern "C"
