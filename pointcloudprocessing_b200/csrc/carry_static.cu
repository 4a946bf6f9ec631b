// carry_static.cu — pcp_carry_forward_static: the stale-record carry of the reference loop (`tempFea` lives outside
// the loop: src/pcaAnalysis.cpp:16,26,39,47-51; pca_aux.cu does it inside pcp_radius_pca) as an entry point that
// ALLOCATES NOTHING.  No counterpart in the reference: added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// pca_aux.cu keeps its scratch (last valid index per record + scan storage) in buffers reserved with cudaMalloc on
// first use.  Here the same scratch is a file-scope static device array of a fixed maximum size:
//     PCP_CARRY_STATIC_MAX_RECORDS = 1 Mi records  ->  4 MiB + 4 KiB, held by every GPU that loads libpcp.so until exit.
// Host code never passes the array's NAME to a kernel (in host code that is a host-side stand-in, not a device
// address): the device address is fetched once per device with cudaGetSymbolAddress, with that device current.
// Every call on a GPU shares the array and finds the previous call's contents in it.  Nothing in it has to start at
// zero: each call overwrites every entry it later reads (block_last[0 .. blocks), last_valid[0 .. n)) — but two calls
// that overlap on one GPU (two contexts / streams) RACE on it; callers serialise them.
#include <mutex>

#include "pcp_internal.cuh"

#define PCP_CARRY_STATIC_MAX_RECORDS (1 << 20)

namespace {

constexpr int CS_BLOCK = 1024;                                         // records per block of the two scan kernels
constexpr int CS_MAX_BLOCKS = PCP_CARRY_STATIC_MAX_RECORDS / CS_BLOCK;  // 1024
constexpr int CS_WORDS = PCP_CARRY_STATIC_MAX_RECORDS + CS_MAX_BLOCKS;

// [0, MAX_RECORDS): last valid index at or before record i (-1: none);  [MAX_RECORDS, +MAX_BLOCKS): per block, the
// last valid index inside it, then (after cs_scan_blocks_kernel) the last valid index BEFORE the block
static __device__ __align__(256) int g_carry_scratch[CS_WORDS];

__device__ __forceinline__ bool cs_valid(const float* __restrict__ rec, int64_t i, int min_pts) {
  return __float_as_int(rec[i * 17 + 13]) >= min_pts;  // ptNum (word 13 of the 68-byte record) against the gate
}

// block-wide inclusive max-scan of one int per thread (1024 threads)
__device__ __forceinline__ int cs_block_scan_max(int v, int* warp_tot /* 32 shared ints */) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v = max(v, t);
  }
  if (lane == 31) warp_tot[w] = v;
  __syncthreads();
  if (w == 0) {
    int t = warp_tot[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t = max(t, u);
    }
    warp_tot[lane] = t;
  }
  __syncthreads();
  if (w > 0) v = max(v, warp_tot[w - 1]);
  return v;
}

// (1) per block: the last valid record inside it
__global__ void __launch_bounds__(CS_BLOCK) cs_block_last_kernel(const float* __restrict__ rec, int64_t n, int min_pts, int* __restrict__ block_last) {
  __shared__ int warp_tot[32];
  const int64_t i = (int64_t)blockIdx.x * CS_BLOCK + threadIdx.x;
  const int v = (i < n && cs_valid(rec, i, min_pts)) ? (int)i : -1;
  const int s = cs_block_scan_max(v, warp_tot);
  if (threadIdx.x == CS_BLOCK - 1) block_last[blockIdx.x] = s;
}

// (2) one block: exclusive max-scan over the (at most 1024) block entries
__global__ void __launch_bounds__(CS_BLOCK) cs_scan_blocks_kernel(int* __restrict__ block_last, int blocks) {
  __shared__ int warp_tot[32];
  const int v = (int)threadIdx.x < blocks ? block_last[threadIdx.x] : -1;
  const int s = cs_block_scan_max(v, warp_tot);
  __shared__ int incl[CS_BLOCK];
  incl[threadIdx.x] = s;
  __syncthreads();
  if ((int)threadIdx.x < blocks) block_last[threadIdx.x] = threadIdx.x == 0 ? -1 : incl[threadIdx.x - 1];
}

// (3) last valid index at or before every record
__global__ void __launch_bounds__(CS_BLOCK) cs_last_valid_kernel(const float* __restrict__ rec, int64_t n, int min_pts,
                                                                  const int* __restrict__ block_before, int* __restrict__ last_valid) {
  __shared__ int warp_tot[32];
  const int64_t i = (int64_t)blockIdx.x * CS_BLOCK + threadIdx.x;
  const int v = (i < n && cs_valid(rec, i, min_pts)) ? (int)i : -1;
  const int s = max(cs_block_scan_max(v, warp_tot), block_before[blockIdx.x]);
  if (i < n) last_valid[i] = s;
}

// (4) one warp per record, lanes 0..16 one word each: a valid record is copied; an invalid one keeps its ptID / ptNum
// (words 12, 13) and takes everything else from the last valid record before it (its own words when there is none:
// the leading records of a cloud).  Sources are valid records, which no warp rewrites: in == out is safe.
__global__ void __launch_bounds__(256) cs_gather_kernel(const float* in, float* out, int64_t n,  // in may equal out
                                                       
                                                        const int* __restrict__ last_valid) {
  const int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (i >= n || lane >= 17) return;
  const int src = last_valid[i];
  const bool own = src < 0 || src == (int)i || lane == 12 || lane == 13;
  const float w = in[(own ? i : (int64_t)src) * 17 + lane];
  if (in != out || !own) out[i * 17 + lane] = w;
}

// device address of the scratch array, fetched once per device with that device current
constexpr int CS_MAX_DEVICES = 64;
int* g_scratch_ptr[CS_MAX_DEVICES] = {nullptr};
std::mutex g_scratch_mutex;

int cs_scratch(pcp_ctx* ctx, int** out) {
  const int dev = ctx->device;
  if (dev < 0 || dev >= CS_MAX_DEVICES) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_carry_forward_static: device index beyond 63");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[dev]) {
    int* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_carry_scratch);  // ctx->device is current (set by the caller below)
    if (e != cudaSuccess || !p) return pcp_fail(ctx, PCP_ERR_CUDA, "pcp_carry_forward_static: cudaGetSymbolAddress", e);
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return PCP_OK;
}

}  // namespace

extern "C" PCP_API int64_t pcp_carry_forward_static_max(void) { return (int64_t)PCP_CARRY_STATIC_MAX_RECORDS; }

extern "C" PCP_API int pcp_carry_forward_static(pcp_ctx* ctx, const pcp_feature68* d_in, pcp_feature68* d_out, int64_t n, int min_pts) {
  if (!ctx) return PCP_ERR_INVALID_ARG;
  if (n < 0 || (n > 0 && (!d_in || !d_out))) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_carry_forward_static: bad argument");
  if (n > (int64_t)PCP_CARRY_STATIC_MAX_RECORDS)
    return pcp_fail(ctx, PCP_ERR_INVALID_ARG,
                    "pcp_carry_forward_static: more than 1048576 records (PCP_CARRY_STATIC_MAX_RECORDS, the size of the static scratch array); "
                    "split the call or use pcp_radius_pca(carry_forward = 1)");
  if (n == 0) return PCP_OK;
  PCP_CUDA(cudaSetDevice(ctx->device));
  int* scratch = nullptr;
  int rc = cs_scratch(ctx, &scratch);
  if (rc) return rc;
  int* last_valid = scratch;
  int* block_last = scratch + PCP_CARRY_STATIC_MAX_RECORDS;
  const int blocks = (int)((n + CS_BLOCK - 1) / CS_BLOCK);
  const float* in = reinterpret_cast<const float*>(d_in);
  float* out = reinterpret_cast<float*>(d_out);
  cs_block_last_kernel<<<blocks, CS_BLOCK, 0, ctx->stream>>>(in, n, min_pts, block_last);
  PCP_LAUNCH_CHECK("cs_block_last_kernel");
  cs_scan_blocks_kernel<<<1, CS_BLOCK, 0, ctx->stream>>>(block_last, blocks);
  PCP_LAUNCH_CHECK("cs_scan_blocks_kernel");
  cs_last_valid_kernel<<<blocks, CS_BLOCK, 0, ctx->stream>>>(in, n, min_pts, block_last, last_valid);
  PCP_LAUNCH_CHECK("cs_last_valid_kernel");
  cs_gather_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, ctx->stream>>>(in, out, n, last_valid);
  PCP_LAUNCH_CHECK("cs_gather_kernel");
  return PCP_OK;
}
