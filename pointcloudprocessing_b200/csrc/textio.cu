// textio.cu — "x y z intensity" text clouds parsed on the device (second half).
//
// Replaces preprocess::txt2pc (src/smoothing.cpp:39-81): getline per line, four `ss >> float` extractions
// into ONE point object that lives across lines, push_back per line.  What a caller can observe of that:
//   * every line yields a point (empty lines too; a last line without '\n' counts, a trailing '\n' adds none);
//   * a field the line does not supply keeps the PREVIOUS line's value (the extraction fails before it
//     writes); a field that does not parse becomes 0 (C++11 num_get) and ends the line; a value beyond
//     the float range becomes +-FLT_MAX and ends the line;
//   * the characters a field takes follow libstdc++'s num_get (_M_extract_float): sign, digits, one
//     decimal point, 'e'/'E' after at least one digit with an optional sign; the collected string must
//     be a complete number for strtof ("1e", "-", "." fail; no "inf"/"nan"/hex);
//   * the value is the correctly rounded float of the decimal (strtof).
// Device: newline ranks by a CUB scan over the bytes -> line starts; one thread per line runs the
// extraction state machine; decimal -> float is exact integer arithmetic (M * 5^E or M / 5^e in 256-bit
// integers, round to nearest even with a sticky bit), so results are bit-identical to strtof for up to 19
// significant digits (longer mantissas round through their 19-digit prefix + sticky); the carry-over of
// unset fields is a max-scan of "last line that set this field" per field.
#include <cfloat>
#include <cstdint>

#include <cub/device/device_scan.cuh>
#include <thrust/iterator/transform_iterator.h>

#include "pcp_internal.cuh"

namespace {

struct IsNewline {
  __host__ __device__ __forceinline__ int operator()(const unsigned char& c) const { return c == '\n' ? 1 : 0; }
};
struct MaxInt {
  __host__ __device__ __forceinline__ int operator()(int a, int b) const { return a > b ? a : b; }
};

// ---- 256-bit unsigned integers, little-endian 32-bit words ----
struct U256 {
  uint32_t w[8];
};
__device__ __forceinline__ void u256_set64(U256& a, unsigned long long v) {
  a.w[0] = (uint32_t)v;
  a.w[1] = (uint32_t)(v >> 32);
#pragma unroll
  for (int i = 2; i < 8; ++i) a.w[i] = 0;
}
__device__ __forceinline__ void u256_mul_small(U256& a, uint32_t m) {
  unsigned long long c = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    c += (unsigned long long)a.w[i] * m;
    a.w[i] = (uint32_t)c;
    c >>= 32;
  }
}
__device__ __forceinline__ int u256_bitlen(const U256& a) {
  for (int i = 7; i >= 0; --i)
    if (a.w[i]) return 32 * i + 32 - __clz(a.w[i]);
  return 0;
}
__device__ __forceinline__ void u256_shl(U256& a, int s) {  // 0 <= s < 256
  const int ws = s >> 5, bs = s & 31;
  for (int i = 7; i >= 0; --i) {
    uint32_t v = 0;
    if (i - ws >= 0) {
      v = a.w[i - ws] << bs;
      if (bs && i - ws - 1 >= 0) v |= a.w[i - ws - 1] >> (32 - bs);
    }
    a.w[i] = v;
  }
}
__device__ __forceinline__ bool u256_ge(const U256& a, const U256& b) {
  for (int i = 7; i >= 0; --i)
    if (a.w[i] != b.w[i]) return a.w[i] > b.w[i];
  return true;
}
__device__ __forceinline__ void u256_sub(U256& a, const U256& b) {
  long long br = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    long long d = (long long)a.w[i] - (long long)b.w[i] - br;
    br = d < 0 ? 1 : 0;
    a.w[i] = (uint32_t)d;
  }
}
__device__ __forceinline__ void u256_shr1(U256& a) {
#pragma unroll
  for (int i = 0; i < 8; ++i) a.w[i] = (a.w[i] >> 1) | (i < 7 ? a.w[i + 1] << 31 : 0u);
}
__device__ __forceinline__ bool u256_bit(const U256& a, int i) { return i >= 0 && i < 256 && ((a.w[i >> 5] >> (i & 31)) & 1u); }
__device__ __forceinline__ bool u256_any_below(const U256& a, int i) {  // any set bit at positions < i
  if (i <= 0) return false;
  const int ws = i >> 5, bs = i & 31;
  for (int k = 0; k < ws && k < 8; ++k)
    if (a.w[k]) return true;
  return ws < 8 && bs && (a.w[ws] & ((1u << bs) - 1u));
}
__device__ __forceinline__ uint32_t u256_extract(const U256& a, int lo, int n) {  // bits [lo, lo+n), n <= 24, lo >= 0
  uint32_t v = 0;
  for (int k = 0; k < n; ++k) v |= (u256_bit(a, lo + k) ? 1u : 0u) << k;
  return v;
}

// round_to_nearest_even((Q + sticky*eps) * 2^x) as float bits (no sign); *overflow when beyond FLT_MAX
__device__ uint32_t round_to_float(const U256& Q, bool sticky, int x, bool* overflow) {
  *overflow = false;
  const int L = u256_bitlen(Q);
  if (L == 0) return 0u;
  const int ex = L - 1 + x;  // value in [2^ex, 2^(ex+1))
  if (ex > 127) {
    *overflow = true;
    return 0x7F800000u;
  }
  int keep = 24;
  if (ex < -126) keep = 24 - (-126 - ex);  // denormal: fewer mantissa bits
  const int shift = L - keep;             // low bits to drop
  uint32_t mant;
  bool half = false, rest = sticky;
  if (shift <= 0) {
    mant = u256_extract(Q, 0, L) << (-shift);
  } else if (keep <= 0) {
    mant = 0;
    if (shift == L) {  // the value lies in [min_denormal / 2, min_denormal)
      half = true;
      rest = rest || u256_any_below(Q, L - 1);
    } else {
      rest = true;
    }
  } else {
    mant = u256_extract(Q, shift, keep);
    half = u256_bit(Q, shift - 1);
    rest = rest || u256_any_below(Q, shift - 1);
  }
  uint32_t bits;
  if (ex < -126) bits = mant;                                  // exponent field 0
  else bits = ((uint32_t)(ex + 127) << 23) + (mant - 0x800000u);  // mant has its implicit bit set
  if (half && (rest || (mant & 1u))) ++bits;                     // a carry walks into the exponent as it should
  if (bits >= 0x7F800000u) {
    *overflow = true;
    return 0x7F800000u;
  }
  return bits;
}

// correctly rounded float of (M + sticky*eps) * 10^E, M != 0
__device__ uint32_t decimal_to_float(unsigned long long M, bool sticky, int E, bool* overflow) {
  *overflow = false;
  if (E > 60) {
    *overflow = true;
    return 0x7F800000u;
  }
  if (E < -70) return 0u;  // M < 10^19: below half the smallest denormal (7e-46)
  U256 Q;
  if (E >= 0) {
    u256_set64(Q, M);
    int e = E;
    while (e >= 13) {
      u256_mul_small(Q, 1220703125u);  // 5^13
      e -= 13;
    }
    while (e-- > 0) u256_mul_small(Q, 5u);
    return round_to_float(Q, sticky, E, overflow);  // M * 5^E * 2^E
  }
  // M / 5^e * 2^-e: 28+ quotient bits by shift-and-subtract, remainder -> sticky
  U256 D;
  u256_set64(D, 1ull);
  int e = -E;
  while (e >= 13) {
    u256_mul_small(D, 1220703125u);
    e -= 13;
  }
  while (e-- > 0) u256_mul_small(D, 5u);
  U256 N;
  u256_set64(N, M);
  const int lN = u256_bitlen(N), lD = u256_bitlen(D);
  int s = lD - lN + 28;  // N * 2^s / D has 28 or 29 bits
  if (s < 0) s = 0;
  u256_shl(N, s);
  // align D under N's top: q bit positions from (lN + s - lD) down to 0
  int top = lN + s - lD;  // >= 28 (or lN - lD when s = 0)
  U256 Dsh = D;
  u256_shl(Dsh, top);
  unsigned long long q = 0;
  for (int b = top; b >= 0; --b) {
    q <<= 1;
    if (u256_ge(N, Dsh)) {
      u256_sub(N, Dsh);
      q |= 1ull;
    }
    u256_shr1(Dsh);
  }
  const bool rem = u256_bitlen(N) != 0;
  U256 Qq;
  u256_set64(Qq, q);
  return round_to_float(Qq, sticky || rem, E - s, overflow);  // q * 2^-s * 2^-e
}

__device__ __forceinline__ bool is_space(unsigned char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }
__device__ __forceinline__ bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }

// line j = bytes [starts[j], end_j): one thread parses its four fields.  set_mask bit f: the field wrote a value.
// XYZ = false: txt2pc (stream extraction, above).  XYZ = true: xyz2pc (src/smoothing.cpp:83-125) — getline(line, 256) +
// sscanf(line, "%f %f %f %s", &x, &y, &z, &intens); pt.intensity = intens: three floats and the FIRST BYTE of the
// fourth token as a signed char; a field that does not parse ends the line WITHOUT being written (it keeps the
// previous line's value, like every field after it), a value beyond the float range is stored as +-inf and the scan
// goes on; a line of 256 characters or more raises *too_long (the reference's loop never ends on it).  Fields in
// inf / nan / hexadecimal notation, which sscanf would accept, are outside the subset offered here (they end the line).
template <bool XYZ>
__global__ void __launch_bounds__(128) txt_parse_kernel(const unsigned char* __restrict__ text, int64_t n_bytes,
                                                        const int64_t* __restrict__ starts, int64_t n_lines, float4* __restrict__ vals,
                                                        int* __restrict__ last_set /* 4 arrays of n_lines */, int* __restrict__ too_long) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n_lines) return;
  int64_t p = starts[j];
  int64_t end = j + 1 < n_lines ? starts[j + 1] - 1 : n_bytes;  // drop the '\n'
  if (j + 1 == n_lines && end > p && text[end - 1] == '\n') --end;
  if (XYZ && end - p >= 256) atomicExch(too_long, 1);
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  int set_mask = 0;
  bool good = true;  // the stringstream's good(): no failbit, no eofbit
  for (int f = 0; f < 4 && good; ++f) {
    while (p < end && is_space(text[p])) ++p;  // sentry: skip whitespace
    if (p >= end) break;                       // eof before any character: failbit, the value is NOT written
    if (XYZ && f == 3) {                       // %s: the token's first byte, as the signed char the reference declares
      v[3] = (float)(signed char)text[p];
      set_mask |= 8;
      break;
    }
    // ---- _M_extract_float: collect the characters of the field
    bool neg = false, found_mantissa = false, found_dec = false, found_sci = false, any_digit = false;
    bool exp_neg = false, exp_digits = false, sticky = false, bad = false;
    unsigned long long M = 0;
    int n_sig = 0, frac_digits = 0, dropped = 0, exp10 = 0;
    unsigned char c = text[p];
    if (c == '+' || c == '-') {
      neg = c == '-';
      ++p;
    }
    while (p < end) {
      c = text[p];
      if (is_digit(c)) {
        if (found_sci) {
          exp_digits = true;
          if (exp10 < 100000) exp10 = exp10 * 10 + (c - '0');
        } else {
          found_mantissa = true;
          any_digit = true;
          if (M == 0 && c == '0') {
            if (found_dec) ++frac_digits;  // leading zeros of a fraction shift the exponent only
          } else if (n_sig < 19) {
            M = M * 10 + (unsigned long long)(c - '0');
            ++n_sig;
            if (found_dec) ++frac_digits;
          } else {
            if (c != '0') sticky = true;
            if (!found_dec) ++dropped;
          }
        }
        ++p;
      } else if (c == '.' && !found_dec && !found_sci) {
        found_dec = true;
        ++p;
      } else if ((c == 'e' || c == 'E') && !found_sci && found_mantissa) {
        found_sci = true;
        ++p;
        if (p < end && (text[p] == '+' || text[p] == '-')) {
          exp_neg = text[p] == '-';
          ++p;
        }
      } else {
        break;
      }
    }
    if (!XYZ && p >= end) good = false;  // the extraction ran into the end of the line: eofbit
    // ---- strtof on the collected string: it must be a complete number
    // (glibc's sscanf converts "1e" / "1e+" as 1 and consumes the exponent marker: only txt2pc's strtof-complete rule rejects them)
    if (!any_digit || (!XYZ && found_sci && !exp_digits)) bad = true;
    uint32_t bits = 0;
    if (bad) {
      if (XYZ) break;  // matching failure: sscanf stops, nothing is stored
      v[f] = 0.f;  // num_get stores 0 and sets failbit
      set_mask |= 1 << f;
      good = false;
      continue;
    }
    if (M != 0) {
      long long E = (long long)(exp_neg ? -exp10 : exp10) - frac_digits + dropped;
      E = E > 100000 ? 100000 : (E < -100000 ? -100000 : E);
      bool ovf;
      bits = decimal_to_float(M, sticky, (int)E, &ovf);
      if (ovf) {  // strtof returns HUGE_VALF: the stream stores +-FLT_MAX and sets failbit; sscanf stores the infinity and goes on
        bits = XYZ ? 0x7F800000u : 0x7F7FFFFFu;
        good = XYZ;
      }
    }
    v[f] = __uint_as_float(bits | (neg ? 0x80000000u : 0u));
    set_mask |= 1 << f;
  }
  vals[j] = make_float4(v[0], v[1], v[2], v[3]);
#pragma unroll
  for (int f = 0; f < 4; ++f) last_set[(int64_t)f * n_lines + j] = (set_mask >> f) & 1 ? (int)j : -1;
}

// last_too: the (empty) line after a final '\n' is a line of its own (xyz2pc's eof() loop)
__global__ void __launch_bounds__(256) txt_starts_kernel(const unsigned char* __restrict__ text, const int* __restrict__ rank, int64_t n_bytes,
                                                         int64_t* __restrict__ starts, int last_too) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n_bytes) return;
  if (i == 0) starts[0] = 0;
  if (text[i] == '\n' && (i + 1 < n_bytes || last_too)) starts[rank[i] + 1] = i + 1;
}

// field f of line j = the value of the last line <= j that set it (0 when none: pcl::PointXYZI() starts at 0)
__global__ void __launch_bounds__(256) txt_fill_kernel(const float4* __restrict__ vals, const int* __restrict__ last_set, int64_t n_lines,
                                                       float4* __restrict__ out) {
  const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (j >= n_lines) return;
  float r[4];
#pragma unroll
  for (int f = 0; f < 4; ++f) {
    const int src = last_set[(int64_t)f * n_lines + j];
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (src >= 0) s = vals[src];
    r[f] = f == 0 ? s.x : (f == 1 ? s.y : (f == 2 ? s.z : s.w));
  }
  out[j] = make_float4(r[0], r[1], r[2], r[3]);
}

}  // namespace

// d_text: n_bytes on the device.  Returns the number of lines through *n_lines_out (host) and leaves the points in
// ctx->txt_out (float4 per line).  Synchronises once (the line count sizes everything that follows).
// xyz != 0: the xyz2pc form (one more point after a final newline; *too_long_out set when a line has 255+ characters).
int pcp_parse_txt_impl(pcp_ctx* ctx, const unsigned char* d_text, int64_t n_bytes, int64_t* n_lines_out, int xyz, int* too_long_out) {
  *n_lines_out = 0;
  if (too_long_out) *too_long_out = 0;
  if (n_bytes <= 0) return PCP_OK;
  if (n_bytes > 0x7FFFFFF0ll) return pcp_fail(ctx, PCP_ERR_INVALID_ARG, "pcp_set_cloud_txt: more than 2 GiB of text per call");
  int rc;
  if ((rc = pcp_reserve(ctx, ctx->txt_rank, (size_t)(n_bytes + 1) * sizeof(int)))) return rc;
  int* rank = (int*)ctx->txt_rank.p;
  auto flags = thrust::make_transform_iterator(d_text, IsNewline());
  size_t tmp = 0;
  PCP_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp, flags, rank, (int)n_bytes, ctx->stream));
  if ((rc = pcp_reserve(ctx, ctx->scan_tmp, tmp))) return rc;
  PCP_CUDA(cub::DeviceScan::ExclusiveSum(ctx->scan_tmp.p, tmp, flags, rank, (int)n_bytes, ctx->stream));
  ctx->launches += 2;
  int last_rank = 0;
  unsigned char last_byte = 0;
  PCP_CUDA(cudaMemcpyAsync(&last_rank, rank + n_bytes - 1, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaMemcpyAsync(&last_byte, d_text + n_bytes - 1, 1, cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cudaStreamSynchronize(ctx->stream));
  // getline: every '\n' ends a line; bytes after the last '\n' are one more line
  // rank of the last byte = newlines before it; the last byte closes or is in a line.  xyz2pc's `while (!eof())` loop
  // reads one more (empty) line after a final newline and pushes the previous point again
  const int64_t n_lines = (int64_t)last_rank + 1 + ((xyz && last_byte == '\n') ? 1 : 0);
  if ((rc = pcp_reserve(ctx, ctx->txt_starts, (size_t)(n_lines + 1) * sizeof(int64_t)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->txt_vals, (size_t)n_lines * sizeof(float4)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->txt_last, (size_t)n_lines * 4 * sizeof(int)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->txt_out, (size_t)n_lines * sizeof(float4)))) return rc;
  if ((rc = pcp_reserve(ctx, ctx->flags, sizeof(int)))) return rc;
  PCP_CUDA(cudaMemsetAsync(ctx->flags.p, 0, sizeof(int), ctx->stream));
  int64_t* starts = (int64_t*)ctx->txt_starts.p;
  txt_starts_kernel<<<(unsigned)((n_bytes + 255) / 256), 256, 0, ctx->stream>>>(d_text, rank, n_bytes, starts, xyz ? 1 : 0);
  PCP_LAUNCH_CHECK("txt_starts_kernel");
  int* last = (int*)ctx->txt_last.p;
  if (xyz)
    txt_parse_kernel<true><<<(unsigned)((n_lines + 127) / 128), 128, 0, ctx->stream>>>(d_text, n_bytes, starts, n_lines, (float4*)ctx->txt_vals.p, last,
                                                                                         (int*)ctx->flags.p);
  else
    txt_parse_kernel<false><<<(unsigned)((n_lines + 127) / 128), 128, 0, ctx->stream>>>(d_text, n_bytes, starts, n_lines, (float4*)ctx->txt_vals.p, last,
                                                                                          (int*)ctx->flags.p);
  PCP_LAUNCH_CHECK("txt_parse_kernel");
  if (xyz && too_long_out) PCP_CUDA(cudaMemcpyAsync(too_long_out, ctx->flags.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  PCP_CUDA(cub::DeviceScan::InclusiveScan(nullptr, tmp, last, last, MaxInt(), (int)n_lines, ctx->stream));
  if ((rc = pcp_reserve(ctx, ctx->scan_tmp, tmp))) return rc;
  for (int f = 0; f < 4; ++f) {
    PCP_CUDA(cub::DeviceScan::InclusiveScan(ctx->scan_tmp.p, tmp, last + (int64_t)f * n_lines, last + (int64_t)f * n_lines, MaxInt(), (int)n_lines,
                                            ctx->stream));
    ctx->launches += 1;
  }
  txt_fill_kernel<<<(unsigned)((n_lines + 255) / 256), 256, 0, ctx->stream>>>((const float4*)ctx->txt_vals.p, last, n_lines, (float4*)ctx->txt_out.p);
  PCP_LAUNCH_CHECK("txt_fill_kernel");
  *n_lines_out = n_lines;
  return PCP_OK;
}
