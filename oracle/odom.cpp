// oracle/odom.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// lidarOdometryClass::getRangeAndptIdImage (src/lidarOdometry.cpp:21-113) and
// lidarOdometryClass::calculateSmoothness (src/lidarOdometry.cpp:121-187) restated for one frame.
// Mixed precision follows the C++ promotions of the reference expressions: the file includes
// "math.h" in C++, so sqrt(float) / atan(float) pick the float overloads; `* 180.0 / M_PI`
// promotes to double, left to right.  Images are returned row-major rows x cols (the reference's
// Eigen::MatrixXf is column-major; the layout of the container is not part of the result).
//
// Behaviour the reference leaves undefined and what is done here (the CUDA path does the same):
//  * row >= 16 or col >= 900 (vertical span > 32 deg; horizontal span of exactly 180 deg) writes
//    outside the 16 x 900 matrix upstream (:96-100) -> the point is skipped;
//  * NaN angles (x = y = 0) make int(NaN) upstream -> the point is skipped, and NaN never
//    updates the min/max (every comparison with it is false, as upstream);
//  * the smoothness window of the last columns reads rangeImg(row, 900..907) (:141-143), past the
//    matrix -> read as 0 (not valid).
#include "pcp_oracle.h"

#include <algorithm>
#include <cmath>
#include <vector>

namespace {
// The reference's libm (glibc 2.23 on its Ubuntu 16.04) and this image's (2.39) both evaluate atanf with the
// fdlibm algorithm restated below; later glibc versions round correctly instead, so the oracle calls the
// restatement and orc_atanf_mismatches() proves it equal to the local libm where that still holds.
float atanf_fdlibm(float x);
inline double vert_angle(float x, float y, float z) {
  // atan(z / sqrt(x*x + y*y)) * 180.0 / M_PI   (:46-47, :80-81)
  return (double)atanf_fdlibm(z / sqrtf(x * x + y * y)) * 180.0 / M_PI;
}
inline double hori_angle(float x, float y) {
  // atan(y / x) * 180.0 / M_PI                  (:49, :83)
  return (double)atanf_fdlibm(y / x) * 180.0 / M_PI;
}
}  // namespace

extern "C" void orc_odom_range_image(const float* xyzi, int64_t n, const int32_t* nb_counts, int min_neighbors, int n_rows,
                                     int n_cols, double vert_res, double hori_res, float* range_img, float* ptid_img,
                                     int32_t* pt_index, int32_t* stats2, double* minmax4) {
  const int64_t cells = (int64_t)n_rows * n_cols;
  for (int64_t c = 0; c < cells; ++c) {
    range_img[c] = 0.f;  // MatrixXf::Zero (:31-32)
    ptid_img[c] = 0.f;
    if (pt_index) pt_index[c] = -1;
  }
  double vmax = 0, hmax = 0, vmin = 1000, hmin = 1000;  // (:40-41)
  for (int64_t i = 0; i < n; ++i) {
    const float x = xyzi[4 * i], y = xyzi[4 * i + 1], z = xyzi[4 * i + 2];
    const double v = vert_angle(x, y, z), h = hori_angle(x, y);
    if (v > vmax) vmax = v;
    if (v < vmin) vmin = v;
    if (h > hmax) hmax = h;
    if (h < hmin) hmin = h;
  }
  if (minmax4) {
    minmax4[0] = vmin;
    minmax4[1] = vmax;
    minmax4[2] = hmin;
    minmax4[3] = hmax;
  }
  int32_t lost = 0, in_img = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (nb_counts[i] < min_neighbors) continue;  // radiusSearch(i, 3.0).size() < 6 (:76-78)
    const float x = xyzi[4 * i], y = xyzi[4 * i + 1], z = xyzi[4 * i + 2];
    const double v = vert_angle(x, y, z) - vmin, h = hori_angle(x, y) - hmin;
    const double rd = v / vert_res, cd = h / hori_res;  // (:85-86)
    if (!(rd > -1.0 && rd < 2.0e9) || !(cd > -1.0 && cd < 2.0e9)) continue;  // NaN / not an int
    const int row = (int)rd, col = (int)cd;
    if (row < 0 || row >= n_rows || col < 0 || col >= n_cols) continue;  // outside the matrix: UB upstream
    const int64_t c = (int64_t)row * n_cols + col;
    if (range_img[c] != 0.f) ++lost;  // "losing points num" (:95-96)
    range_img[c] = sqrtf(x * x + y * y + z * z);  // (:97-99)
    ptid_img[c] = (float)(int)i;                  // (:100)
    if (pt_index) pt_index[c] = (int32_t)i;
  }
  for (int64_t c = 0; c < cells; ++c)
    if (range_img[c] > 0.f) ++in_img;  // "points num in range image" (:106-110)
  if (stats2) {
    stats2[0] = lost;
    stats2[1] = in_img;
  }
}

extern "C" void orc_odom_smoothness(const float* range_img, int n_rows, int n_cols, int n_blocks, int half_window, float near_range,
                                    int min_valid, int pick, int min_block_valid, float* smooth_img, float* type_img) {
  const int64_t cells = (int64_t)n_rows * n_cols;
  for (int64_t c = 0; c < cells; ++c) smooth_img[c] = type_img[c] = 0.f;  // (:123-124)
  const int bw = n_cols / n_blocks;                                         // 900 / 5 = 180 (:131)
  auto R = [&](int row, int col) -> float { return col < n_cols ? range_img[(int64_t)row * n_cols + col] : 0.f; };
  std::vector<double> srow;
  for (int row = 0; row < n_rows; ++row) {
    for (int b = 0; b < n_blocks; ++b) {
      srow.clear();
      for (int col = bw * b + half_window; col < bw * (b + 1); ++col) {
        if (R(row, col) < near_range) continue;  // (:136-137)
        int valid = 0;
        double sum = 0;
        for (int i = col - half_window; i < col + half_window; ++i) {  // (:141)
          if (R(row, i) > near_range) {
            ++valid;
            sum += R(row, i);
          }
        }
        if (valid >= min_valid) {
          double s = sum - R(row, col) * valid;        // float * int -> float, then double (:151)
          s = fabs(s / (valid * R(row, col)));         // (:152)
          smooth_img[(int64_t)row * n_cols + col] = (float)s;
          srow.push_back(s);
        }
      }
      const size_t m = srow.size();
      std::sort(srow.begin(), srow.end());
      if (m > (size_t)min_block_valid) {  // (:162)
        for (int col = bw * b + half_window; col < bw * (b + 1); ++col) {
          const double s = smooth_img[(int64_t)row * n_cols + col];
          if (s >= srow[m - pick]) type_img[(int64_t)row * n_cols + col] = 1.f;  // (:167-169)
          if (s < srow[pick]) type_img[(int64_t)row * n_cols + col] = 2.f;      // (:170-172)
        }
      }
    }
  }
}

// ---- fdlibm atanf restated (the algorithm of glibc's flt-32/s_atanf.c before 2.41), the form the CUDA
// path evaluates.  orc_atanf_mismatches walks every `stride`-th float bit pattern and counts the inputs
// where it differs from the C library's atanf (0 expected on this image: glibc 2.39; stride 1 = all 2^32
// inputs, about half a minute).
namespace {
float atanf_fdlibm(float x) {
  static const float hi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
  static const float lo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
  int32_t hx;
  __builtin_memcpy(&hx, &x, 4);
  const int32_t ix = hx & 0x7fffffff;
  if (ix >= 0x4c000000) {
    if (ix > 0x7f800000) return x + x;
    return hx > 0 ? hi[3] + lo[3] : -hi[3] - lo[3];
  }
  int id;
  if (ix < 0x3ee00000) {
    if (ix < 0x31000000) return x;
    id = -1;
  } else {
    x = fabsf(x);
    if (ix < 0x3f980000) {
      if (ix < 0x3f300000) {
        id = 0;
        x = (2.0f * x - 1.0f) / (2.0f + x);
      } else {
        id = 1;
        x = (x - 1.0f) / (x + 1.0f);
      }
    } else if (ix < 0x401c0000) {
      id = 2;
      x = (x - 1.5f) / (1.0f + 1.5f * x);
    } else {
      id = 3;
      x = -1.0f / x;
    }
  }
  const float z = x * x, w = z * z;
  const float s1 = z * (3.3333334327e-01f + w * (1.4285714924e-01f + w * (9.0908870101e-02f + w * (6.6610731184e-02f + w * (4.9768779427e-02f + w * 1.6285819933e-02f)))));
  const float s2 = w * (-2.0000000298e-01f + w * (-1.1111110449e-01f + w * (-7.6918758452e-02f + w * (-5.8335702866e-02f + w * -3.6531571299e-02f))));
  if (id < 0) return x - x * (s1 + s2);
  const float zz = hi[id] - ((x * (s1 + s2) - lo[id]) - x);
  return hx < 0 ? -zz : zz;
}
}  // namespace

extern "C" int64_t orc_atanf_mismatches(uint32_t start, uint32_t stride) {
  int64_t bad = 0;
  for (uint64_t u = start; u < 0x100000000ull; u += stride) {
    const uint32_t v = (uint32_t)u;
    float x;
    __builtin_memcpy(&x, &v, 4);
    const float a = atanf(x), b = atanf_fdlibm(x);
    uint32_t ua, ub;
    __builtin_memcpy(&ua, &a, 4);
    __builtin_memcpy(&ub, &b, 4);
    if (ua != ub && !(a != a && b != b)) ++bad;
  }
  return bad;
}
