// oracle/ring.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// RingProjector::projectPointCloud restated (src/ringProjection.cpp:24-81; constants
// include/ringProjection.h:16-20).  Mixed precision follows the C++ promotions of the
// reference expressions: atan2/sqrt on floats pick the float overloads, `* 180` stays
// float, `/ M_PI` and `- 90.0` promote to double, assignments narrow back to float.
// Row/column are size_t in the reference, so the `< 0` tests are dead code: a negative
// float truncates toward zero (−0.3 → 0) or wraps to a huge value that the `>=` tests
// reject; restated with int64 truncation + explicit range checks.
#include "pcp_oracle.h"

#include <cfloat>
#include <cmath>

extern "C" int32_t orc_ring_project(const float* xyzi, int64_t n, int n_scan, int horizon, float ang_res_x,
                                    float ang_res_y, float ang_bottom, float* range_img, int32_t* pt_index,
                                    float* full_intensity) {
  const int64_t cells = (int64_t)n_scan * horizon;
  for (int64_t c = 0; c < cells; ++c) {
    if (range_img) range_img[c] = FLT_MAX;  // rangeMat init, include/ringProjection.h:142
    if (pt_index) pt_index[c] = -1;
    if (full_intensity) full_intensity[c] = -1.f;
  }
  int32_t long_range = 0;
  for (int64_t i = 0; i < n; ++i) {
    float x = xyzi[4 * i], y = xyzi[4 * i + 1], z = xyzi[4 * i + 2];
    float vertical = (float)((double)(atan2f(z, sqrtf(x * x + y * y)) * 180.0f) / M_PI);
    float rowf = (vertical + ang_bottom) / ang_res_y;
    if (!(rowf > -9.0e18f && rowf < 9.0e18f)) continue;  // NaN / out of int64: rejected upstream by the >= test
    int64_t row = (int64_t)rowf;
    if (row < 0 || row >= n_scan) continue;

    float horizonAngle = (float)((double)(atan2f(x, y) * 180.0f) / M_PI);
    double cold = -round(((double)horizonAngle - 90.0) / (double)ang_res_x) + (double)(horizon / 2);
    if (!(cold > -9.0e18 && cold < 9.0e18)) continue;
    int64_t col = (int64_t)cold;
    if (col >= horizon) col -= horizon;
    if (col < 0 || col >= horizon) continue;

    float range = sqrtf(x * x + y * y + z * z);
    int64_t index = col + row * horizon;
    if (range_img) range_img[index] = range;  // later points overwrite earlier ones
    if (pt_index) pt_index[index] = (int32_t)i;
    if (full_intensity) full_intensity[index] = (float)((double)(float)row + (double)(float)col / 10000.0);
    if (range > 60.f) ++long_range;
  }
  return long_range;
}
