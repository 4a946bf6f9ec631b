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

// RingProjector::groundRemoval + cloudSegmentation + labelComponents restated literally
// (src/ringProjection.cpp:83-136, 138-197, 199-289; constants include/ringProjection.h:21-28), for one
// frame whose projection (range image + winning point index per pixel) is given.  Quirks kept:
//  * `neighborIterator` is declared vector<pair<uint8_t,uint8_t>> (include/ringProjection.h:69) but filled
//    with -1 entries (:119-122): (-1,0) becomes (255,0), always off the image; (0,-1) becomes (0,255), i.e.
//    the pixel 255 columns to the RIGHT, or column 0 when that passes the last column (`thisIndY >=
//    Horizon_SCAN -> 0`, :226-227; the `< 0` branch is dead).  The flood therefore grows right, far-right
//    and down only, and its result depends on the raster order of the seeds.
//  * lineCountFlag is set for pushed pixels only (:254), never for the seed's own row.
//  * groundMat(i,j) = -1 of pair i can overwrite the 1 that pair i-1 gave the same pixel (:94-98).
//  * maxGroundrangeOfcol is `new double[]`, never initialised (include/ringProjection.h:130); a fresh
//    RingProjector is made per frame (src/dynaObjTest.cpp:103): defined here as zeros.
//  * float overloads everywhere (`using namespace std` + <math.h>, include/commontools.h:10,44):
//    atan2f, sqrtf, sinf, cosf, fabsf; `* 180` in float, `/ M_PI` in double, narrowed on assignment.
extern "C" int32_t orc_ring_segment(const float* xyzi, int n_scan, int horizon, float ang_res_x, float ang_res_y, int ground_scan_ind,
                                    float sensor_mount_angle, float segment_theta, int valid_point_num, int valid_line_num,
                                    const float* range_img, const int32_t* pt_index, int8_t* ground, int32_t* label,
                                    float* col_max_ground, int32_t* above_idx, int32_t* n_labels) {
  const int64_t cells = (int64_t)n_scan * horizon;
  for (int64_t c = 0; c < cells; ++c) {
    ground[c] = 0;
    label[c] = 0;
  }
  double* max_ground = new double[horizon];
  for (int j = 0; j < horizon; ++j) max_ground[j] = 0.0;
  // ---- groundRemoval, first loop (:88-114)
  for (int j = 0; j < horizon; ++j)
    for (int i = 0; i < ground_scan_ind; ++i) {
      const int64_t lower = j + (int64_t)i * horizon, upper = j + (int64_t)(i + 1) * horizon;
      if (pt_index[lower] < 0 || pt_index[upper] < 0) {  // intensity == -1: the nanPoint
        ground[lower] = -1;
        continue;
      }
      const float* pu = xyzi + 4 * (int64_t)pt_index[upper];
      const float* pl = xyzi + 4 * (int64_t)pt_index[lower];
      const float dx = pu[0] - pl[0], dy = pu[1] - pl[1], dz = pu[2] - pl[2];
      const float angle = (float)((double)(atan2f(dz, sqrtf(dx * dx + dy * dy)) * 180.0f) / M_PI);
      if (fabsf(angle - sensor_mount_angle) <= 10.0f) {
        ground[lower] = 1;
        ground[upper] = 1;
      }
    }
  // ---- second loop (:118-128)
  for (int i = 0; i < n_scan; ++i)
    for (int j = 0; j < horizon; ++j) {
      const int64_t c = j + (int64_t)i * horizon;
      if (ground[c] == 1) {
        label[c] = -1;
        if ((double)range_img[c] > max_ground[j]) max_ground[j] = (double)range_img[c];
      }
      if (range_img[c] == FLT_MAX) label[c] = -1;
    }
  // ---- cloudSegmentation: flood from every unlabelled pixel in raster order (:140-143, 199-289)
  const float alpha_x = (float)((double)ang_res_x / 180.0 * M_PI), alpha_y = (float)((double)ang_res_y / 180.0 * M_PI);
  const int nb_row[4] = {255, 0, 0, 1}, nb_col[4] = {0, 1, 255, 0};
  int32_t* qx = new int32_t[cells];
  int32_t* qy = new int32_t[cells];
  bool* line_flag = new bool[n_scan];
  int label_count = 1;
  for (int row = 0; row < n_scan; ++row)
    for (int col = 0; col < horizon; ++col) {
      if (label[col + (int64_t)row * horizon] != 0) continue;
      for (int i = 0; i < n_scan; ++i) line_flag[i] = false;
      qx[0] = row;
      qy[0] = col;
      int64_t q_size = 1, q_start = 0, q_end = 1;  // the queue array doubles as allPushedInd
      while (q_size > 0) {
        const int fx = qx[q_start], fy = qy[q_start];
        --q_size;
        ++q_start;
        label[fy + (int64_t)fx * horizon] = label_count;
        for (int t = 0; t < 4; ++t) {
          int tx = fx + nb_row[t], ty = fy + nb_col[t];
          if (tx < 0 || tx >= n_scan) continue;
          if (ty < 0) ty = horizon - 1;
          if (ty >= horizon) ty = 0;
          if (label[ty + (int64_t)tx * horizon] != 0) continue;
          const float ra = range_img[fy + (int64_t)fx * horizon], rb = range_img[ty + (int64_t)tx * horizon];
          const float d1 = ra > rb ? ra : rb, d2 = ra < rb ? ra : rb;  // std::max / std::min
          const float alpha = nb_row[t] == 0 ? alpha_x : alpha_y;
          const float angle = atan2f(d2 * sinf(alpha), d1 - d2 * cosf(alpha));
          if (angle > segment_theta) {
            qx[q_end] = tx;
            qy[q_end] = ty;
            ++q_size;
            ++q_end;
            label[ty + (int64_t)tx * horizon] = label_count;
            line_flag[tx] = true;
          }
        }
      }
      bool feasible = false;
      if (q_end >= 30) feasible = true;
      else if (q_end >= valid_point_num) {
        int lines = 0;
        for (int i = 0; i < n_scan; ++i) lines += line_flag[i] ? 1 : 0;
        feasible = lines >= valid_line_num;
      }
      if (feasible) ++label_count;
      else
        for (int64_t i = 0; i < q_end; ++i) label[qy[i] + (int64_t)qx[i] * horizon] = 999999;
    }
  // ---- cloudaboveGround (:186-189): raster order, original point indices
  int32_t n_above = 0;
  for (int i = 0; i < n_scan; ++i)
    for (int j = 0; j < horizon; ++j) {
      const int64_t c = j + (int64_t)i * horizon;
      if (ground[c] != 1 && (double)range_img[c] < max_ground[j]) above_idx[n_above++] = pt_index[c];
    }
  for (int64_t c = n_above; c < cells; ++c) above_idx[c] = -1;
  for (int j = 0; j < horizon; ++j) col_max_ground[j] = (float)max_ground[j];
  *n_labels = label_count - 1;
  delete[] max_ground;
  delete[] qx;
  delete[] qy;
  delete[] line_flag;
  return n_above;
}
