// oracle/eigen3.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// fp32 restatements of the two upstream 3×3 symmetric eigen-solvers the reference path
// reaches; neither library is vendored in the reference tree:
//   * Eigen::SelfAdjointEigenSolver<Matrix3f>::compute — called at
//     src/pcaAnalysis.cpp:57-59, :98-100, :131 (Eigen 3.3 semantics: deflation threshold
//     2*epsilon; Eigen 3.2 used dummy_precision 1e-5 instead — version unpinned upstream).
//   * pcl::eigen33 (PCL 1.8.1 common/impl/eigen.hpp) — inside pcl::NormalEstimation,
//     src/smoothing.cpp:239-246.
// Built with -ffp-contract=off on baseline x86-64 so no FMA is formed, like the
// reference's own build (CMakeLists.txt:4-12 sets no -march).
#include "pcp_oracle.h"

#include <cfloat>
#include <cmath>
#include <utility>

namespace {

struct Givens {
  float c, s;
};

// Eigen::JacobiRotation<float>::makeGivens(p, q) — real branch
inline Givens make_givens(float p, float q) {
  Givens g;
  if (q == 0.f) {
    g.c = p < 0.f ? -1.f : 1.f;
    g.s = 0.f;
  } else if (p == 0.f) {
    g.c = 0.f;
    g.s = q < 0.f ? 1.f : -1.f;
  } else if (std::fabs(p) > std::fabs(q)) {
    float t = q / p;
    float u = std::sqrt(1.f + t * t);
    if (p < 0.f) u = -u;
    g.c = 1.f / u;
    g.s = -t * g.c;
  } else {
    float t = p / q;
    float u = std::sqrt(1.f + t * t);
    if (q < 0.f) u = -u;
    g.s = -1.f / u;
    g.c = -t * g.s;
  }
  return g;
}

// Eigen numext::hypot for positive reals: p*sqrt(1+(q/p)^2)
inline float eig_hypot(float x, float y) {
  float ax = std::fabs(x), ay = std::fabs(y);
  float p = ax > ay ? ax : ay;
  if (p == 0.f) return 0.f;
  float qp = (ax > ay ? ay : ax) / p;
  return p * std::sqrt(1.f + qp * qp);
}

// one implicit symmetric QR step with Wilkinson shift on the block [start, end]
void qr_step(float* diag, float* sub, int start, int end, float* Q /* row-major 3x3 */) {
  float td = (diag[end - 1] - diag[end]) * 0.5f;
  float e = sub[end - 1];
  float mu = diag[end];
  if (td == 0.f) {
    mu -= std::fabs(e);
  } else {
    float e2 = e * e;
    float h = eig_hypot(td, e);
    if (e2 == 0.f) mu -= (e / (td + (td > 0.f ? 1.f : -1.f))) * (e / h);
    else mu -= e2 / (td + (td > 0.f ? h : -h));
  }
  float x = diag[start] - mu;
  float z = sub[start];
  for (int k = start; k < end; ++k) {
    Givens r = make_givens(x, z);
    float sdk = r.s * diag[k] + r.c * sub[k];
    float dkp1 = r.s * sub[k] + r.c * diag[k + 1];
    diag[k] = r.c * (r.c * diag[k] - r.s * sub[k]) - r.s * (r.c * sub[k] - r.s * diag[k + 1]);
    diag[k + 1] = r.s * sdk + r.c * dkp1;
    sub[k] = r.c * sdk - r.s * dkp1;
    if (k > start) sub[k - 1] = r.c * sub[k - 1] - r.s * z;
    x = sub[k];
    if (k < end - 1) {
      z = -r.s * sub[k + 1];
      sub[k + 1] = r.c * sub[k + 1];
    }
    // Q = Q * G on columns k, k+1:  x' = c*x − s*y ; y' = s*x + c*y
    for (int i = 0; i < 3; ++i) {
      float xi = Q[3 * i + k], yi = Q[3 * i + k + 1];
      Q[3 * i + k] = r.c * xi - r.s * yi;
      Q[3 * i + k + 1] = r.s * xi + r.c * yi;
    }
  }
}

}  // namespace

extern "C" int orc_eigen_sym3_qr(const float C[9], float evals[3], float evecs[9]) {
  // lower triangle, scaled to [-1, 1]
  float m00 = C[0], m10 = C[3], m11 = C[4], m20 = C[6], m21 = C[7], m22 = C[8];
  float scale = 0.f;
  const float lower[6] = {m00, m10, m11, m20, m21, m22};
  for (int i = 0; i < 6; ++i) {
    float a = std::fabs(lower[i]);
    if (a > scale) scale = a;  // NaN never wins, like Eigen's maxCoeff on finite data
  }
  if (scale == 0.f) scale = 1.f;
  m00 /= scale; m10 /= scale; m11 /= scale; m20 /= scale; m21 /= scale; m22 /= scale;

  float diag[3], sub[2];
  float Q[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  // 3×3 tridiagonalisation special case
  diag[0] = m00;
  float v1norm2 = m20 * m20;
  if (v1norm2 <= FLT_MIN) {
    diag[1] = m11;
    diag[2] = m22;
    sub[0] = m10;
    sub[1] = m21;
  } else {
    float beta = std::sqrt(m10 * m10 + v1norm2);
    float invBeta = 1.f / beta;
    float m01 = m10 * invBeta;
    float m02 = m20 * invBeta;
    float q = 2.f * m01 * m21 + m02 * (m22 - m11);
    diag[1] = m11 + m02 * q;
    diag[2] = m22 - m02 * q;
    sub[0] = beta;
    sub[1] = m21 - m01 * q;
    Q[4] = m01; Q[5] = m02;
    Q[7] = m02; Q[8] = -m01;
  }

  const int n = 3, max_iter = 30;
  const float tiny = FLT_MIN;
  const float precision = 2.f * FLT_EPSILON;
  int end = n - 1, start = 0, iter = 0;
  while (end > 0) {
    for (int i = start; i < end; ++i) {
      float as = std::fabs(sub[i]);
      if (as <= (std::fabs(diag[i]) + std::fabs(diag[i + 1])) * precision || as <= tiny) sub[i] = 0.f;
    }
    while (end > 0 && sub[end - 1] == 0.f) --end;
    if (end <= 0) break;
    ++iter;
    if (iter > max_iter * n) break;
    start = end - 1;
    while (start > 0 && sub[start - 1] != 0.f) --start;
    qr_step(diag, sub, start, end, Q);
  }
  // ascending selection sort, swapping the columns of Q alongside
  if (iter <= max_iter * n) {
    for (int i = 0; i < n - 1; ++i) {
      int k = 0;
      for (int j = 1; j < n - i; ++j)
        if (diag[i + j] < diag[i + k]) k = j;
      if (k > 0) {
        std::swap(diag[i], diag[i + k]);
        for (int r = 0; r < 3; ++r) std::swap(Q[3 * r + i], Q[3 * r + i + k]);
      }
    }
  }
  for (int i = 0; i < 3; ++i) evals[i] = diag[i] * scale;
  for (int i = 0; i < 9; ++i) evecs[i] = Q[i];
  return iter;
}

namespace {

void pcl_roots2(float b, float c, float roots[3]) {
  roots[0] = 0.f;
  float d = b * b - 4.f * c;
  if (d < 0.f) d = 0.f;
  float sd = std::sqrt(d);
  roots[2] = 0.5f * (b + sd);
  roots[1] = 0.5f * (b - sd);
}

void pcl_roots(const float m[9], float roots[3]) {
  float c0 = m[0] * m[4] * m[8] + 2.f * m[1] * m[2] * m[5] - m[0] * m[5] * m[5] - m[4] * m[2] * m[2] -
             m[8] * m[1] * m[1];
  float c1 = m[0] * m[4] - m[1] * m[1] + m[0] * m[8] - m[2] * m[2] + m[4] * m[8] - m[5] * m[5];
  float c2 = m[0] + m[4] + m[8];
  if (std::fabs(c0) < FLT_EPSILON) {
    pcl_roots2(c2, c1, roots);
    return;
  }
  const float s_inv3 = (float)(1.0 / 3.0);
  const float s_sqrt3 = std::sqrt(3.0f);
  float c2_over_3 = c2 * s_inv3;
  float a_over_3 = (c1 - c2 * c2_over_3) * s_inv3;
  if (a_over_3 > 0.f) a_over_3 = 0.f;
  float half_b = 0.5f * (c0 + c2_over_3 * (2.f * c2_over_3 * c2_over_3 - c1));
  float q = half_b * half_b + a_over_3 * a_over_3 * a_over_3;
  if (q > 0.f) q = 0.f;
  float rho = std::sqrt(-a_over_3);
  float theta = std::atan2(std::sqrt(-q), half_b) * s_inv3;
  float cos_theta = std::cos(theta);
  float sin_theta = std::sin(theta);
  roots[0] = c2_over_3 + 2.f * rho * cos_theta;
  roots[1] = c2_over_3 - rho * (cos_theta + s_sqrt3 * sin_theta);
  roots[2] = c2_over_3 - rho * (cos_theta - s_sqrt3 * sin_theta);
  if (roots[0] >= roots[1]) std::swap(roots[0], roots[1]);
  if (roots[1] >= roots[2]) {
    std::swap(roots[1], roots[2]);
    if (roots[0] >= roots[1]) std::swap(roots[0], roots[1]);
  }
  if (roots[0] <= 0.f) pcl_roots2(c2, c1, roots);
}

inline void cross3(const float* a, const float* b, float* o) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

}  // namespace

extern "C" void orc_pcl_eigen33(const float C[9], float* eval, float evec[3]) {
  float scale = 0.f;
  for (int i = 0; i < 9; ++i) {
    float a = std::fabs(C[i]);
    if (a > scale) scale = a;
  }
  if (scale <= FLT_MIN) scale = 1.f;
  float m[9];
  for (int i = 0; i < 9; ++i) m[i] = C[i] / scale;
  float roots[3];
  pcl_roots(m, roots);
  *eval = roots[0] * scale;
  m[0] -= roots[0];
  m[4] -= roots[0];
  m[8] -= roots[0];
  float v1[3], v2[3], v3[3];
  cross3(m + 0, m + 3, v1);
  cross3(m + 0, m + 6, v2);
  cross3(m + 3, m + 6, v3);
  float l1 = v1[0] * v1[0] + v1[1] * v1[1] + v1[2] * v1[2];
  float l2 = v2[0] * v2[0] + v2[1] * v2[1] + v2[2] * v2[2];
  float l3 = v3[0] * v3[0] + v3[1] * v3[1] + v3[2] * v3[2];
  const float* v;
  float l;
  if (l1 >= l2 && l1 >= l3) { v = v1; l = l1; }
  else if (l2 >= l1 && l2 >= l3) { v = v2; l = l2; }
  else { v = v3; l = l3; }
  float s = std::sqrt(l);
  for (int i = 0; i < 3; ++i) evec[i] = v[i] / s;
}

extern "C" void orc_structure_features(float la1, float la2, float la3, float out3[3]) {
  // src/pcaAnalysis.cpp:29-31 — sqrt(float) picks the float overload; fp32 throughout.
  float s1 = std::sqrt(la1), s2 = std::sqrt(la2), s3 = std::sqrt(la3);
  out3[0] = (s1 - s2) / s1;
  out3[1] = (s2 - s3) / s1;
  out3[2] = s3 / s1;
}
