// oracle/pca.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// Restated control flow of src/pcaAnalysis.cpp:11-121 and of the PCL normal/curvature twin
// (src/smoothing.cpp:236-248) on top of search.cpp / eigen3.cpp.  Timed loops do not print
// (the reference prints nothing inside calculatePCAofPointCloud either).
#include "pcp_oracle.h"

#include <cmath>
#include <cstring>
#include <vector>

extern "C" void orc_covariance_about(const float* xyzi, const int32_t* idx, int64_t m, const float centre[3],
                                     float C[9]) {
  // pcl::computeCovarianceMatrixNormalized, dense branch — SURVEY A.2 accumulation order.
  float c00 = 0.f, c01 = 0.f, c02 = 0.f, c11 = 0.f, c12 = 0.f, c22 = 0.f;
  for (int64_t j = 0; j < m; ++j) {
    const float* p = xyzi + 4 * (int64_t)idx[j];
    float px = p[0] - centre[0];
    float py = p[1] - centre[1];
    float pz = p[2] - centre[2];
    c11 += py * py;
    c12 += py * pz;
    c22 += pz * pz;
    float t0 = px * px, t1 = py * px, t2 = pz * px;
    c00 += t0;
    c01 += t1;
    c02 += t2;
  }
  C[0] = c00; C[1] = c01; C[2] = c02;
  C[3] = c01; C[4] = c11; C[5] = c12;
  C[6] = c02; C[7] = c12; C[8] = c22;
  if (m != 0) {
    float fn = (float)m;
    for (int i = 0; i < 9; ++i) C[i] = C[i] / fn;
  }
}

namespace {

// calculatePCAofPoint body after the gate (src/pcaAnalysis.cpp:53-86): moment about
// `centre`, QR eigen, λ ascending into la1..la3, eigenvector columns into v1..v3.
void pca_fill(const float* xyzi, const int32_t* idx, int64_t m, const float centre[3], orc_feature68* f) {
  float C[9], ev[3], Q[9];
  orc_covariance_about(xyzi, idx, m, centre, C);
  orc_eigen_sym3_qr(C, ev, Q);
  f->la1 = ev[0];
  f->la2 = ev[1];
  f->la3 = ev[2];
  for (int r = 0; r < 3; ++r) {
    f->v1[r] = Q[3 * r + 0];
    f->v2[r] = Q[3 * r + 1];
    f->v3[r] = Q[3 * r + 2];
  }
}

void features_fill(orc_feature68* f) {
  float s[3];
  orc_structure_features(f->la1, f->la2, f->la3, s);
  f->lineStruc = s[0];
  f->planeStruc = s[1];
  f->scatterStruc = s[2];
}

// stale-tempFea semantics (Q4): invalid records repeat everything but ptID/ptNum from the
// previous record; leading ones are zero (the reference leaves them uninitialised).
void apply_carry(orc_feature68* out, const uint8_t* valid, int64_t n) {
  orc_feature68 last;
  std::memset(&last, 0, sizeof(last));
  for (int64_t i = 0; i < n; ++i) {
    if (valid[i]) {
      last = out[i];
    } else {
      int32_t id = out[i].ptID, num = out[i].ptNum;
      out[i] = last;
      out[i].ptID = id;
      out[i].ptNum = num;
    }
  }
}

}  // namespace

extern "C" void orc_radius_pca_cloud(const float* xyzi, int64_t n, float radius, int min_pts,
                                     int carry_forward, orc_feature68* out, uint8_t* valid_out,
                                     int nthreads) {
  orc_kdtree* tree = orc_kdtree_build(xyzi, n, 15);  // src/pcaAnalysis.cpp:14-15
  const float r2 = orc_radius2(radius);
  std::vector<uint8_t> valid((size_t)n, 0);
  if (nthreads <= 1) {
    // Faithful form: one tempFea outside the loop, result vectors re-created per point,
    // indices handed over by value (Q8).
    orc_feature68 temp;
    std::memset(&temp, 0, sizeof(temp));
    for (int64_t i = 0; i < n; ++i) {
      std::vector<int32_t> nn;
      std::vector<float> nd;
      int64_t m = orc_kdtree_radius(tree, i, r2, nullptr, nullptr, 0);
      nn.resize((size_t)m);
      nd.resize((size_t)m);
      orc_kdtree_radius(tree, i, r2, nn.data(), nd.data(), m);
      std::vector<int32_t> copy(nn);  // by-value parameter
      temp.ptID = (int32_t)i;
      temp.ptNum = (int32_t)copy.size();
      if (temp.ptNum >= min_pts) {
        pca_fill(xyzi, copy.data(), m, xyzi + 4 * i, &temp);
        valid[(size_t)i] = 1;
      } else if (!carry_forward) {
        int32_t id = temp.ptID, num = temp.ptNum;
        std::memset(&temp, 0, sizeof(temp));
        temp.ptID = id;
        temp.ptNum = num;
      }
      // :29-31 recompute the features from (possibly stale) λ every iteration; for a stale
      // record that reproduces the stale features bit for bit, so only do it when valid
      // (keeps the zero-filled leading records zero instead of 0/0).
      if (valid[(size_t)i]) features_fill(&temp);
      out[i] = temp;
    }
  } else {
#pragma omp parallel for schedule(dynamic, 512) num_threads(nthreads)
    for (int64_t i = 0; i < n; ++i) {
      std::vector<int32_t> nn;
      std::vector<float> nd;
      int64_t m = orc_kdtree_radius(tree, i, r2, nullptr, nullptr, 0);
      nn.resize((size_t)m);
      nd.resize((size_t)m);
      orc_kdtree_radius(tree, i, r2, nn.data(), nd.data(), m);
      orc_feature68 f;
      std::memset(&f, 0, sizeof(f));
      f.ptID = (int32_t)i;
      f.ptNum = (int32_t)m;
      if (m >= min_pts) {
        pca_fill(xyzi, nn.data(), m, xyzi + 4 * i, &f);
        features_fill(&f);
        valid[(size_t)i] = 1;
      }
      out[i] = f;
    }
    if (carry_forward) apply_carry(out, valid.data(), n);
  }
  if (valid_out) std::memcpy(valid_out, valid.data(), (size_t)n);
  orc_kdtree_free(tree);
}

extern "C" void orc_knn_pca_cloud(const float* xyzi, int64_t n, int k, int center_mode, orc_feature68* out,
                                  float* normals4, int nthreads) {
  orc_kdtree* tree = orc_kdtree_build(xyzi, n, 15);
#pragma omp parallel for schedule(dynamic, 512) num_threads(nthreads > 0 ? nthreads : 1)
  for (int64_t i = 0; i < n; ++i) {
    std::vector<int32_t> nn((size_t)k);
    std::vector<float> nd((size_t)k);
    int64_t m = orc_kdtree_knn(tree, i, k, nn.data(), nd.data());
    if (out) {
      orc_feature68 f;
      std::memset(&f, 0, sizeof(f));
      f.ptID = (int32_t)i;
      f.ptNum = (int32_t)m;
      float centre[3] = {xyzi[4 * i], xyzi[4 * i + 1], xyzi[4 * i + 2]};
      if (center_mode == 1) {
        // pcl::compute3DCentroid: fp32 running sum in list order, then / n
        float sx = 0.f, sy = 0.f, sz = 0.f;
        for (int64_t j = 0; j < m; ++j) {
          const float* p = xyzi + 4 * (int64_t)nn[(size_t)j];
          sx += p[0];
          sy += p[1];
          sz += p[2];
        }
        centre[0] = sx / (float)m;
        centre[1] = sy / (float)m;
        centre[2] = sz / (float)m;
      }
      pca_fill(xyzi, nn.data(), m, centre, &f);
      features_fill(&f);
      out[i] = f;
    }
    if (normals4) {
      // pcl::NormalEstimation::computePointNormal — SURVEY A.4
      float* o = normals4 + 4 * i;
      if (m < 3) {
        o[0] = o[1] = o[2] = o[3] = NAN;
        continue;
      }
      float a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      for (int64_t j = 0; j < m; ++j) {
        const float* p = xyzi + 4 * (int64_t)nn[(size_t)j];
        a[0] += p[0] * p[0];
        a[1] += p[0] * p[1];
        a[2] += p[0] * p[2];
        a[3] += p[1] * p[1];
        a[4] += p[1] * p[2];
        a[5] += p[2] * p[2];
        a[6] += p[0];
        a[7] += p[1];
        a[8] += p[2];
      }
      for (int t = 0; t < 9; ++t) a[t] /= (float)m;
      float C[9];
      C[0] = a[0] - a[6] * a[6];
      C[1] = a[1] - a[6] * a[7];
      C[2] = a[2] - a[6] * a[8];
      C[4] = a[3] - a[7] * a[7];
      C[5] = a[4] - a[7] * a[8];
      C[8] = a[5] - a[8] * a[8];
      C[3] = C[1];
      C[6] = C[2];
      C[7] = C[5];
      float ev, v[3];
      orc_pcl_eigen33(C, &ev, v);
      float tr = C[0] + C[4] + C[8];
      float curv = tr != 0.f ? std::fabs(ev / tr) : 0.f;
      const float* p = xyzi + 4 * i;
      float cs = (0.f - p[0]) * v[0] + (0.f - p[1]) * v[1] + (0.f - p[2]) * v[2];
      if (cs < 0.f) {
        v[0] = -v[0];
        v[1] = -v[1];
        v[2] = -v[2];
      }
      o[0] = v[0];
      o[1] = v[1];
      o[2] = v[2];
      o[3] = curv;
    }
  }
  orc_kdtree_free(tree);
}

extern "C" void orc_segment_pca(const float* xyzi, const float* centres_xyz, const int64_t* seg_offsets,
                                const int32_t* seg_idx, int64_t n_seg, orc_feature68* out) {
  for (int64_t s = 0; s < n_seg; ++s) {
    orc_feature68 f;
    std::memset(&f, 0, sizeof(f));
    int64_t b = seg_offsets[s], e = seg_offsets[s + 1];
    f.ptNum = (int32_t)(e - b);  // ptID is not set by this overload (src/pcaAnalysis.cpp:92)
    pca_fill(xyzi, seg_idx + b, e - b, centres_xyz + 3 * s, &f);
    features_fill(&f);
    out[s] = f;
  }
}

// pcaAnalysist::getMiniBoundingBox (src/pcaAnalysis.cpp:123-148), statement by statement, with the result returned
// instead of discarded (and a real cloud where the reference dereferences a null pointer).
// out: centroid[3], eigval[3], axes[9] (rows: eigenvector 0, eigenvector 1, their cross product), min[3], max[3],
// mean_diag[3] = 24 floats.
extern "C" void orc_bbox_pca(const float* xyzi, int64_t n, float* out) {
  // pcl::compute3DCentroid (:125-126): fp32 running sums in index order, divided by n
  float sx = 0.f, sy = 0.f, sz = 0.f;
  for (int64_t i = 0; i < n; ++i) {
    sx += xyzi[4 * i];
    sy += xyzi[4 * i + 1];
    sz += xyzi[4 * i + 2];
  }
  const float centre[3] = {sx / (float)n, sy / (float)n, sz / (float)n};
  // pcl::computeCovarianceMatrixNormalized(cloud, centroid, cov) (:128-129): the whole-cloud overload, same loop
  std::vector<int32_t> all((size_t)n);
  for (int64_t i = 0; i < n; ++i) all[(size_t)i] = (int32_t)i;
  float C[9], ev[3], Q[9];
  orc_covariance_about(xyzi, all.data(), n, centre, C);
  orc_eigen_sym3_qr(C, ev, Q);  // :131-132, columns ascending
  float a[3], b[3], c[3];
  for (int r = 0; r < 3; ++r) {
    a[r] = Q[3 * r + 0];
    b[r] = Q[3 * r + 1];
  }
  c[0] = a[1] * b[2] - a[2] * b[1];  // pcaVecs.col(2) = col(0).cross(col(1)) (:133)
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
  const float* R[3] = {a, b, c};  // transMatrix.block<3,3>(0,0) = pcaVecs.transpose() (:138)
  float t[3];
  for (int k = 0; k < 3; ++k) t[k] = -1.f * (R[k][0] * centre[0] + R[k][1] * centre[1] + R[k][2] * centre[2]);  // :139
  float lo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, hi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
  for (int64_t i = 0; i < n; ++i) {  // pcl::transformPointCloud + pcl::getMinMax3D (:142-145)
    const float* p = xyzi + 4 * i;
    for (int k = 0; k < 3; ++k) {
      const float v = R[k][0] * p[0] + R[k][1] * p[1] + R[k][2] * p[2] + t[k];
      lo[k] = v < lo[k] ? v : lo[k];
      hi[k] = v > hi[k] ? v : hi[k];
    }
  }
  for (int k = 0; k < 3; ++k) {
    out[k] = centre[k];
    out[3 + k] = ev[k];
    out[6 + k] = a[k];
    out[9 + k] = b[k];
    out[12 + k] = c[k];
    out[15 + k] = lo[k];
    out[18 + k] = hi[k];
    out[21 + k] = 0.5f * (hi[k] + lo[k]);  // meanDiag (:146)
  }
}
