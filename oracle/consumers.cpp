// oracle/consumers.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).
//
// The two consumers of radiusSearch next to the PCA stage, restated literally:
//  * DynaTools::clusterDBSCAN (include/dbscanCLuster.h:256-345): sequential seed loop + list-growing expansion,
//    then per-cluster centroid / min / max.
//  * buildingFacadeExtractor::meanShiftClustering (src/buildingFacadeExtraction.cpp:59-132).
#include "pcp_oracle.h"

#include <cfloat>
#include <cmath>
#include <vector>

struct orc_kdtree;
void orc_kdtree_radius_vec(const orc_kdtree* t, const float* q, float r2, std::vector<int32_t>& out);

// state: 0 unvisited, -1 visited, -2 noise, > 0 cluster id (:262).  The caller passes the cloud the reference
// searches in (cloud2d: z already zeroed, :264-267).  Returns clusterID.
extern "C" int32_t orc_dbscan_d(const float* xyzi, int64_t n, double radius, int min_pts, int32_t* state);
extern "C" int32_t orc_dbscan(const float* xyzi, int64_t n, float radius, int min_pts, int32_t* state) {
  return orc_dbscan_d(xyzi, n, (double)radius, min_pts, state);
}
// the reference's own parameter type: double r (include/dbscanCLuster.h:256); pcl::KdTreeFLANN::radiusSearch(p, double
// radius, ...) hands FLANN static_cast<float>(radius * radius)
extern "C" int32_t orc_dbscan_d(const float* xyzi, int64_t n, double radius, int min_pts, int32_t* state) {
  orc_kdtree* tree = orc_kdtree_build(xyzi, n, 15);
  const float r2 = (float)(radius * radius);
  for (int64_t i = 0; i < n; ++i) state[i] = 0;
  int32_t cluster = 0;
  std::vector<int32_t> nb, tmp;
  for (int64_t i = 0; i < n; ++i) {
    if (state[i] != 0) continue;
    state[i] = -1;  // (:278)
    orc_kdtree_radius_vec(tree, xyzi + 4 * i, r2, nb);
    if ((int64_t)nb.size() < min_pts) {
      state[i] = -2;  // (:280-281)
      continue;
    }
    ++cluster;
    state[i] = cluster;
    for (size_t j = 0; j < nb.size(); ++j) {  // the list grows while it is walked (:285-296)
      const int32_t q = nb[j];
      if (state[q] == 0) {
        state[q] = -1;
        orc_kdtree_radius_vec(tree, xyzi + 4 * (int64_t)q, r2, tmp);
        if ((int64_t)tmp.size() >= min_pts) nb.insert(nb.end(), tmp.begin(), tmp.end());
      }
      if (state[q] == -1) state[q] = cluster;
    }
  }
  orc_kdtree_free(tree);
  return cluster;
}

// per cluster (ids 1..n_clusters in `state`): pcl::compute3DCentroid (fp32 sums in index order, / n) and
// pcl::getMinMax3D over the cluster's points of the ORIGINAL cloud (:312-325).
extern "C" void orc_cluster_boxes(const float* xyzi, int64_t n, const int32_t* state, int32_t n_clusters, float* centroid3,
                                  float* min3, float* max3, int32_t* counts) {
  for (int32_t c = 0; c < n_clusters; ++c) {
    float s[3] = {0.f, 0.f, 0.f}, lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    int32_t m = 0;
    for (int64_t i = 0; i < n; ++i) {
      if (state[i] != c + 1) continue;
      ++m;
      for (int a = 0; a < 3; ++a) {
        const float v = xyzi[4 * i + a];
        s[a] += v;
        if (v < lo[a]) lo[a] = v;
        if (v > hi[a]) hi[a] = v;
      }
    }
    for (int a = 0; a < 3; ++a) {
      centroid3[3 * c + a] = m ? s[a] / (float)m : 0.f;
      min3[3 * c + a] = lo[a];
      max3[3 * c + a] = hi[a];
    }
    counts[c] = m;
  }
}

// One sweep of the do-while body (:72-123): the tree holds the ORIGINAL positions (FLANN copies the cloud at
// setInputCloud, :63-64), the query is the point's CURRENT position (radiusSearch(i, ...) reads input_->points[i],
// which the loop moves, :79), and the shift is accumulated from the CURRENT positions of the neighbours found
// (:96-104) as fp32 `tmp += diff / size`, in (d2, index) order.  next = cur + shift (:112-117).
extern "C" void orc_mean_shift_step(const float* orig_xyzi, const float* cur_xyzi, int64_t n, float radius, int min_pts,
                                    float* next_xyzi, float* shift_xyz, int32_t* counts) {
  orc_kdtree* tree = orc_kdtree_build(orig_xyzi, n, 15);
  const float r2 = orc_radius2(radius);
#pragma omp parallel
  {
    std::vector<int32_t> nb;
#pragma omp for schedule(dynamic, 256)
    for (int64_t i = 0; i < n; ++i) {
      float t[3] = {0.f, 0.f, 0.f};
      orc_kdtree_radius_vec(tree, cur_xyzi + 4 * i, r2, nb);
      if ((int64_t)nb.size() >= min_pts) {  // (:81-85)
        const float m = (float)nb.size();
        for (size_t j = 0; j < nb.size(); ++j)
          for (int a = 0; a < 3; ++a) t[a] += (cur_xyzi[4 * (int64_t)nb[j] + a] - cur_xyzi[4 * i + a]) / m;
      }
      for (int a = 0; a < 3; ++a) {
        if (shift_xyz) shift_xyz[3 * i + a] = t[a];
        next_xyzi[4 * i + a] = cur_xyzi[4 * i + a] + t[a];
      }
      next_xyzi[4 * i + 3] = cur_xyzi[4 * i + 3];
      if (counts) counts[i] = (int32_t)nb.size();
    }
  }
  orc_kdtree_free(tree);
}
