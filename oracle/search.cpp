// oracle/search.cpp — TEST INFRASTRUCTURE (see pcp_oracle.h).  Exact fp32 radius / k-NN search.
//
// Restates the behaviour of pcl::KdTreeFLANN<PointXYZI> as called at
// src/pcaAnalysis.cpp:14-15,25: distance = FLANN L2_Simple<float>,
// fp32, ((dx*dx)+dy*dy)+dz*dz, no FMA; radius membership is strict d2 < r2 with
// r2 = (float)((double)r*(double)r); results ordered by (d2, index); k-NN = first k of that
// order (north-star tie rule).  The brute-force functions DEFINE the result; the kd-tree
// (single tree, leaf 15, built once like FLANN's KDTreeSingleIndex) must reproduce it
// exactly and is what the CPU baseline times.
#include "pcp_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <queue>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

typedef std::pair<float, int32_t> DI;  // (d2, index): operator< is exactly FLANN's DistanceIndex order

inline float dist2(const float* a, const float* b) {
  // L2_Simple<float>: result += diff*diff for d = 0,1,2 — separate multiply and add.
  float dx = a[0] - b[0];
  float dy = a[1] - b[1];
  float dz = a[2] - b[2];
  float r = dx * dx;
  r = r + dy * dy;
  r = r + dz * dz;
  return r;
}

}  // namespace

extern "C" float orc_radius2(float radius) { return (float)((double)radius * (double)radius); }

extern "C" int64_t orc_radius_brute(const float* xyzi, int64_t n, int64_t q, float r2, int32_t* idx,
                                    float* d2, int64_t cap) {
  std::vector<DI> hits;
  const float* pq = xyzi + 4 * q;
  for (int64_t j = 0; j < n; ++j) {
    float d = dist2(pq, xyzi + 4 * j);
    if (d < r2) hits.push_back(DI(d, (int32_t)j));
  }
  std::sort(hits.begin(), hits.end());
  int64_t m = (int64_t)hits.size();
  for (int64_t j = 0; j < m && j < cap; ++j) {
    if (idx) idx[j] = hits[j].second;
    if (d2) d2[j] = hits[j].first;
  }
  return m;
}

extern "C" int64_t orc_knn_brute(const float* xyzi, int64_t n, int64_t q, int k, int32_t* idx, float* d2) {
  std::vector<DI> all((size_t)n);
  const float* pq = xyzi + 4 * q;
  for (int64_t j = 0; j < n; ++j) all[(size_t)j] = DI(dist2(pq, xyzi + 4 * j), (int32_t)j);
  int64_t m = std::min<int64_t>(k, n);
  std::partial_sort(all.begin(), all.begin() + m, all.end());
  for (int64_t j = 0; j < m; ++j) {
    if (idx) idx[j] = all[(size_t)j].second;
    if (d2) d2[j] = all[(size_t)j].first;
  }
  return m;
}

// ------------------------------------------------------------------------------------
// kd-tree: median split on the widest box dimension, leaves of <= leaf_max points, points
// re-ordered into tree order (FLANN reorder=true).  Pruning uses the exact distance to a
// child's bounding box evaluated in double with a relative slack, so it can never drop a
// point the brute-force predicate would keep.
struct orc_kdtree {
  struct Node {
    int32_t left, right;  // children (−1 for a leaf)
    int32_t begin, end;   // point range in tree order
    float lo[3], hi[3];
  };
  const float* cloud;
  int64_t n;
  int leaf_max;
  std::vector<Node> nodes;
  std::vector<int32_t> order;  // tree position -> original index
  std::vector<float> pts;      // tree-ordered xyz (3 floats per point)
};

namespace {

int32_t build_rec(orc_kdtree* t, int32_t begin, int32_t end) {
  orc_kdtree::Node nd;
  nd.left = nd.right = -1;
  nd.begin = begin;
  nd.end = end;
  for (int d = 0; d < 3; ++d) {
    nd.lo[d] = INFINITY;
    nd.hi[d] = -INFINITY;
  }
  for (int32_t i = begin; i < end; ++i) {
    const float* p = t->cloud + 4 * (int64_t)t->order[(size_t)i];
    for (int d = 0; d < 3; ++d) {
      nd.lo[d] = std::min(nd.lo[d], p[d]);
      nd.hi[d] = std::max(nd.hi[d], p[d]);
    }
  }
  int32_t me = (int32_t)t->nodes.size();
  t->nodes.push_back(nd);
  if (end - begin > t->leaf_max) {
    int dim = 0;
    float span = nd.hi[0] - nd.lo[0];
    for (int d = 1; d < 3; ++d)
      if (nd.hi[d] - nd.lo[d] > span) {
        span = nd.hi[d] - nd.lo[d];
        dim = d;
      }
    if (span > 0.f) {
      int32_t mid = begin + (end - begin) / 2;
      const float* c = t->cloud;
      std::nth_element(t->order.begin() + begin, t->order.begin() + mid, t->order.begin() + end,
                       [c, dim](int32_t a, int32_t b) {
                         float va = c[4 * (int64_t)a + dim], vb = c[4 * (int64_t)b + dim];
                         return va < vb || (va == vb && a < b);
                       });
      int32_t l = build_rec(t, begin, mid);
      int32_t r = build_rec(t, mid, end);
      t->nodes[(size_t)me].left = l;
      t->nodes[(size_t)me].right = r;
    }
  }
  return me;
}

inline double box_dist2(const orc_kdtree::Node& nd, const float* q) {
  double s = 0.0;
  for (int d = 0; d < 3; ++d) {
    double v = 0.0;
    if (q[d] < nd.lo[d]) v = (double)nd.lo[d] - (double)q[d];
    else if (q[d] > nd.hi[d]) v = (double)q[d] - (double)nd.hi[d];
    s += v * v;
  }
  return s * (1.0 - 1e-6);  // slack: never prune what fp32 d2 might still accept
}

void radius_rec(const orc_kdtree* t, int32_t ni, const float* q, float r2, std::vector<DI>& hits) {
  const orc_kdtree::Node& nd = t->nodes[(size_t)ni];
  if (nd.left < 0) {
    for (int32_t i = nd.begin; i < nd.end; ++i) {
      float d = dist2(q, &t->pts[3 * (size_t)i]);
      if (d < r2) hits.push_back(DI(d, t->order[(size_t)i]));
    }
    return;
  }
  if (!(box_dist2(t->nodes[(size_t)nd.left], q) >= (double)r2)) radius_rec(t, nd.left, q, r2, hits);
  if (!(box_dist2(t->nodes[(size_t)nd.right], q) >= (double)r2)) radius_rec(t, nd.right, q, r2, hits);
}

typedef std::priority_queue<DI> MaxHeap;

void knn_rec(const orc_kdtree* t, int32_t ni, const float* q, size_t k, MaxHeap& heap) {
  const orc_kdtree::Node& nd = t->nodes[(size_t)ni];
  if (nd.left < 0) {
    for (int32_t i = nd.begin; i < nd.end; ++i) {
      DI c(dist2(q, &t->pts[3 * (size_t)i]), t->order[(size_t)i]);
      if (heap.size() < k) heap.push(c);
      else if (c < heap.top()) {
        heap.pop();
        heap.push(c);
      }
    }
    return;
  }
  double dl = box_dist2(t->nodes[(size_t)nd.left], q);
  double dr = box_dist2(t->nodes[(size_t)nd.right], q);
  int32_t first = nd.left, second = nd.right;
  double d1 = dl, d2v = dr;
  if (dr < dl) {
    first = nd.right;
    second = nd.left;
    d1 = dr;
    d2v = dl;
  }
  if (heap.size() < k || !(d1 > (double)heap.top().first)) knn_rec(t, first, q, k, heap);
  if (heap.size() < k || !(d2v > (double)heap.top().first)) knn_rec(t, second, q, k, heap);
}

}  // namespace

extern "C" orc_kdtree* orc_kdtree_build(const float* xyzi, int64_t n, int leaf_max) {
  orc_kdtree* t = new orc_kdtree();
  t->cloud = xyzi;
  t->n = n;
  t->leaf_max = leaf_max > 0 ? leaf_max : 15;
  t->order.resize((size_t)n);
  for (int64_t i = 0; i < n; ++i) t->order[(size_t)i] = (int32_t)i;
  t->nodes.reserve((size_t)(2 * n / t->leaf_max + 16));
  if (n > 0) build_rec(t, 0, (int32_t)n);
  t->pts.resize(3 * (size_t)n);
  for (int64_t i = 0; i < n; ++i)
    for (int d = 0; d < 3; ++d) t->pts[3 * (size_t)i + d] = xyzi[4 * (int64_t)t->order[(size_t)i] + d];
  return t;
}

extern "C" void orc_kdtree_free(orc_kdtree* t) { delete t; }

extern "C" int64_t orc_kdtree_radius(const orc_kdtree* t, int64_t q, float r2, int32_t* idx, float* d2,
                                     int64_t cap) {
  std::vector<DI> hits;
  if (t->n > 0) radius_rec(t, 0, t->cloud + 4 * q, r2, hits);
  std::sort(hits.begin(), hits.end());
  int64_t m = (int64_t)hits.size();
  for (int64_t j = 0; j < m && j < cap; ++j) {
    if (idx) idx[j] = hits[(size_t)j].second;
    if (d2) d2[j] = hits[(size_t)j].first;
  }
  return m;
}

// radius search around an arbitrary position (pcl::KdTreeFLANN::radiusSearch(const PointT&, ...)): indices in
// (d2, index) order.  C++ linkage: used by consumers.cpp.
void orc_kdtree_radius_vec(const orc_kdtree* t, const float* q, float r2, std::vector<int32_t>& out) {
  std::vector<DI> hits;
  if (t->n > 0) radius_rec(t, 0, q, r2, hits);
  std::sort(hits.begin(), hits.end());
  out.resize(hits.size());
  for (size_t j = 0; j < hits.size(); ++j) out[j] = hits[j].second;
}

extern "C" int64_t orc_kdtree_knn(const orc_kdtree* t, int64_t q, int k, int32_t* idx, float* d2) {
  MaxHeap heap;
  if (t->n > 0 && k > 0) knn_rec(t, 0, t->cloud + 4 * q, (size_t)k, heap);
  int64_t m = (int64_t)heap.size();
  for (int64_t j = m - 1; j >= 0; --j) {
    if (idx) idx[j] = heap.top().second;
    if (d2) d2[j] = heap.top().first;
    heap.pop();
  }
  return m;
}

extern "C" void orc_kdtree_knn_all(const orc_kdtree* t, int k, int32_t* idx, float* d2, int nthreads) {
  int64_t n = t->n;
#pragma omp parallel for schedule(dynamic, 1024) num_threads(nthreads > 0 ? nthreads : 1)
  for (int64_t i = 0; i < n; ++i) {
    int64_t m = orc_kdtree_knn(t, i, k, idx ? idx + i * k : nullptr, d2 ? d2 + i * k : nullptr);
    for (int64_t j = m; j < k; ++j) {
      if (idx) idx[i * k + j] = -1;
      if (d2) d2[i * k + j] = INFINITY;
    }
  }
}

extern "C" void orc_kdtree_radius_counts(const orc_kdtree* t, float r2, int32_t* counts, int nthreads) {
  int64_t n = t->n;
#pragma omp parallel for schedule(dynamic, 1024) num_threads(nthreads > 0 ? nthreads : 1)
  for (int64_t i = 0; i < n; ++i) counts[i] = (int32_t)orc_kdtree_radius(t, i, r2, nullptr, nullptr, 0);
}

extern "C" void orc_kdtree_radius_all(const orc_kdtree* t, float r2, const int64_t* offsets, int32_t* idx,
                                      float* d2, int nthreads) {
  int64_t n = t->n;
#pragma omp parallel for schedule(dynamic, 1024) num_threads(nthreads > 0 ? nthreads : 1)
  for (int64_t i = 0; i < n; ++i)
    orc_kdtree_radius(t, i, r2, idx ? idx + offsets[i] : nullptr, d2 ? d2 + offsets[i] : nullptr,
                      offsets[i + 1] - offsets[i]);
}

extern "C" int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
