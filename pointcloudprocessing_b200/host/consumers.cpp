// consumers.cpp — thin adapters from the reference's signatures to the pcp_* C-ABI (include/pcp.h → libpcp.so).
// Replaces src/lidarOdometry.cpp:21-187, include/dbscanCLuster.h:256-345 and
// src/buildingFacadeExtraction.cpp:59-132.  Host work is packing and hand-over only.
#include "consumers.h"

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "pcp.h"

namespace {

struct ThreadCtx {
  pcp_ctx* ctx = nullptr;
  bool failed = false;
  ~ThreadCtx() {
    if (ctx) pcp_destroy(ctx);
  }
};

pcp_ctx* context() {
  static thread_local ThreadCtx t;
  if (t.ctx || t.failed) return t.ctx;
  const char* dev = std::getenv("PCP_DEVICE");
  const int rc = pcp_create(&t.ctx, dev ? std::atoi(dev) : 0);
  if (rc != PCP_OK) {
    std::fprintf(stderr, "pcp_create failed (%d): no sm_100 GPU / libpcp; there is no CPU path\n", rc);
    t.failed = true;
    t.ctx = nullptr;
  }
  return t.ctx;
}

bool fail(pcp_ctx* ctx, const char* where, int rc) {
  std::fprintf(stderr, "%s: pcp error %d: %s\n", where, rc, pcp_last_error(ctx));
  return false;
}

// the cloud as it is (32-byte PointXYZI records) + a grid with the given cell
int upload(pcp_ctx* ctx, const pcXYZI& cloud, float cell) {
  int rc = pcp_set_cloud_host(ctx, cloud.points.data(), (int64_t)cloud.points.size(), (int)sizeof(pcl::PointXYZI));
  if (rc) return rc;
  return pcp_build_grid(ctx, cell, 0);
}

}  // namespace

namespace lidarOdometry {

bool lidarOdometryClass::getRangeAndptIdImage(pcXYZIptr inputCloud, Eigen::MatrixXf& rangeImg, Eigen::MatrixXf& ptIdImg) {
  pcp_odom_params prm;
  pcp_odom_params_default(&prm);
  rangeImg = Eigen::MatrixXf::Zero(prm.n_rows, prm.n_cols);  // (:29-32)
  ptIdImg = Eigen::MatrixXf::Zero(prm.n_rows, prm.n_cols);
  if (!inputCloud || inputCloud->points.empty()) return false;  // the reference reads points[0] (:34)
  pcp_ctx* ctx = context();
  if (!ctx) return false;
  int rc = upload(ctx, *inputCloud, prm.radius);
  if (rc) return fail(ctx, "getRangeAndptIdImage", rc);
  const size_t cells = (size_t)prm.n_rows * prm.n_cols;
  std::vector<float> rng(cells), pid(cells);
  rc = pcp_odom_range_image(ctx, &prm, nullptr, 1, rng.data(), pid.data(), nullptr, nullptr, nullptr, 0);
  if (rc) return fail(ctx, "getRangeAndptIdImage", rc);
  for (int i = 0; i < prm.n_rows; ++i)
    for (int j = 0; j < prm.n_cols; ++j) {
      rangeImg(i, j) = rng[(size_t)i * prm.n_cols + j];
      ptIdImg(i, j) = pid[(size_t)i * prm.n_cols + j];
    }
  return true;
}

bool lidarOdometryClass::calculateSmoothness(Eigen::MatrixXf rangeImg, Eigen::MatrixXf& ptTypeImg) {
  pcp_odom_params prm;
  pcp_odom_params_default(&prm);
  ptTypeImg = Eigen::MatrixXf::Zero(prm.n_rows, prm.n_cols);  // (:124)
  if (rangeImg.rows() != prm.n_rows || rangeImg.cols() != prm.n_cols) return false;
  pcp_ctx* ctx = context();
  if (!ctx) return false;
  const size_t cells = (size_t)prm.n_rows * prm.n_cols;
  std::vector<float> rng(cells), ty(cells);
  for (int i = 0; i < prm.n_rows; ++i)
    for (int j = 0; j < prm.n_cols; ++j) rng[(size_t)i * prm.n_cols + j] = rangeImg(i, j);
  const int rc = pcp_odom_smoothness(ctx, &prm, rng.data(), 0, 1, nullptr, ty.data(), 0);
  if (rc) return fail(ctx, "calculateSmoothness", rc);
  for (int i = 0; i < prm.n_rows; ++i)
    for (int j = 0; j < prm.n_cols; ++j) ptTypeImg(i, j) = ty[(size_t)i * prm.n_cols + j];
  return true;
}

}  // namespace lidarOdometry

namespace DynaTools {

int clusterDBSCAN(pcXYZIptr inCloud, double r, int minptNum, pcXYZIptr pcClusterCentos, pcXYZIptr pcClusterMinpts,
                  pcXYZIptr pcClusterMaxpts) {
  if (!inCloud || inCloud->points.empty()) return 0;
  pcp_ctx* ctx = context();
  if (!ctx) return 0;
  const size_t n = inCloud->points.size();
  // cloud2d: a copy with z = 0 (:264-267)
  std::vector<float> flat(4 * n);
  for (size_t i = 0; i < n; ++i) {
    flat[4 * i] = inCloud->points[i].x;
    flat[4 * i + 1] = inCloud->points[i].y;
    flat[4 * i + 2] = 0.f;
    flat[4 * i + 3] = inCloud->points[i].intensity;
  }
  int rc = pcp_set_cloud_host(ctx, flat.data(), (int64_t)n, 16);
  if (!rc) rc = pcp_build_grid(ctx, (float)(r * 1.0001), 0);  // a cell just above r: a 3 x 3 x 3 stencil covers the radius
  std::vector<int32_t> state(n);
  int32_t clusterID = 0;
  if (!rc) rc = pcp_dbscan_d(ctx, r, minptNum, state.data(), &clusterID, 0);  // r stays the double the reference declares
  if (rc) {
    fail(ctx, "clusterDBSCAN", rc);
    return 0;
  }
  if (clusterID > 3) {  // (:308)
    std::vector<float> cen(3 * (size_t)clusterID), lo(3 * (size_t)clusterID), hi(3 * (size_t)clusterID);
    rc = pcp_set_cloud_host(ctx, inCloud->points.data(), (int64_t)n, (int)sizeof(pcl::PointXYZI));
    if (!rc) rc = pcp_cluster_boxes(ctx, state.data(), 0, clusterID, cen.data(), lo.data(), hi.data(), nullptr);
    if (rc) {
      fail(ctx, "clusterDBSCAN", rc);
      return clusterID;
    }
    for (int c = 0; c < clusterID; ++c) {
      const float *mn = &lo[3 * (size_t)c], *mx = &hi[3 * (size_t)c];
      if (mx[0] - mn[0] > 4 || mx[1] - mn[1] > 4 || mx[2] - mn[2] < 0.3) continue;  // (:318-319)
      if (pcClusterCentos) pcClusterCentos->push_back(pcl::PointXYZI(cen[3 * (size_t)c], cen[3 * (size_t)c + 1], cen[3 * (size_t)c + 2]));
      if (pcClusterMinpts) pcClusterMinpts->push_back(pcl::PointXYZI(mn[0], mn[1], mn[2]));
      if (pcClusterMaxpts) pcClusterMaxpts->push_back(pcl::PointXYZI(mx[0], mx[1], mx[2]));
    }
  }
  return clusterID;
}

}  // namespace DynaTools

bool buildingFacadeExtractor::makeSuperVoxels(voxel* voxelArrptr, int voxelNums, pcXYZIptr inputCloud, std::vector<voxel>& superVoxels,
                                              std::vector<pcaAnalysist::pcaFeature>* features) {
  if (features) features->clear();
  if (voxelNums <= 0) return true;
  if (!voxelArrptr || !inputCloud || inputCloud->points.empty()) return false;
  pcp_ctx* ctx = context();
  if (!ctx) return false;
  int rc = pcp_set_cloud_host(ctx, inputCloud->points.data(), (int64_t)inputCloud->points.size(), (int)sizeof(pcl::PointXYZI));
  if (rc) return fail(ctx, "makeSuperVoxels", rc);
  // CSR batch: every voxel is one segment about ITS centre (the voxel centre, not the data mean: :216-218)
  std::vector<float> centres(3 * (size_t)voxelNums);
  std::vector<int64_t> offs((size_t)voxelNums + 1, 0);
  for (int v = 0; v < voxelNums; ++v) offs[(size_t)v + 1] = offs[(size_t)v] + (int64_t)voxelArrptr[v].ptIndices.size();
  std::vector<int32_t> idx((size_t)offs[(size_t)voxelNums]);
  for (int v = 0; v < voxelNums; ++v) {
    centres[3 * (size_t)v] = voxelArrptr[v].centroid.x;
    centres[3 * (size_t)v + 1] = voxelArrptr[v].centroid.y;
    centres[3 * (size_t)v + 2] = voxelArrptr[v].centroid.z;
    std::copy(voxelArrptr[v].ptIndices.begin(), voxelArrptr[v].ptIndices.end(), idx.begin() + offs[(size_t)v]);
  }
  static_assert(sizeof(pcp_feature68) == sizeof(pcaAnalysist::pcaFeature), "layout");
  std::vector<pcaAnalysist::pcaFeature> rec((size_t)voxelNums);
  rc = pcp_segment_pca(ctx, centres.data(), offs.data(), idx.data(), voxelNums, reinterpret_cast<pcp_feature68*>(rec.data()));
  if (rc) return fail(ctx, "makeSuperVoxels", rc);
  for (int v = 0; v < voxelNums; ++v)
    if (rec[(size_t)v].planeStruc > 0.6) superVoxels.push_back(voxelArrptr[v]);  // (:219-220; never true upstream, SURVEY Q3: kept)
  if (features) features->swap(rec);
  return true;
}

bool buildingFacadeExtractor::meanShiftClustering(pcXYZIptr inputcloud, int iterTimes) {
  if (!inputcloud || inputcloud->points.empty()) return true;
  pcp_ctx* ctx = context();
  if (!ctx) return false;
  const size_t n = inputcloud->points.size();
  int rc = upload(ctx, *inputcloud, 2.0f);
  std::vector<float> out(4 * n);
  if (!rc) rc = pcp_mean_shift(ctx, 2.0f, 15, iterTimes, nullptr, 0, out.data(), nullptr, 0);  // (:79-80)
  if (rc) return fail(ctx, "meanShiftClustering", rc);
  for (size_t i = 0; i < n; ++i) {  // (:112-117)
    inputcloud->points[i].x = out[4 * i];
    inputcloud->points[i].y = out[4 * i + 1];
    inputcloud->points[i].z = out[4 * i + 2];
  }
  return true;
}

static bool text_cloud(const std::string& txtpath, pcXYZI& origiCloud, bool xyz) {
  FILE* f = std::fopen(txtpath.c_str(), "rb");
  if (!f) {
    std::fprintf(stderr, xyz ? "Wrong *.xyz file! \n" : "Wrong txt file! \n");  // (:47-49, :91-94); the reference still returns true
    return true;
  }
  std::vector<char> text;
  char chunk[1 << 16];
  size_t got;
  while ((got = std::fread(chunk, 1, sizeof(chunk), f)) > 0) text.insert(text.end(), chunk, chunk + got);
  std::fclose(f);
  pcp_ctx* ctx = context();
  if (!ctx) return false;
  size_t lines = 2;
  for (char c : text) lines += c == '\n';
  std::vector<float> pts(4 * lines);
  int64_t n = 0;
  const int rc = xyz ? pcp_set_cloud_xyz(ctx, text.data(), (int64_t)text.size(), 0, &n, pts.data(), (int64_t)lines, 0)
                     : pcp_set_cloud_txt(ctx, text.data(), (int64_t)text.size(), 0, &n, pts.data(), (int64_t)lines, 0);
  if (rc) return fail(ctx, xyz ? "xyz2pc" : "txt2pc", rc);
  for (int64_t i = 0; i < n; ++i) origiCloud.push_back(pcl::PointXYZI(pts[4 * i], pts[4 * i + 1], pts[4 * i + 2], pts[4 * i + 3]));
  return true;
}

bool preprocess::txt2pc(std::string txtpath, pcXYZI& origiCloud) { return text_cloud(txtpath, origiCloud, false); }
// getline(line, 256) + sscanf("%f %f %f %s") per line (src/smoothing.cpp:83-125); intensity = first byte of the 4th token
bool preprocess::xyz2pc(std::string txtpath, pcXYZI& origiCloud) { return text_cloud(txtpath, origiCloud, true); }
