// consumers_test.cpp — drives the drop-in consumers like a reference caller.  Used by tests/test_gpu_dropin.py:
//   consumers_test odom  cloud.bin out_range.bin out_ptid.bin out_type.bin     (16 x 900 floats each, row-major)
//   consumers_test dbscan cloud.bin r minpts out.bin      (int32 clusterID, int32 n_kept, then n_kept x 9 floats)
//   consumers_test meanshift cloud.bin iters out.bin      (n x 3 floats)
//   consumers_test txt|xyz file out.bin                   (1 + n points x 4 floats; the first is the caller's own)
//   consumers_test supervoxels cloud.bin cell out.bin     (makeSuperVoxels over voxels of `cell` metres: see the test)
// cloud.bin: n x 4 float32 (x, y, z, intensity).
#include <cstdint>
#include <cstdio>
#include <utility>
#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "consumers.h"

static pcXYZIptr load(const char* path) {
  pcXYZIptr cloud(new pcXYZI());
  FILE* f = std::fopen(path, "rb");
  if (!f) return cloud;
  float p[4];
  while (std::fread(p, sizeof(float), 4, f) == 4) cloud->push_back(pcl::PointXYZI(p[0], p[1], p[2], p[3]));
  std::fclose(f);
  return cloud;
}

static void dump(const char* path, const Eigen::MatrixXf& m) {
  FILE* o = std::fopen(path, "wb");
  for (int i = 0; i < m.rows(); ++i)
    for (int j = 0; j < m.cols(); ++j) {
      const float v = m(i, j);
      std::fwrite(&v, 4, 1, o);
    }
  std::fclose(o);
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  pcXYZIptr cloud = std::strcmp(argv[1], "txt") ? load(argv[2]) : pcXYZIptr(new pcXYZI());
  if (!std::strcmp(argv[1], "odom") && argc >= 6) {
    lidarOdometry::lidarOdometryClass lo;
    Eigen::MatrixXf rng, pid, ty;
    if (!lo.getRangeAndptIdImage(cloud, rng, pid)) return 1;
    if (!lo.calculateSmoothness(rng, ty)) return 1;
    dump(argv[3], rng);
    dump(argv[4], pid);
    dump(argv[5], ty);
    return 0;
  }
  if (!std::strcmp(argv[1], "dbscan") && argc >= 6) {
    pcXYZIptr cen(new pcXYZI()), mn(new pcXYZI()), mx(new pcXYZI());
    const int32_t id = DynaTools::clusterDBSCAN(cloud, std::atof(argv[3]), std::atoi(argv[4]), cen, mn, mx);
    FILE* o = std::fopen(argv[5], "wb");
    const int32_t kept = (int32_t)cen->points.size();
    std::fwrite(&id, 4, 1, o);
    std::fwrite(&kept, 4, 1, o);
    for (int32_t c = 0; c < kept; ++c) {
      const float v[9] = {cen->points[c].x, cen->points[c].y, cen->points[c].z, mn->points[c].x, mn->points[c].y,
                          mn->points[c].z,  mx->points[c].x,  mx->points[c].y,  mx->points[c].z};
      std::fwrite(v, 4, 9, o);
    }
    std::fclose(o);
    return 0;
  }
  if ((!std::strcmp(argv[1], "txt") || !std::strcmp(argv[1], "xyz")) && argc >= 4) {  // consumers_test txt|xyz file out.bin  (n x 4 floats)
    preprocess pre;
    pcXYZI cl;
    cl.push_back(pcl::PointXYZI(-1.f, -2.f, -3.f, -4.f));  // txt2pc / xyz2pc append
    if (!(argv[1][0] == 'x' ? pre.xyz2pc(argv[2], cl) : pre.txt2pc(argv[2], cl))) return 1;
    FILE* o = std::fopen(argv[3], "wb");
    for (size_t i = 0; i < cl.points.size(); ++i) {
      const float v[4] = {cl.points[i].x, cl.points[i].y, cl.points[i].z, cl.points[i].intensity};
      std::fwrite(v, 4, 4, o);
    }
    std::fclose(o);
    return 0;
  }
  if (!std::strcmp(argv[1], "supervoxels") && argc >= 5) {  // consumers_test supervoxels cloud.bin cell out.bin: voxels of `cell` metres, >= 20 points
    const float cell = (float)std::atof(argv[3]);
    std::vector<buildingFacadeExtractor::voxel> vox;
    {
      std::vector<std::pair<long long, int> > key(cloud->points.size());
      for (size_t i = 0; i < cloud->points.size(); ++i) {
        const pcl::PointXYZI& p = cloud->points[i];
        const long long a = (long long)std::floor(p.x / cell), b = (long long)std::floor(p.y / cell), c = (long long)std::floor(p.z / cell);
        key[i] = std::make_pair((a * 100003LL + b) * 100003LL + c, (int)i);
      }
      std::sort(key.begin(), key.end());
      for (size_t i = 0; i < key.size();) {
        size_t j = i;
        buildingFacadeExtractor::voxel v;
        while (j < key.size() && key[j].first == key[i].first) v.ptIndices.push_back(key[j++].second);
        if (v.ptIndices.size() >= 20) {  // src/buildingFacadeExtraction.cpp:172
          const pcl::PointXYZI& p = cloud->points[(size_t)v.ptIndices[0]];
          v.id = (int)vox.size();
          v.ptNums = (int)v.ptIndices.size();
          v.centroid = pcl::PointXYZI((std::floor(p.x / cell) + 0.5f) * cell, (std::floor(p.y / cell) + 0.5f) * cell, (std::floor(p.z / cell) + 0.5f) * cell, 0.f);
          vox.push_back(v);
        }
        i = j;
      }
    }
    buildingFacadeExtractor bfe;
    std::vector<buildingFacadeExtractor::voxel> super;
    std::vector<pcaAnalysist::pcaFeature> fea;
    if (!bfe.makeSuperVoxels(vox.data(), (int)vox.size(), cloud, super, &fea)) return 1;
    // the same voxels one by one through the class method the reference calls (:212-219)
    pcaAnalysist pca;
    FILE* o = std::fopen(argv[4], "wb");
    const int32_t nv = (int32_t)vox.size(), ns = (int32_t)super.size();
    std::fwrite(&nv, 4, 1, o);
    std::fwrite(&ns, 4, 1, o);
    for (size_t v = 0; v < vox.size(); ++v) {
      pcaAnalysist::pcaFeature one;
      std::memset(&one, 0, sizeof(one));
      if (v < 64 && !pca.calculatePCAofPoint(cloud, vox[v].centroid, vox[v].ptIndices, one)) return 1;
      std::fwrite(&fea[v], sizeof(one), 1, o);
      std::fwrite(&one, sizeof(one), 1, o);
      std::fwrite(&vox[v].centroid.x, 4, 3, o);
      const int32_t m = (int32_t)vox[v].ptIndices.size();
      std::fwrite(&m, 4, 1, o);
      std::fwrite(vox[v].ptIndices.data(), 4, (size_t)m, o);
    }
    std::fclose(o);
    return 0;
  }
  if (!std::strcmp(argv[1], "meanshift") && argc >= 5) {
    buildingFacadeExtractor bfe;
    if (!bfe.meanShiftClustering(cloud, std::atoi(argv[3]))) return 1;
    FILE* o = std::fopen(argv[4], "wb");
    for (size_t i = 0; i < cloud->points.size(); ++i) std::fwrite(&cloud->points[i].x, 4, 3, o);
    std::fclose(o);
    return 0;
  }
  return 2;
}
