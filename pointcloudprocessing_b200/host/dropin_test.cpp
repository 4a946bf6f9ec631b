// dropin_test.cpp — exercises the pcaAnalysist class exactly as a reference caller would
// (main.cpp:115-117; src/buildingFacadeExtraction.cpp:212-219).  Used by tests/test_gpu_dropin.py:
//   dropin_test cloud.bin radius out_cloud.bin [segments.bin out_segments.bin]
// cloud.bin: n × 4 float32 (x,y,z,intensity); segments.bin: int64 n_seg, then per segment
// {float cx,cy,cz; int64 m; int32 idx[m]}.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "pcaAnalysis.h"

static std::vector<char> slurp(const char* path) {
  std::vector<char> b;
  FILE* f = std::fopen(path, "rb");
  if (!f) return b;
  std::fseek(f, 0, SEEK_END);
  long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  b.resize((size_t)sz);
  if (sz > 0 && std::fread(b.data(), 1, (size_t)sz, f) != (size_t)sz) b.clear();
  std::fclose(f);
  return b;
}

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  std::vector<char> raw = slurp(argv[1]);
  const size_t n = raw.size() / 16;
  const float* xyzi = reinterpret_cast<const float*>(raw.data());
  pcXYZIptr cloud(new pcXYZI());
  for (size_t i = 0; i < n; ++i) cloud->push_back(pcl::PointXYZI(xyzi[4 * i], xyzi[4 * i + 1], xyzi[4 * i + 2], xyzi[4 * i + 3]));
  pcaAnalysist pca;
  std::vector<pcaAnalysist::pcaFeature> feats;
  if (!pca.calculatePCAofPointCloud(cloud, (float)std::atof(argv[2]), feats)) return 1;
  if (feats.size() != n) return 3;
  FILE* o = std::fopen(argv[3], "wb");
  std::fwrite(feats.data(), sizeof(pcaAnalysist::pcaFeature), feats.size(), o);
  std::fclose(o);
  if (argc >= 6) {
    std::vector<char> sb = slurp(argv[4]);
    const char* p = sb.data();
    int64_t n_seg = *reinterpret_cast<const int64_t*>(p);
    p += 8;
    std::vector<pcaAnalysist::pcaFeature> out((size_t)n_seg);
    for (int64_t s = 0; s < n_seg; ++s) {
      const float* c = reinterpret_cast<const float*>(p);
      p += 12;
      int64_t m = *reinterpret_cast<const int64_t*>(p);
      p += 8;
      std::vector<int> idx(reinterpret_cast<const int32_t*>(p), reinterpret_cast<const int32_t*>(p) + m);
      p += 4 * m;
      out[(size_t)s].ptID = -7;  // must survive: this overload never writes ptID
      if (!pca.calculatePCAofPoint(cloud, pcl::PointXYZI(c[0], c[1], c[2]), idx, out[(size_t)s])) return 4;
    }
    // the (cloud, int, idx) overload on the first segment's list, as calculatePCAofPointCloud would call it
    FILE* so = std::fopen(argv[5], "wb");
    std::fwrite(out.data(), sizeof(pcaAnalysist::pcaFeature), out.size(), so);
    std::fclose(so);
  }
  if (!pca.getMiniBoundingBox(cloud)) return 5;
  std::printf("dropin ok: %zu records\n", feats.size());
  return 0;
}
