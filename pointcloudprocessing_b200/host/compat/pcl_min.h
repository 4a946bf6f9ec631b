// compat/pcl_min.h — the few PCL / Eigen / Boost types the pcaAnalysist interface mentions, for
// machines without PCL (this image and the GPU box have none).  With real PCL installed, build with
// -DPCP_HAVE_PCL and this header is not used.  Layouts follow PCL 1.8: pcl::PointXYZI is 32 bytes,
// 16-byte aligned {x,y,z,1 | intensity,pad×3}; Eigen::Vector3f is three packed floats, so
// sizeof(pcaAnalysist::pcaFeature) == 68 exactly as in the reference build.
#ifndef PCP_COMPAT_PCL_MIN_H_
#define PCP_COMPAT_PCL_MIN_H_
#include <cstddef>
#include <memory>
#include <vector>

namespace Eigen {
struct Vector3f {
  float v[3];
  float& operator()(int r, int = 0) { return v[r]; }
  float operator()(int r, int = 0) const { return v[r]; }
  float& operator[](int r) { return v[r]; }
  float operator[](int r) const { return v[r]; }
};
// column-major dynamic float matrix: the slice of Eigen::MatrixXf that lidarOdometryClass's signatures use
struct MatrixXf {
  int r = 0, c = 0;
  std::vector<float> d;
  MatrixXf() {}
  MatrixXf(int rows_, int cols_) : r(rows_), c(cols_), d((std::size_t)rows_ * cols_, 0.f) {}
  void resize(int rows_, int cols_) {
    r = rows_;
    c = cols_;
    d.assign((std::size_t)rows_ * cols_, 0.f);
  }
  static MatrixXf Zero(int rows_, int cols_) { return MatrixXf(rows_, cols_); }
  int rows() const { return r; }
  int cols() const { return c; }
  float& operator()(int i, int j) { return d[(std::size_t)j * r + i]; }
  float operator()(int i, int j) const { return d[(std::size_t)j * r + i]; }
};
}  // namespace Eigen

namespace boost {
template <typename T>
using shared_ptr = std::shared_ptr<T>;
}

namespace pcl {
struct alignas(16) PointXYZI {
  float x, y, z, pad_ = 1.f;
  float intensity = 0.f, pad2_[3] = {0.f, 0.f, 0.f};
  PointXYZI() : x(0.f), y(0.f), z(0.f) {}
  PointXYZI(float x_, float y_, float z_, float i_ = 0.f) : x(x_), y(y_), z(z_), intensity(i_) {}
};
static_assert(sizeof(PointXYZI) == 32, "pcl::PointXYZI is 32 bytes");

template <typename PointT>
struct PointCloud {
  typedef boost::shared_ptr<PointCloud<PointT> > Ptr;
  std::vector<PointT> points;
  unsigned width = 0, height = 1;
  bool is_dense = true;
  void push_back(const PointT& p) {
    points.push_back(p);
    width = (unsigned)points.size();
  }
  std::size_t size() const { return points.size(); }
  void clear() {
    points.clear();
    width = 0;
  }
};
}  // namespace pcl
#endif
