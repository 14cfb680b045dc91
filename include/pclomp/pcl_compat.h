// pcl_compat.h -- the few PCL / Eigen types include/pclomp/ndt_omp.h needs, for builds WITHOUT PCL.
//
// With PCL installed (the xchu_slam workspace) ndt_omp.h includes the real <pcl/point_cloud.h>,
// <pcl/point_types.h> and <Eigen/Core> and this file is not used.  This container and the GPU box
// have neither, so the shim is exercised against these stand-ins: same member names
// (points, size(), push_back, width, height, is_dense, header, Ptr / ConstPtr) and the same memory
// layout for the point types (pcl::PointXYZ 16 B, pcl::PointXYZI 32 B, x y z first) and for
// Eigen::Matrix4f (16 floats, column-major).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace boost { using std::shared_ptr; }   // the reference holds clouds and the NDT by boost::shared_ptr

namespace Eigen {
struct Matrix4f {
  float m[16];   // column-major, like Eigen
  Matrix4f() { std::memset(m, 0, sizeof(m)); }
  static Matrix4f Identity() { Matrix4f r; r.setIdentity(); return r; }
  void setIdentity() { std::memset(m, 0, sizeof(m)); m[0] = m[5] = m[10] = m[15] = 1.f; }
  float& operator()(int r, int c) { return m[c * 4 + r]; }
  float operator()(int r, int c) const { return m[c * 4 + r]; }
  float* data() { return m; }
  const float* data() const { return m; }
  Matrix4f operator*(const Matrix4f& o) const {
    Matrix4f r;
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) {
        float s = 0.f;
        for (int k = 0; k < 4; ++k) s += (*this)(i, k) * o(k, j);
        r(i, j) = s;
      }
    return r;
  }
  bool operator==(const Matrix4f& o) const { return std::memcmp(m, o.m, sizeof(m)) == 0; }
  bool operator!=(const Matrix4f& o) const { return !(*this == o); }
};
}  // namespace Eigen

namespace pcl {
struct alignas(16) PointXYZ {
  float x = 0, y = 0, z = 0, pad_ = 1.f;
};
struct alignas(16) PointXYZI {
  float x = 0, y = 0, z = 0, pad_ = 1.f;
  float intensity = 0, pad2_[3] = {0, 0, 0};
};
static_assert(sizeof(PointXYZ) == 16 && sizeof(PointXYZI) == 32, "PCL point layouts");

struct PCLHeader {
  std::uint32_t seq = 0;
  std::uint64_t stamp = 0;
  std::string frame_id;
};

template <typename PointT>
class PointCloud {
 public:
  typedef boost::shared_ptr<PointCloud<PointT>> Ptr;
  typedef boost::shared_ptr<const PointCloud<PointT>> ConstPtr;
  PCLHeader header;
  std::vector<PointT> points;
  std::uint32_t width = 0, height = 0;
  bool is_dense = true;
  std::size_t size() const { return points.size(); }
  bool empty() const { return points.empty(); }
  void clear() { points.clear(); width = height = 0; }
  void resize(std::size_t n) { points.resize(n); width = (std::uint32_t)n; height = 1; }
  void push_back(const PointT& p) { points.push_back(p); width = (std::uint32_t)points.size(); height = 1; }
  PointT& operator[](std::size_t i) { return points[i]; }
  const PointT& operator[](std::size_t i) const { return points[i]; }
  PointCloud& operator+=(const PointCloud& o) {   // `localmap += *transformed_scan_ptr`, odom_node.cpp:336
    points.insert(points.end(), o.points.begin(), o.points.end());
    width = (std::uint32_t)points.size();
    height = 1;
    return *this;
  }
};
}  // namespace pcl
