// Stand-in for the parts of PCL 1.7 (point types, PointCloud, console macros, field helpers, boost
// shims) that the reference's NDT sources name.  TEST INFRASTRUCTURE ONLY: PCL/Boost are not in this
// image; this lets /root/reference/xchu_mapping/include/pclomp/* compile unmodified into oracle/_ref.
// Semantics follow SURVEY.md appendix B (disclosed decisions, shared with oracle/ndt_oracle.cpp).
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <iostream>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include <Eigen/Dense>

#define PCL_EXPORTS
#define PCL_WARN(...) fprintf(stderr, __VA_ARGS__)
#define PCL_ERROR(...) fprintf(stderr, __VA_ARGS__)
#define PCL_DEBUG(...) ((void)0)
#define pcl_isfinite(x) std::isfinite(x)

namespace boost {
using std::shared_ptr;
using std::make_shared;
template <class Sig> using function = std::function<Sig>;
typedef std::mt19937 mt19937;
template <class T = double> using normal_distribution = std::normal_distribution<T>;
template <class Engine, class Dist>
struct variate_generator {
  typename std::remove_reference<Engine>::type& e;
  Dist d;
  variate_generator(typename std::remove_reference<Engine>::type& e_, const Dist& d_) : e(e_), d(d_) {}
  typename Dist::result_type operator()() { return d(e); }
};
namespace mpl {
template <class L> struct size { enum { value = L::value }; };
}  // namespace mpl
}  // namespace boost

namespace pcl {

struct PCLHeader { uint32_t seq = 0; uint64_t stamp = 0; std::string frame_id; };

struct PointXYZ {
  union { float data[4]; struct { float x, y, z; }; };
  PointXYZ() : x(0), y(0), z(0) { data[3] = 1.0f; }
  PointXYZ(float x_, float y_, float z_) : x(x_), y(y_), z(z_) { data[3] = 1.0f; }
};
struct alignas(16) PointXYZI {
  union { float data[4]; struct { float x, y, z; }; };
  union { struct { float intensity; }; float data_c[4]; };
  PointXYZI() : x(0), y(0), z(0) { data[3] = 1.0f; intensity = 0; data_c[1] = data_c[2] = data_c[3] = 0; }
};

struct PCLPointField { std::string name; uint32_t offset = 0; uint8_t datatype = 0; uint32_t count = 0; };
struct PointIndices {
  PCLHeader header;
  std::vector<int> indices;
  typedef boost::shared_ptr<PointIndices> Ptr;
  typedef boost::shared_ptr<const PointIndices> ConstPtr;
};
typedef boost::shared_ptr<std::vector<int>> IndicesPtr;
typedef boost::shared_ptr<const std::vector<int>> IndicesConstPtr;

namespace traits {
template <class PointT> struct fieldList { struct type { enum { value = 4 }; }; };
}  // namespace traits

template <class PointT>
class PointCloud {
 public:
  PCLHeader header;
  std::vector<PointT> points;
  uint32_t width = 0, height = 0;
  bool is_dense = true;
  typedef boost::shared_ptr<PointCloud<PointT>> Ptr;
  typedef boost::shared_ptr<const PointCloud<PointT>> ConstPtr;
  typedef PointT PointType;
  size_t size() const { return points.size(); }
  bool empty() const { return points.empty(); }
  void clear() { points.clear(); width = height = 0; }
  void push_back(const PointT& p) { points.push_back(p); width = (uint32_t)points.size(); height = 1; }
  PointT& back() { return points.back(); }
  const PointT& back() const { return points.back(); }
  PointT& operator[](size_t i) { return points[i]; }
  const PointT& operator[](size_t i) const { return points[i]; }
  PointCloud& operator+=(const PointCloud& o) {
    points.insert(points.end(), o.points.begin(), o.points.end());
    width = (uint32_t)points.size();
    height = 1;
    return *this;
  }
};

// neither PointXYZ nor PointXYZI carries an rgb/rgba field
template <class PointT>
int getFieldIndex(const PointCloud<PointT>&, const std::string&, std::vector<PCLPointField>& fields) {
  fields.clear();
  return -1;
}
// downsample_all_data_ is false on this path (VGC_H:214): these are compiled, never executed
template <class PointT> struct NdCopyPointEigenFunctor { NdCopyPointEigenFunctor(const PointT&, Eigen::VectorXf&) {} };
template <class PointT> struct NdCopyEigenPointFunctor { NdCopyEigenPointFunctor(const Eigen::VectorXf&, PointT&) {} };
template <class List, class F> void for_each_type(F) {}

// pcl::getMinMax3D(cloud, min, max): component-wise float min / max over all (finite) points
template <class PointT>
void getMinMax3D(const PointCloud<PointT>& cloud, Eigen::Vector4f& min_pt, Eigen::Vector4f& max_pt) {
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (size_t i = 0; i < cloud.points.size(); ++i) {
    const PointT& p = cloud.points[i];
    if (!cloud.is_dense && (!std::isfinite(p.x) || !std::isfinite(p.y) || !std::isfinite(p.z))) continue;
    mn[0] = p.x < mn[0] ? p.x : mn[0]; mn[1] = p.y < mn[1] ? p.y : mn[1]; mn[2] = p.z < mn[2] ? p.z : mn[2];
    mx[0] = p.x > mx[0] ? p.x : mx[0]; mx[1] = p.y > mx[1] ? p.y : mx[1]; mx[2] = p.z > mx[2] ? p.z : mx[2];
  }
  min_pt = Eigen::Vector4f(mn[0], mn[1], mn[2], 0.f);
  max_pt = Eigen::Vector4f(mx[0], mx[1], mx[2], 0.f);
}
// field-limited overload: the path never sets a filter field (VGC_I:68-69); present so the branch compiles
template <class PointT>
void getMinMax3D(const typename PointCloud<PointT>::ConstPtr& cloud, const std::string&, float, float, Eigen::Vector4f& min_pt,
                 Eigen::Vector4f& max_pt, bool = false) {
  getMinMax3D<PointT>(*cloud, min_pt, max_pt);
}

// pcl::transformPointCloud(in, out, Matrix4f): x' = ((m00 x + m01 y) + m02 z) + m03 in float; other fields copied
template <class PointT>
void transformPointCloud(const PointCloud<PointT>& in, PointCloud<PointT>& out, const Eigen::Matrix4f& T) {
  if (&in != &out) {
    out.header = in.header; out.width = in.width; out.height = in.height; out.is_dense = in.is_dense;
    out.points.resize(in.points.size());
  }
  const float m00 = T(0, 0), m01 = T(0, 1), m02 = T(0, 2), m03 = T(0, 3), m10 = T(1, 0), m11 = T(1, 1), m12 = T(1, 2), m13 = T(1, 3),
              m20 = T(2, 0), m21 = T(2, 1), m22 = T(2, 2), m23 = T(2, 3);
  for (size_t i = 0; i < in.points.size(); ++i) {
    PointT p = in.points[i];
    const float x = p.x, y = p.y, z = p.z;
    p.x = ((m00 * x + m01 * y) + m02 * z) + m03;
    p.y = ((m10 * x + m11 * y) + m12 * z) + m13;
    p.z = ((m20 * x + m21 * y) + m22 * z) + m23;
    out.points[i] = p;
  }
}

// pcl::getAllNeighborCellIndices(): the 13 cells of getHalfNeighborCellIndices() followed by their
// negations, i.e. {-1,0,1}^3 without (0,0,0)
inline Eigen::MatrixXi getAllNeighborCellIndices() {
  Eigen::MatrixXi half = Eigen::MatrixXi::Zero(3, 13);
  int idx = 0;
  for (int i = -1; i < 2; i++)            // 0 - 8
    for (int j = -1; j < 2; j++) { half(0, idx) = i; half(1, idx) = j; half(2, idx) = -1; idx++; }
  for (int i = -1; i < 2; i++) { half(0, idx) = i; half(1, idx) = -1; half(2, idx) = 0; idx++; }   // 9 - 11
  half(0, idx) = -1; half(1, idx) = 0; half(2, idx) = 0;                                           // 12
  Eigen::MatrixXi all = Eigen::MatrixXi::Zero(3, 26);
  for (int c = 0; c < 13; ++c)
    for (int r = 0; r < 3; ++r) { all(r, c) = half(r, c); all(r, 13 + c) = -half(r, c); }
  return all;
}

}  // namespace pcl
