// Stand-in for <pcl/filters/voxel_grid.h>: the pcl::Filter / pcl::VoxelGrid members that
// pclomp::VoxelGridCovariance inherits (VGC_H:58-78).  TEST INFRASTRUCTURE ONLY -- see pcl/point_types.h.
#pragma once
#include <pcl/point_types.h>

namespace pcl {

template <class PointT>
class Filter {
 public:
  typedef pcl::PointCloud<PointT> PointCloud;
  typedef typename PointCloud::Ptr PointCloudPtr;
  typedef typename PointCloud::ConstPtr PointCloudConstPtr;
  virtual ~Filter() {}
  virtual void setInputCloud(const PointCloudConstPtr& cloud) { input_ = cloud; }
  PointCloudConstPtr getInputCloud() const { return input_; }

 protected:
  PointCloudConstPtr input_;
  IndicesPtr indices_;
  std::string filter_name_;
  const std::string& getClassName() const { return filter_name_; }
};

template <class PointT>
class VoxelGrid : public Filter<PointT> {
 public:
  VoxelGrid()
      : downsample_all_data_(true), save_leaf_layout_(false), filter_limit_min_(-FLT_MAX), filter_limit_max_(FLT_MAX),
        filter_limit_negative_(false) {
    this->filter_name_ = "VoxelGrid";
  }
  // pcl::VoxelGrid::setLeafSize: leaf (x, y, z, 1), inverse_leaf_size_ = 1.0f / leaf_size_ (float, per component)
  void setLeafSize(float lx, float ly, float lz) {
    leaf_size_[0] = lx; leaf_size_[1] = ly; leaf_size_[2] = lz;
    if (leaf_size_[3] == 0) leaf_size_[3] = 1;
    for (int a = 0; a < 4; ++a) inverse_leaf_size_[a] = 1.0f / leaf_size_[a];
  }
  Eigen::Vector3f getLeafSize() const { return Eigen::Vector3f(leaf_size_[0], leaf_size_[1], leaf_size_[2]); }

 protected:
  Eigen::Vector4f leaf_size_;
  Eigen::Array4f inverse_leaf_size_;
  bool downsample_all_data_, save_leaf_layout_;
  std::vector<int> leaf_layout_;
  Eigen::Vector4i min_b_, max_b_, div_b_, divb_mul_;
  std::string filter_field_name_;
  double filter_limit_min_, filter_limit_max_;
  bool filter_limit_negative_;
};

}  // namespace pcl
