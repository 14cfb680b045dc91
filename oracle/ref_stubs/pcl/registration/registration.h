// Stand-in for <pcl/registration/registration.h>: the pcl::Registration base-class plumbing that
// pclomp::NormalDistributionsTransform inherits (NDT_H:71, 232-249) and odom_node calls
// (ODOM:264-283).  TEST INFRASTRUCTURE ONLY -- see pcl/point_types.h.  Behaviour per SURVEY.md
// appendix B (PCL 1.7 registration.hpp as understood by the surveyor; disclosed, not verifiable offline).
#pragma once
#include <pcl/kdtree/kdtree_flann.h>
#include <pcl/point_types.h>

namespace pcl {

template <class PointSource, class PointTarget, class Scalar = float>
class Registration {
 public:
  typedef Eigen::Matrix<Scalar, 4, 4> Matrix4;
  typedef pcl::PointCloud<PointSource> PointCloudSource;
  typedef typename PointCloudSource::Ptr PointCloudSourcePtr;
  typedef typename PointCloudSource::ConstPtr PointCloudSourceConstPtr;
  typedef pcl::PointCloud<PointTarget> PointCloudTarget;
  typedef typename PointCloudTarget::Ptr PointCloudTargetPtr;
  typedef typename PointCloudTarget::ConstPtr PointCloudTargetConstPtr;
  typedef boost::function<void(const PointCloudSource&, const std::vector<int>&, const PointCloudTarget&, const std::vector<int>&)>
      UpdateVisualizerCallbackSignature;

  Registration()
      : nr_iterations_(0), max_iterations_(10), final_transformation_(Matrix4::Identity()), transformation_(Matrix4::Identity()),
        previous_transformation_(Matrix4::Identity()), transformation_epsilon_(0.0), corr_dist_threshold_(std::sqrt(DBL_MAX)),
        inlier_threshold_(0.05), converged_(false), target_cloud_updated_(true) {}
  virtual ~Registration() {}

  virtual void setInputSource(const PointCloudSourceConstPtr& cloud) { input_ = cloud; }
  virtual void setInputTarget(const PointCloudTargetConstPtr& cloud) { target_ = cloud; target_cloud_updated_ = true; }
  void setMaximumIterations(int n) { max_iterations_ = n; }
  int getMaximumIterations() const { return max_iterations_; }
  void setTransformationEpsilon(double e) { transformation_epsilon_ = e; }
  double getTransformationEpsilon() const { return transformation_epsilon_; }
  Matrix4 getFinalTransformation() const { return final_transformation_; }
  Matrix4 getLastIncrementalTransformation() const { return transformation_; }
  bool hasConverged() const { return converged_; }

  void align(PointCloudSource& output) { align(output, Matrix4::Identity()); }
  // output := copy of the source with w = 1; state reset; computeTransformation(output, guess)
  void align(PointCloudSource& output, const Matrix4& guess) {
    if (!input_ || !target_) return;
    output.header = input_->header;
    output.points = input_->points;
    output.width = (uint32_t)output.points.size();
    output.height = 1;
    output.is_dense = input_->is_dense;
    for (size_t i = 0; i < output.points.size(); ++i) output.points[i].data[3] = 1.0f;
    converged_ = false;
    final_transformation_ = transformation_ = previous_transformation_ = Matrix4::Identity();
    computeTransformation(output, guess);
  }
  // mean squared 1-NN distance (float squared distances, double mean) of the transformed source to the target
  double getFitnessScore(double max_range = std::numeric_limits<double>::max()) {
    if (target_cloud_updated_ || !tree_) {
      tree_.reset(new pcl::KdTreeFLANN<PointTarget>());
      tree_->setInputCloud(target_);
      target_cloud_updated_ = false;
    }
    PointCloudSource t;
    transformPointCloud(*input_, t, final_transformation_);
    double sum = 0.0;
    int nr = 0;
    std::vector<int> idx(1);
    std::vector<float> d2(1);
    for (size_t i = 0; i < t.points.size(); ++i) {
      PointTarget q;
      q.x = t.points[i].x; q.y = t.points[i].y; q.z = t.points[i].z;
      if (tree_->nearestKSearch(q, 1, idx, d2) < 1) continue;
      if (d2[0] <= max_range) { sum += d2[0]; nr++; }
    }
    return nr > 0 ? sum / nr : std::numeric_limits<double>::max();
  }

 protected:
  std::string reg_name_;
  PointCloudSourceConstPtr input_;
  IndicesPtr indices_;
  PointCloudTargetConstPtr target_;
  int nr_iterations_, max_iterations_;
  Matrix4 final_transformation_, transformation_, previous_transformation_;
  double transformation_epsilon_, corr_dist_threshold_, inlier_threshold_;
  bool converged_;
  bool target_cloud_updated_;
  boost::shared_ptr<pcl::KdTreeFLANN<PointTarget>> tree_;
  boost::function<void(const PointCloudSource&, const std::vector<int>&, const PointCloudTarget&, const std::vector<int>&)> update_visualizer_;

  const std::string& getClassName() const { return reg_name_; }
  virtual void computeTransformation(PointCloudSource& output, const Matrix4& guess) = 0;
};

}  // namespace pcl
