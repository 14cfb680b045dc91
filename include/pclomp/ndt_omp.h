// include/pclomp/ndt_omp.h -- drop-in for xchu_slam's xchu_mapping/include/pclomp/ndt_omp.h.
//
// Same namespace, class template, enum, ::Ptr and public members as the reference
// (xchu_mapping/include/pclomp/ndt_omp.h:52-57 enum, :70-71 class, :98-99 Ptr, :110-204 setters /
// getters, plus the pcl::Registration members odom_node uses: setInputSource, setInputTarget,
// setTransformationEpsilon, setMaximumIterations, align, getFinalTransformation, hasConverged,
// getFitnessScore -- src/odom_node.cpp:70-83, 278-283, 349), so that odom_node.cpp's matcher
// selection and localmap refresh compile unchanged.  Everything underneath is the C ABI of
// libndt_b200.so (include/ndt_b200/ndt_c_api.h): CUDA only, no CPU fallback -- construction throws
// std::runtime_error when no device is usable, and every failing ABI call throws with
// ndtb_last_error() (the reference itself never fails: it has no device to lose).
//
// Builds with PCL + Eigen when they are installed (the xchu_slam workspace); otherwise against the
// stand-ins of pcl_compat.h (this repository's tests).
#pragma once

#if defined(__has_include)
#if __has_include(<pcl/point_cloud.h>) && __has_include(<Eigen/Core>) && !defined(NDT_B200_NO_PCL)
#define NDT_B200_WITH_PCL 1
#endif
#endif

#ifdef NDT_B200_WITH_PCL
#include <Eigen/Core>
#if __has_include(<Eigen/Geometry>)
#include <Eigen/Geometry>
#endif
#if __has_include(<boost/shared_ptr.hpp>)
#include <boost/shared_ptr.hpp>
#endif
#include <pcl/point_cloud.h>
#include <pcl/point_types.h>
#include <pcl/registration/registration.h>
#else
#include "pcl_compat.h"
#endif

#include <cfloat>
#include <limits>
#include <stdexcept>
#include <string>
#include <vector>

#include "../ndt_b200/ndt_c_api.h"

namespace pclomp {

enum NeighborSearchMethod { KDTREE, DIRECT26, DIRECT7, DIRECT1 };   // ndt_omp.h:52-57

// The base class: pcl::Registration itself when PCL is there (ndt_omp.h:70-71 -- a pclomp NDT can then be held through a
// pcl::Registration<PointSource, PointTarget>::Ptr like any other PCL matcher, and align() is the base's: copy the source into
// `output`, reset the state, call the virtual computeTransformation()); without PCL a stand-in with the same members and the
// same align() protocol, so that the class body below is one text for both builds.
#ifdef NDT_B200_WITH_PCL
template <typename PointSource, typename PointTarget>
using RegistrationBase = pcl::Registration<PointSource, PointTarget>;
#else
template <typename PointSource, typename PointTarget>
class RegistrationBase {
 public:
  typedef Eigen::Matrix4f Matrix4;
  typedef pcl::PointCloud<PointSource> PointCloudSource;
  typedef typename PointCloudSource::ConstPtr PointCloudSourceConstPtr;
  typedef pcl::PointCloud<PointTarget> PointCloudTarget;
  typedef typename PointCloudTarget::ConstPtr PointCloudTargetConstPtr;
  RegistrationBase() { final_transformation_.setIdentity(); }
  virtual ~RegistrationBase() {}
  virtual void setInputSource(const PointCloudSourceConstPtr& cloud) { input_ = cloud; }
  virtual void setInputTarget(const PointCloudTargetConstPtr& cloud) { target_ = cloud; }
  void setMaximumIterations(int n) { max_iterations_ = n; }
  int getMaximumIterations() const { return max_iterations_; }
  void setTransformationEpsilon(double e) { transformation_epsilon_ = e; }
  double getTransformationEpsilon() const { return transformation_epsilon_; }
  Matrix4 getFinalTransformation() const { return final_transformation_; }
  bool hasConverged() const { return converged_; }
  void align(PointCloudSource& output) { align(output, Matrix4::Identity()); }
  void align(PointCloudSource& output, const Matrix4& guess) {   // pcl::Registration::align
    if (!input_) return;
    if (&output != input_.get()) {
      output.points = input_->points;
      output.header = input_->header;
      output.width = (decltype(output.width))output.points.size();
      output.height = 1;
      output.is_dense = input_->is_dense;
    }
    converged_ = false;
    final_transformation_.setIdentity();
    computeTransformation(output, guess);
  }

 protected:
  virtual void computeTransformation(PointCloudSource& output, const Matrix4& guess) = 0;
  std::string reg_name_;
  PointCloudSourceConstPtr input_;
  PointCloudTargetConstPtr target_;
  int nr_iterations_ = 0, max_iterations_ = 10;     // pcl::Registration() defaults
  double transformation_epsilon_ = 0.0;
  Matrix4 final_transformation_;
  bool converged_ = false;
};
#endif

template <typename PointSource, typename PointTarget>
class NormalDistributionsTransform : public RegistrationBase<PointSource, PointTarget> {
 protected:
  typedef RegistrationBase<PointSource, PointTarget> Base;
  using Base::reg_name_;            // ndt_omp.h:237-249
  using Base::input_;
  using Base::target_;
  using Base::nr_iterations_;
  using Base::max_iterations_;
  using Base::transformation_epsilon_;
  using Base::final_transformation_;
  using Base::converged_;

 public:
  typedef pcl::PointCloud<PointSource> PointCloudSource;
  typedef typename PointCloudSource::Ptr PointCloudSourcePtr;
  typedef typename PointCloudSource::ConstPtr PointCloudSourceConstPtr;
  typedef pcl::PointCloud<PointTarget> PointCloudTarget;
  typedef typename PointCloudTarget::Ptr PointCloudTargetPtr;
  typedef typename PointCloudTarget::ConstPtr PointCloudTargetConstPtr;
  typedef boost::shared_ptr<NormalDistributionsTransform<PointSource, PointTarget>> Ptr;
  typedef boost::shared_ptr<const NormalDistributionsTransform<PointSource, PointTarget>> ConstPtr;
  typedef Eigen::Matrix4f Matrix4;

  // ndt_omp_impl.hpp:46-69: resolution 1.0, step 0.1, outlier 0.55, eps 0.1, 35 iterations, DIRECT7
  explicit NormalDistributionsTransform(int device = 0) : search_method(DIRECT7) {
    if (ndtb_create(device, &h_) != 0) throw std::runtime_error(std::string("pclomp(B200): ") + ndtb_last_error(nullptr));
    ndtb_get_params(h_, &p_);
    reg_name_ = "NormalDistributionsTransform";      // ndt_omp_impl.hpp:56
    transformation_epsilon_ = p_.trans_eps;          // ndt_omp_impl.hpp:67-68: 0.1, 35
    max_iterations_ = p_.max_iter;
    final_transformation_.setIdentity();
  }
  virtual ~NormalDistributionsTransform() { ndtb_destroy(h_); }
  NormalDistributionsTransform(const NormalDistributionsTransform&) = delete;
  NormalDistributionsTransform& operator=(const NormalDistributionsTransform&) = delete;

  void setNumThreads(int) {}   // ndt_omp.h:110-112: OpenMP threads have no meaning on the device

  // ndt_omp.h:117-122: stores the target and rebuilds the whole voxel grid (init()).
  // Incremental refresh: when `cloud` extends the previous target (odom_node only ever appends
  // between window resets, src/odom_node.cpp:336, 349) only the appended tail is merged.
  inline void setInputTarget(const PointCloudTargetConstPtr& cloud) override {
    Base::setInputTarget(cloud);
    push_search_method();
    int inc = 0;
    ck(ndtb_set_target_auto(h_, cloud->points.data(), cloud->points.size(), sizeof(PointTarget), &inc));
    last_target_incremental_ = inc != 0;
    target_set_ = true;
  }
  // cpu::NormalDistributionsTransform::updateVoxelGrid (include/ndt_cpu/NormalDistributionsTransform.h:39)
  inline void updateVoxelGrid(const PointCloudTargetConstPtr& new_points) {
    ck(ndtb_update_target(h_, new_points->points.data(), new_points->points.size(), sizeof(PointTarget)));
  }
  inline void setInputSource(const PointCloudSourceConstPtr& cloud) override {
    Base::setInputSource(cloud);
    ck(ndtb_set_source(h_, cloud->points.data(), cloud->points.size(), sizeof(PointSource)));
  }

  inline void setResolution(float resolution) {           // ndt_omp.h:127-137
    if (p_.resolution != resolution) { p_.resolution = resolution; push(); }
  }
  inline float getResolution() const { return p_.resolution; }
  inline double getStepSize() const { return p_.step_size; }
  inline void setStepSize(double step_size) { p_.step_size = step_size; push(); }      // ndt_omp.h:160-164
  inline double getOulierRatio() const { return p_.outlier_ratio; }                      // (sic) ndt_omp.h:169-173
  inline void setOulierRatio(double outlier_ratio) { p_.outlier_ratio = outlier_ratio; push(); }
  inline void setNeighborhoodSearchMethod(NeighborSearchMethod method) { search_method = method; push_search_method(); }
  // setTransformationEpsilon / setMaximumIterations / their getters, align(), getFinalTransformation() and hasConverged() are
  // the (non-virtual) base's: they act on transformation_epsilon_, max_iterations_, final_transformation_, converged_, which
  // computeTransformation() below reads and writes -- so they work through a pcl::Registration pointer too.
  inline double getTransformationProbability() const { return result_.trans_probability; }   // ndt_omp.h:191-195
  inline int getFinalNumIteration() const { return nr_iterations_; }                         // ndt_omp.h:200-204

  // pcl::Registration::align, restated here so that a call through the pclomp type (odom_node's, src/odom_node.cpp:278) skips
  // the base's initCompute(): PCL's builds a host kd-tree over the whole target after every setInputTarget, which neither this
  // class nor the reference's DIRECT7 path ever queries.  Through a pcl::Registration pointer the base's align() runs instead
  // and reaches the same computeTransformation().
  inline void align(PointCloudSource& output) { align(output, Matrix4::Identity()); }
  inline void align(PointCloudSource& output, const Matrix4& guess) {
    if (!input_ || !target_set_) throw std::runtime_error("pclomp(B200): align() needs setInputSource and setInputTarget");
    if (&output != input_.get()) {
      output.points = input_->points;   // the base copies the source fields first
      output.header = input_->header;
      output.width = (decltype(output.width))output.points.size();
      output.height = 1;
      output.is_dense = input_->is_dense;
    }
    computeTransformation(output, guess);
  }
  inline double getFitnessScore(double max_range = std::numeric_limits<double>::max()) {
    double out = 0;
    ck(ndtb_fitness(h_, max_range, nullptr, &out));
    return out;
  }
  // ndt_omp.h:231-233 / ndt_omp_impl.hpp:919-952: negative log likelihood of an ALREADY TRANSFORMED cloud (lower is better).
  // The cloud takes the source slot of the handle for the evaluation and the input source is put back afterwards -- two
  // uploads; a caller that scores the current source under a transform uses the overload below and uploads nothing.
  inline double calculateScore(const PointCloudSource& cloud) {
    double out = 0;
    Matrix4 I = Matrix4::Identity();
    ck(ndtb_set_source(h_, cloud.points.data(), cloud.points.size(), sizeof(PointSource)));
    const int rc = ndtb_calculate_score(h_, I.data(), &out);
    if (input_) ck(ndtb_set_source(h_, input_->points.data(), input_->points.size(), sizeof(PointSource)));
    ck(rc);
    return out;
  }
  inline double calculateScore(const Matrix4& cloud_transform) {       // not in the reference: cloud = T * source
    double out = 0;
    ck(ndtb_calculate_score(h_, cloud_transform.data(), &out));
    return out;
  }
  // ndt_omp.h:210-229: [x y z roll pitch yaw] -> Translation * AngleAxis(X) * AngleAxis(Y) * AngleAxis(Z), float
  template <typename Vector6>
  static void convertTransform(const Vector6& x, Matrix4& trans) {
    const double p[6] = {(double)x(0), (double)x(1), (double)x(2), (double)x(3), (double)x(4), (double)x(5)};
    float T[16];
    ndtb_pose_to_matrix(p, T);
    for (int i = 0; i < 16; ++i) trans.data()[i] = T[i];
  }
#ifdef NDT_B200_WITH_PCL
  static void convertTransform(const Eigen::Matrix<double, 6, 1>& x, Eigen::Affine3f& trans) {
    Matrix4 m;
    convertTransform(x, m);
    trans.matrix() = m;
  }
#endif

  // ---- not in the reference ---------------------------------------------------------------
  bool lastTargetUpdateWasIncremental() const { return last_target_incremental_; }
  // align() without the copy of the moved source back to the host (odom_node never reads output_cloud, ODOM:262-279)
  inline void alignNoOutput(const Matrix4& guess) { run(guess); }
  // filter_node's front end on the device (ndt_c_api.h "filter_node" section; reference src/filter_node.cpp:236-262): the
  // filtered scan becomes the source, as if setInputSource had been called with it; returns its size
  inline size_t filterScan(const PointCloudSource& raw, double r_min = 1.0, double r_max = 60.0, float leaf = 0.5f, int mean_k = 30,
                           double stddev_mul = 1.0) {
    ndtb_filter_params fp;
    fp.r_min = r_min; fp.r_max = r_max; fp.leaf = leaf; fp.mean_k = mean_k; fp.stddev_mul = stddev_mul;
    size_t c[3] = {0, 0, 0};
    ck(ndtb_filter_scan(h_, raw.points.data(), raw.points.size(), sizeof(PointSource), &fp, c));
    input_.reset(new PointCloudSource());          // a placeholder: the source lives on the device
    return c[2];
  }
  // odom_node's localmap kept in HBM (ndt_c_api.h "localmap" section; reference src/odom_node.cpp:218-233, 289-290, 329-356)
  inline void localmapClear() { ck(ndtb_localmap_clear(h_)); }
  // localmap (+ tmp_map) += [voxel-downsampled at `leaf` if > 0] T * current source; returns the points appended
  inline size_t localmapAppendSource(const Matrix4& T, float leaf, bool also_tmp) {
    size_t n = 0;
    ck(ndtb_localmap_append_source(h_, T.data(), leaf, also_tmp ? 1 : 0, &n));
    return n;
  }
  inline void localmapSnapshot() { ck(ndtb_localmap_snapshot(h_)); }                 // pc_target_.reset(new Cloud(localmap))
  inline void localmapSetTarget(bool force_full = false) {                          // setInputTarget(pc_target_)
    push_search_method();
    int inc = 0;
    ck(ndtb_localmap_set_target(h_, force_full ? 1 : 0, &inc));
    last_target_incremental_ = inc != 0;
    target_set_ = true;
  }
  inline void localmapRoll() { ck(ndtb_localmap_roll(h_)); }                         // localmap = tmp_map; tmp_map.clear()
  // the keyframe-count window kept in HBM (ndt_c_api.h "sliding window" section; reference src/odom_node.cpp:737-802)
  inline int windowAddSource(const Matrix4& T, float leaf) {                         // tf_cloud = downSizeFilterLocalmap(T * source)
    int id = -1;
    ck(ndtb_window_add_source(h_, T.data(), leaf, &id, nullptr));
    return id;
  }
  inline void windowDrop(int id) { ck(ndtb_window_drop(h_, id)); }
  template <typename IdContainer>
  inline size_t windowSetTarget(const IdContainer& ids, float leaf) {                // pc_target_ = VoxelGrid(concat); setInputTarget
    std::vector<int> v(ids.begin(), ids.end());
    push_search_method();
    size_t n = 0;
    ck(ndtb_window_set_target(h_, v.data(), (int)v.size(), leaf, &n));
    last_target_incremental_ = false;
    target_set_ = true;
    return n;
  }
  inline size_t localmapSize(int which = 0) { size_t s[3]; ck(ndtb_localmap_sizes(h_, s)); return s[which]; }
  void setStrictMath(bool on) { p_.strict_math = on ? 1 : 0; push(); }
  ndtb_handle* handle() { return h_; }

  NeighborSearchMethod search_method;   // public member in the reference too (ndt_omp.h:494)

 protected:
  void ck(int rc) const {
    if (rc != 0) throw std::runtime_error(std::string("pclomp(B200): ") + ndtb_last_error(h_));
  }
  void push() { ck(ndtb_set_params(h_, &p_)); }
  // ndt_omp_impl.hpp:73-170.  `output` arrives as a copy of the source (the base's align()); it leaves moved by the result.
  void computeTransformation(PointCloudSource& output, const Matrix4& guess) override {
    run(guess);
    if (!output.points.empty() && output.points.size() == input_->points.size())
      ck(ndtb_get_output(h_, output.points.data(), sizeof(PointSource)));
  }
  void run(const Matrix4& guess) {
    if (!input_ || !target_set_) throw std::runtime_error("pclomp(B200): align() needs setInputSource and setInputTarget");
    if (p_.trans_eps != transformation_epsilon_ || p_.max_iter != max_iterations_) {   // set through the base's setters
      p_.trans_eps = transformation_epsilon_;
      p_.max_iter = max_iterations_;
      push();
    }
    push_search_method();
    ck(ndtb_align(h_, guess.data(), &result_));
    for (int i = 0; i < 16; ++i) final_transformation_.data()[i] = result_.T[i];
    converged_ = result_.converged != 0;
    nr_iterations_ = result_.iterations;
  }
  void push_search_method() {   // `search_method` may have been assigned directly
    if (p_.search_method != (int)search_method) { p_.search_method = (int)search_method; push(); }
  }

  ndtb_handle* h_ = nullptr;
  ndtb_params p_;
  ndtb_result result_ = ndtb_result();
  bool last_target_incremental_ = false;
  bool target_set_ = false;
};

}  // namespace pclomp
