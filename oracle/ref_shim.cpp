// ref_shim.cpp -- C entry points around the reference's OWN NDT sources, compiled unmodified from
// /root/reference/xchu_mapping/include/pclomp/{ndt_omp.h, ndt_omp_impl.hpp,
// voxel_grid_covariance_omp.h, voxel_grid_covariance_omp_impl.hpp} into oracle/_ref/libndt_ref.so.
//
// TEST INFRASTRUCTURE ONLY (the checker's checker): tests/ use it to pin oracle/ndt_oracle.cpp against
// the reference's code.  The reference's third-party dependencies (PCL, Eigen, FLANN, Boost) are not
// in this image; they are replaced by the stand-in headers of oracle/ref_stubs/, whose arithmetic is
// the shared restatement oracle/thirdparty_math.h.  So: NDT code = the
// reference's, linear-algebra / PCL primitives = restated.  Nothing here is shipped or timed as product.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <unordered_map>
#include <vector>

#include <Eigen/Dense>
#include <pcl/filters/voxel_grid.h>
#include <pcl/kdtree/kdtree_flann.h>
#include <pcl/point_types.h>
#include <pcl/registration/registration.h>

// the C entry points below read protected state (target_cells_, leaves_, min_b_, ...) and call protected members
#define protected public
#define private public
#include <pclomp/ndt_omp.h>
#include <pclomp/ndt_omp_impl.hpp>
#include <pclomp/voxel_grid_covariance_omp_impl.hpp>
#undef protected
#undef private

namespace {

typedef pcl::PointXYZI P;
typedef pcl::PointCloud<P> Cloud;
typedef pclomp::NormalDistributionsTransform<P, P> RefNdt;
typedef pclomp::VoxelGridCovariance<P> RefGrid;

struct Ref {
  RefNdt ndt;
  Cloud::Ptr target{new Cloud()}, source{new Cloud()};
  Cloud output;
  std::unordered_map<const RefGrid::Leaf*, size_t> leaf_key;
  void index_leaves() {
    leaf_key.clear();
    for (auto& kv : ndt.target_cells_.leaves_) leaf_key[&kv.second] = kv.first;
  }
  void key_to_ijk(size_t key, int ijk[3]) const {
    const RefGrid& g = ndt.target_cells_;
    const int dx = g.div_b_[0], dy = g.div_b_[1];
    const int idx = (int)key;
    ijk[0] = idx % dx + g.min_b_[0];
    ijk[1] = (idx / dx) % dy + g.min_b_[1];
    ijk[2] = idx / (dx * dy) + g.min_b_[2];
  }
  // NDT_I:80-87 (the reference recomputes these only inside computeTransformation)
  void gauss() {
    const double c1 = 10 * (1 - ndt.outlier_ratio_);
    const double c2 = ndt.outlier_ratio_ / pow(ndt.resolution_, 3);
    ndt.gauss_d3_ = -log(c2);
    ndt.gauss_d1_ = -log(c1 + c2) - ndt.gauss_d3_;
    ndt.gauss_d2_ = -2 * log((-log(c1 * exp(-0.5) + c2) - ndt.gauss_d3_) / ndt.gauss_d1_);
  }
};

void load(Cloud& c, const float* pts, size_t n, size_t stride_floats, bool append) {
  if (!append) c.points.clear();
  for (size_t i = 0; i < n; ++i) {
    P p;
    p.x = pts[i * stride_floats]; p.y = pts[i * stride_floats + 1]; p.z = pts[i * stride_floats + 2];
    c.points.push_back(p);
  }
  c.width = (uint32_t)c.points.size();
  c.height = 1;
  c.is_dense = true;
}
Eigen::Matrix4f mat_from_colmajor(const float* T) {
  Eigen::Matrix4f m;
  for (int c = 0; c < 4; ++c)
    for (int r = 0; r < 4; ++r) m(r, c) = T[c * 4 + r];
  return m;
}

}  // namespace

extern "C" {

void* ref_create() { return new Ref(); }
void ref_destroy(void* h) { delete (Ref*)h; }

void ref_set_params(void* h, float resolution, double step_size, double trans_eps, int max_iter, double outlier_ratio,
                    int search_method, int num_threads) {
  Ref* r = (Ref*)h;
  r->ndt.setNumThreads(num_threads > 0 ? num_threads : 1);
  r->ndt.setTransformationEpsilon(trans_eps);
  r->ndt.setStepSize(step_size);
  r->ndt.setResolution(resolution);
  r->ndt.setMaximumIterations(max_iter);
  r->ndt.setOulierRatio(outlier_ratio);
  r->ndt.setNeighborhoodSearchMethod((pclomp::NeighborSearchMethod)search_method);
  // setResolution re-inits only `if (input_)` (NDT_H:131-136, the source pointer); make a changed resolution
  // take effect on an existing target whatever the call order, as the oracle's set_params does
  if (r->target->points.size() && r->ndt.target_ && r->ndt.target_cells_.leaf_size_[0] != resolution) {
    r->ndt.init();
    r->index_leaves();
  }
}
void ref_set_voxel_params(void* h, int min_points, double eig_mult) {
  Ref* r = (Ref*)h;
  r->ndt.target_cells_.setMinPointPerVoxel(min_points);
  r->ndt.target_cells_.setCovEigValueInflationRatio(eig_mult);
}
void ref_set_target(void* h, const float* pts, size_t n, size_t stride_floats) {
  Ref* r = (Ref*)h;
  r->target.reset(new Cloud());
  load(*r->target, pts, n, stride_floats, false);
  r->ndt.setInputTarget(r->target);
  r->index_leaves();
}
void ref_set_source(void* h, const float* pts, size_t n, size_t stride_floats) {
  Ref* r = (Ref*)h;
  r->source.reset(new Cloud());
  load(*r->source, pts, n, stride_floats, false);
  r->ndt.setInputSource(r->source);
}
void ref_grid_bounds(void* h, int* min_b, int* max_b) {
  Ref* r = (Ref*)h;
  for (int a = 0; a < 3; ++a) { min_b[a] = r->ndt.target_cells_.min_b_[a]; max_b[a] = r->ndt.target_cells_.max_b_[a]; }
}
size_t ref_num_leaves(void* h) { return ((Ref*)h)->ndt.target_cells_.leaves_.size(); }
// all leaves in ascending map-key order; any pointer may be null
void ref_dump_leaves(void* h, int* ijk, int* nr_points, double* mean, double* cov, double* icov, float* centroid, double* evals) {
  Ref* r = (Ref*)h;
  size_t i = 0;
  for (auto& kv : r->ndt.target_cells_.leaves_) {
    const RefGrid::Leaf& l = kv.second;
    if (ijk) r->key_to_ijk(kv.first, ijk + i * 3);
    if (nr_points) nr_points[i] = l.nr_points;
    if (mean) for (int a = 0; a < 3; ++a) mean[i * 3 + a] = l.mean_[a];
    if (cov) for (int a = 0; a < 9; ++a) cov[i * 9 + a] = l.cov_(a / 3, a % 3);
    if (icov) for (int a = 0; a < 9; ++a) icov[i * 9 + a] = l.icov_(a / 3, a % 3);
    if (centroid) for (int a = 0; a < 3; ++a) centroid[i * 3 + a] = l.centroid.size() >= 3 ? l.centroid[a] : 0.f;
    if (evals) for (int a = 0; a < 3; ++a) evals[i * 3 + a] = l.evals_[a];
    ++i;
  }
}
size_t ref_num_centroids(void* h) {
  Ref* r = (Ref*)h;
  return r->ndt.target_cells_.voxel_centroids_ ? r->ndt.target_cells_.voxel_centroids_->points.size() : 0;
}
// neighbour sets of n already-transformed points, in the order the reference returns them
void ref_neighbors(void* h, const float* pts, size_t n, int method, int* out_cnt, int* out_ijk) {
  Ref* r = (Ref*)h;
  std::vector<RefGrid::LeafConstPtr> nb;
  std::vector<float> d;
  for (size_t i = 0; i < n; ++i) {
    P q;
    q.x = pts[i * 3]; q.y = pts[i * 3 + 1]; q.z = pts[i * 3 + 2];
    switch (method) {
      case pclomp::KDTREE: r->ndt.target_cells_.radiusSearch(q, r->ndt.resolution_, nb, d); break;
      case pclomp::DIRECT26: r->ndt.target_cells_.getNeighborhoodAtPoint(q, nb); break;
      case pclomp::DIRECT1: r->ndt.target_cells_.getNeighborhoodAtPoint1(q, nb); break;
      default: r->ndt.target_cells_.getNeighborhoodAtPoint7(q, nb); break;
    }
    out_cnt[i] = (int)nb.size();
    for (size_t k = 0; k < nb.size() && k < 27; ++k) r->key_to_ijk(r->leaf_key.at(nb[k]), out_ijk + (i * 27 + k) * 3);
  }
}
void ref_gauss(void* h, double* d1, double* d2, double* d3) {
  Ref* r = (Ref*)h;
  r->gauss();
  *d1 = r->ndt.gauss_d1_; *d2 = r->ndt.gauss_d2_; *d3 = r->ndt.gauss_d3_;
}
// one computeDerivatives (NDT_I:175-283) of the source transformed by T (column-major) at pose vector p
double ref_eval_derivatives(void* h, const float* T, const double* p, int compute_hessian, double* g, double* H) {
  Ref* r = (Ref*)h;
  r->gauss();
  Cloud tc;
  pcl::transformPointCloud(*r->source, tc, mat_from_colmajor(T));
  Eigen::Matrix<double, 6, 1> pv, gv;
  Eigen::Matrix<double, 6, 6> Hm;
  for (int i = 0; i < 6; ++i) pv[i] = p[i];
  const double s = r->ndt.computeDerivatives(gv, Hm, tc, pv, compute_hessian != 0);
  for (int i = 0; i < 6; ++i) {
    g[i] = gv[i];
    for (int j = 0; j < 6; ++j) H[i * 6 + j] = Hm(i, j);
  }
  return s;
}
// computeHessian (NDT_I:551-606) after computeAngleDerivatives(p)
void ref_compute_hessian(void* h, const float* T, const double* p, double* H) {
  Ref* r = (Ref*)h;
  r->gauss();
  Cloud tc;
  pcl::transformPointCloud(*r->source, tc, mat_from_colmajor(T));
  Eigen::Matrix<double, 6, 1> pv;
  Eigen::Matrix<double, 6, 6> Hm;
  for (int i = 0; i < 6; ++i) pv[i] = p[i];
  r->ndt.computeAngleDerivatives(pv);
  r->ndt.computeHessian(Hm, tc, pv);
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) H[i * 6 + j] = Hm(i, j);
}
double ref_calculate_score(void* h, const float* T) {
  Ref* r = (Ref*)h;
  r->gauss();
  Cloud tc;
  pcl::transformPointCloud(*r->source, tc, mat_from_colmajor(T));
  return r->ndt.calculateScore(tc);
}
struct ref_result { float T[16]; int iterations; int converged; double score; double trans_probability; };
void ref_align(void* h, const float* guess, ref_result* out) {
  Ref* r = (Ref*)h;
  r->ndt.align(r->output, mat_from_colmajor(guess));
  const Eigen::Matrix4f F = r->ndt.getFinalTransformation();
  for (int c = 0; c < 4; ++c)
    for (int rr = 0; rr < 4; ++rr) out->T[c * 4 + rr] = F(rr, c);
  out->iterations = r->ndt.getFinalNumIteration();
  out->converged = r->ndt.hasConverged() ? 1 : 0;
  out->trans_probability = r->ndt.getTransformationProbability();
  out->score = out->trans_probability * (double)r->source->points.size();
}
void ref_get_output(void* h, float* out) {
  Ref* r = (Ref*)h;
  for (size_t i = 0; i < r->output.points.size(); ++i) {
    out[i * 3] = r->output.points[i].x; out[i * 3 + 1] = r->output.points[i].y; out[i * 3 + 2] = r->output.points[i].z;
  }
}
double ref_fitness(void* h, double max_range) { return ((Ref*)h)->ndt.getFitnessScore(max_range); }
void ref_pose_to_matrix(const double* p, float* T) {
  Eigen::Matrix<double, 6, 1> x;
  for (int i = 0; i < 6; ++i) x[i] = p[i];
  Eigen::Matrix4f m;
  RefNdt::convertTransform(x, m);
  for (int c = 0; c < 4; ++c)
    for (int r = 0; r < 4; ++r) T[c * 4 + r] = m(r, c);
}
double ref_mt_trial(void* h, double a_l, double f_l, double g_l, double a_u, double f_u, double g_u, double a_t, double f_t, double g_t) {
  return ((Ref*)h)->ndt.trialValueSelectionMT(a_l, f_l, g_l, a_u, f_u, g_u, a_t, f_t, g_t);
}
int ref_mt_update(void* h, double* s /*a_l f_l g_l a_u f_u g_u*/, double a_t, double f_t, double g_t) {
  return ((Ref*)h)->ndt.updateIntervalMT(s[0], s[1], s[2], s[3], s[4], s[5], a_t, f_t, g_t) ? 1 : 0;
}
void ref_ctor_defaults(void* h, float* resolution, double* step, double* outlier, double* eps, int* max_iter, int* method, int* min_pts,
                       double* eig_mult) {
  Ref* r = (Ref*)h;
  *resolution = r->ndt.getResolution(); *step = r->ndt.getStepSize(); *outlier = r->ndt.getOulierRatio();
  *eps = r->ndt.getTransformationEpsilon(); *max_iter = r->ndt.getMaximumIterations(); *method = (int)r->ndt.search_method;
  *min_pts = r->ndt.target_cells_.getMinPointPerVoxel(); *eig_mult = r->ndt.target_cells_.getCovEigValueInflationRatio();
}

}  // extern "C"
