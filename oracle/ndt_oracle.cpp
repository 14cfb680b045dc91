// ndt_oracle.cpp -- CPU restatement of xchu_slam's ndt_omp scan-matching path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing in the product (xchu_slam_b200/, include/) may
// link, import or call this file.  It is the checker for tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs.
//
// PARITY PIN: the reference ships no tests, golden vectors or fixtures for this path (SURVEY.md
// section 8c), and its CMake build needs PCL, Eigen, FLANN, Boost and ROS, none of which are in
// this image.  What pins this file instead is the reference's OWN NDT code: oracle/Makefile target
// `ref` compiles /root/reference/xchu_mapping/include/pclomp/{ndt_omp.h, ndt_omp_impl.hpp,
// voxel_grid_covariance_omp.h, voxel_grid_covariance_omp_impl.hpp} unmodified, in place, against
// stand-in Eigen/PCL headers (oracle/ref_stubs/) into oracle/_ref/libndt_ref.so, and
// tests/test_reference_build.py requires this restatement to equal it BIT FOR BIT at every stage
// (voxel grid, neighbour sets, computeDerivatives, computeHessian, calculateScore, the Newton /
// More-Thuente loop, fitness).  What stays a disclosed restatement is the third-party arithmetic
// underneath both (oracle/thirdparty_math.h, shared by this file and the stand-in headers; list
// below) plus the Gauss-constant known answers of SURVEY.md appendix C.
//
// File:line citations use the shorthands
//   NDT_I = xchu_mapping/include/pclomp/ndt_omp_impl.hpp
//   NDT_H = xchu_mapping/include/pclomp/ndt_omp.h
//   VGC_I = xchu_mapping/include/pclomp/voxel_grid_covariance_omp_impl.hpp
//   VGC_H = xchu_mapping/include/pclomp/voxel_grid_covariance_omp.h
//
// Disclosed decisions for arithmetic that lives in un-vendored dependencies
// (PCL 1.7 / Eigen 3.3 / FLANN 1.8):
//   * every small fixed-size float product is evaluated coefficient-wise with the inner
//     index ascending and no FMA contraction (compile with -ffp-contract=off);
//   * sinf/cosf of Eigen::AngleAxis<float> := (float)sin((double)a), (float)cos((double)a);
//     the unqualified exp(float) of NDT_I:507 := (float)exp((double)a);
//   * pcl::transformPointCloud: x' = ((m00*x + m01*y) + m02*z) + m03 in float;
//   * Eigen::SelfAdjointEigenSolver<Matrix3d>: cyclic Jacobi on the lower triangle,
//     eigenvalues ascending; Matrix3d::inverse(): cofactor formula;
//   * Eigen::JacobiSVD::solve: one-sided Jacobi SVD pseudo-inverse solve, singular values
//     below 6*eps*sigma_max treated as zero;
//   * KdTreeFLANN::radiusSearch over voxel centroids := exact scan of the 27 surrounding
//     cells, squared float distance ((dx^2+dy^2)+dz^2) < float(r*r), ascending distance;
//   * Registration::getFitnessScore := exact 1-NN (grid ring search) over the raw target.
//
// Build: see oracle/Makefile (g++ -O2 -fopenmp -ffp-contract=off -shared -fPIC).

#include "thirdparty_math.h"
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <unordered_map>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// third-party (Eigen/PCL) arithmetic: shared with the stand-in headers of oracle/_ref
using namespace tpm;

static inline void transform_point(const float T[16], const float* x, float* y) {
  y[0] = ((T[0] * x[0] + T[4] * x[1]) + T[8] * x[2]) + T[12];
  y[1] = ((T[1] * x[0] + T[5] * x[1]) + T[9] * x[2]) + T[13];
  y[2] = ((T[2] * x[0] + T[6] * x[1]) + T[10] * x[2]) + T[14];
}

// ----------------------------------------------------------------------------------------
// VoxelGridCovariance restatement
// ----------------------------------------------------------------------------------------
struct Leaf {                   // VGC_H:92-187
  int nr_points = 0;
  double mean[3] = {0, 0, 0};
  float centroid[3] = {0, 0, 0};
  M3 cov{{{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}};  // Identity, VGC_H:101
  M3 icov{{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
  double evals[3] = {0, 0, 0};
  int n_raw = 0;                // points actually inserted (nr_points may become -1)
  int ijk_abs[3] = {0, 0, 0};   // absolute cell coordinate (min_b + ijk)
};

struct Grid {
  float leaf = 1.f, inv_leaf = 1.f;
  int min_b[3] = {0, 0, 0}, max_b[3] = {0, 0, 0}, div_b[3] = {0, 0, 0}, divb_mul[3] = {0, 0, 0};
  int min_points = 6;
  double eig_mult = 0.01;
  std::map<int64_t, Leaf> leaves;                 // ordered, VGC_H:195
  std::vector<int64_t> centroid_leaf_idx;         // voxel_centroids_leaf_indices_
  bool valid = false;

  void build(const float* pts, size_t n, size_t stride) {  // VGC_I:48-370
    leaves.clear();
    centroid_leaf_idx.clear();
    valid = false;
    if (n == 0) return;
    inv_leaf = 1.0f / leaf;  // pcl::VoxelGrid::setLeafSize
    float mn[3], mx[3];
    for (int a = 0; a < 3; ++a) { mn[a] = std::numeric_limits<float>::max(); mx[a] = -std::numeric_limits<float>::max(); }
    for (size_t i = 0; i < n; ++i)
      for (int a = 0; a < 3; ++a) {
        const float v = pts[i * stride + a];
        mn[a] = std::min(mn[a], v);
        mx[a] = std::max(mx[a], v);
      }
    const int64_t dx = (int64_t)((mx[0] - mn[0]) * inv_leaf) + 1;
    const int64_t dy = (int64_t)((mx[1] - mn[1]) * inv_leaf) + 1;
    const int64_t dz = (int64_t)((mx[2] - mn[2]) * inv_leaf) + 1;
    if (dx * dy * dz > (int64_t)std::numeric_limits<int32_t>::max()) return;  // VGC_I:79-84
    for (int a = 0; a < 3; ++a) {
      min_b[a] = (int)std::floor(mn[a] * inv_leaf);
      max_b[a] = (int)std::floor(mx[a] * inv_leaf);
      div_b[a] = max_b[a] - min_b[a] + 1;
    }
    divb_mul[0] = 1; divb_mul[1] = div_b[0]; divb_mul[2] = div_b[0] * div_b[1];
    valid = true;
    for (size_t cp = 0; cp < n; ++cp) {  // pass 1, VGC_I:209-264
      const float* p = pts + cp * stride;
      const int ijk0 = (int)(std::floor(p[0] * inv_leaf) - (float)min_b[0]);
      const int ijk1 = (int)(std::floor(p[1] * inv_leaf) - (float)min_b[1]);
      const int ijk2 = (int)(std::floor(p[2] * inv_leaf) - (float)min_b[2]);
      const int idx = ijk0 * divb_mul[0] + ijk1 * divb_mul[1] + ijk2 * divb_mul[2];
      Leaf& l = leaves[(int64_t)idx];
      if (l.nr_points == 0) { l.ijk_abs[0] = ijk0 + min_b[0]; l.ijk_abs[1] = ijk1 + min_b[1]; l.ijk_abs[2] = ijk2 + min_b[2]; }
      const double d[3] = {(double)p[0], (double)p[1], (double)p[2]};
      for (int a = 0; a < 3; ++a) l.mean[a] += d[a];
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) l.cov.a[a][b] += d[a] * d[b];
      for (int a = 0; a < 3; ++a) l.centroid[a] += p[a];
      ++l.nr_points;
    }
    for (auto& kv : leaves) {  // pass 2, VGC_I:282-367
      Leaf& l = kv.second;
      l.n_raw = l.nr_points;
      const float nf = (float)l.nr_points;
      for (int a = 0; a < 3; ++a) l.centroid[a] /= nf;
      const double pt_sum[3] = {l.mean[0], l.mean[1], l.mean[2]};
      for (int a = 0; a < 3; ++a) l.mean[a] /= l.nr_points;
      if (l.nr_points < min_points) continue;
      centroid_leaf_idx.push_back(kv.first);
      const double n = (double)l.nr_points;
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b)
          l.cov.a[a][b] = (l.cov.a[a][b] - 2 * (pt_sum[a] * l.mean[b])) / l.nr_points + l.mean[a] * l.mean[b];
      const double f = (l.nr_points - 1.0) / l.nr_points;
      (void)n;
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) l.cov.a[a][b] *= f;
      double w[3];
      M3 V;
      eigh3(l.cov, w, V);
      if (w[0] < 0 || w[1] < 0 || w[2] <= 0) { l.nr_points = -1; continue; }  // VGC_I:337-341
      const double min_ev = eig_mult * w[2];
      if (w[0] < min_ev) {  // VGC_I:346-356
        w[0] = min_ev;
        if (w[1] < min_ev) w[1] = min_ev;
        M3 D{{{w[0], 0, 0}, {0, w[1], 0}, {0, 0, w[2]}}};
        l.cov = mul3(mul3(V, D), inverse3(V));
      }
      for (int a = 0; a < 3; ++a) l.evals[a] = w[a];
      l.icov = inverse3(l.cov);
      double mxc = -INFINITY, mnc = INFINITY;
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) { mxc = std::max(mxc, l.icov.a[a][b]); mnc = std::min(mnc, l.icov.a[a][b]); }
      if (mxc == INFINITY || mnc == -INFINITY) l.nr_points = -1;  // VGC_I:360-364
    }
  }

  // VGC_I:373-404.  offsets: k x 3.  Returns pointers in offset order.
  int neighborhood(const int (*off)[3], int k, const float* pt, const Leaf** out) const {
    if (!valid) return 0;
    const int ijk[3] = {(int)std::floor(pt[0] / leaf), (int)std::floor(pt[1] / leaf), (int)std::floor(pt[2] / leaf)};
    int cnt = 0;
    for (int ni = 0; ni < k; ++ni) {
      bool inside = true;
      for (int a = 0; a < 3; ++a) {
        const int c = ijk[a] + off[ni][a];
        if (c < min_b[a] || c > max_b[a]) inside = false;
      }
      if (!inside) continue;
      const int64_t idx = (int64_t)(ijk[0] + off[ni][0] - min_b[0]) * divb_mul[0] +
                          (int64_t)(ijk[1] + off[ni][1] - min_b[1]) * divb_mul[1] +
                          (int64_t)(ijk[2] + off[ni][2] - min_b[2]) * divb_mul[2];
      auto it = leaves.find((int64_t)(int)idx);
      if (it != leaves.end() && it->second.nr_points >= min_points) out[cnt++] = &it->second;
    }
    return cnt;
  }

  // VGC_H:470-499 (FLANN radius search over the centroids of leaves with n_raw >= min_points,
  // no validity test on nr_points).  Exact 27-cell scan, ascending squared distance.
  int radius_search(const float* pt, double radius, const Leaf** out, float* d2out) const {
    if (!valid) return 0;
    const float r2 = (float)(radius * radius);
    const int ijk[3] = {(int)std::floor(pt[0] / leaf), (int)std::floor(pt[1] / leaf), (int)std::floor(pt[2] / leaf)};
    struct Hit { float d2; int64_t idx; const Leaf* l; };
    Hit hits[27];
    int cnt = 0;
    for (int dz = -1; dz <= 1; ++dz)
      for (int dy = -1; dy <= 1; ++dy)
        for (int dx = -1; dx <= 1; ++dx) {
          const int c[3] = {ijk[0] + dx, ijk[1] + dy, ijk[2] + dz};
          if (c[0] < min_b[0] || c[0] > max_b[0] || c[1] < min_b[1] || c[1] > max_b[1] || c[2] < min_b[2] || c[2] > max_b[2]) continue;
          const int64_t idx = (int64_t)(c[0] - min_b[0]) * divb_mul[0] + (int64_t)(c[1] - min_b[1]) * divb_mul[1] +
                              (int64_t)(c[2] - min_b[2]) * divb_mul[2];
          auto it = leaves.find(idx);
          if (it == leaves.end() || it->second.n_raw < min_points) continue;
          const Leaf& l = it->second;
          const float ex = pt[0] - l.centroid[0], ey = pt[1] - l.centroid[1], ez = pt[2] - l.centroid[2];
          const float d2 = (ex * ex + ey * ey) + ez * ez;
          if (d2 < r2) hits[cnt++] = Hit{d2, idx, &l};
        }
    std::sort(hits, hits + cnt, [](const Hit& a, const Hit& b) { return a.d2 < b.d2 || (a.d2 == b.d2 && a.idx < b.idx); });
    for (int i = 0; i < cnt; ++i) { out[i] = hits[i].l; if (d2out) d2out[i] = hits[i].d2; }
    return cnt;
  }
};

static const int OFF1[1][3] = {{0, 0, 0}};
static const int OFF7[7][3] = {{0, 0, 0}, {1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};  // VGC_I:423-430
static int OFF26[26][3];
static bool off26_init = [] {  // pcl::getAllNeighborCellIndices(): 13 half neighbours, then their negations
  int idx = 0;
  for (int i = -1; i <= 1; ++i)
    for (int j = -1; j <= 1; ++j) { OFF26[idx][0] = i; OFF26[idx][1] = j; OFF26[idx][2] = -1; ++idx; }
  for (int i = -1; i <= 1; ++i) { OFF26[idx][0] = i; OFF26[idx][1] = -1; OFF26[idx][2] = 0; ++idx; }
  OFF26[idx][0] = -1; OFF26[idx][1] = 0; OFF26[idx][2] = 0; ++idx;
  for (int i = 0; i < 13; ++i)
    for (int a = 0; a < 3; ++a) OFF26[13 + i][a] = -OFF26[i][a];
  return true;
}();

enum { KDTREE = 0, DIRECT26 = 1, DIRECT7 = 2, DIRECT1 = 3 };  // NDT_H:52-57

// ----------------------------------------------------------------------------------------
// NDT restatement
// ----------------------------------------------------------------------------------------
struct TraceRec { double p[6]; double score; double g[6]; double H[36]; double step; int compute_hessian; int kind; };

struct Ndt {
  // NDT_I:46-69 defaults
  float resolution = 1.0f;
  double step_size = 0.1, outlier_ratio = 0.55, trans_eps = 0.1;
  int max_iterations = 35, search_method = DIRECT7, num_threads = 1;
  double d1 = 0, d2 = 0, d3 = 0;
  Grid grid;
  std::vector<float> target;   // xyz packed
  std::vector<float> source;   // xyz packed
  // state (pcl::Registration members)
  float final_T[16], transformation[16];
  int nr_iterations = 0;
  bool converged = false;
  double trans_probability = 0;
  std::vector<float> output;   // transformed source after align
  std::vector<TraceRec> trace;
  // angle derivative tables
  float j_ang[8][4], h_ang[16][4];
  double j_ang_d[8][3], h_ang_d[15][3];
  long long n_pairs = 0;  // (point, neighbour) pairs in the last computeDerivatives

  void gauss() {  // NDT_I:80-87
    const double c1 = 10 * (1 - outlier_ratio);
    const double c2 = outlier_ratio / std::pow((double)resolution, 3);
    d3 = -std::log(c2);
    d1 = -std::log(c1 + c2) - d3;
    d2 = -2 * std::log((-std::log(c1 * std::exp(-0.5) + c2) - d3) / d1);
  }
  void init_grid() {  // NDT_H:271-278
    grid.leaf = resolution;
    grid.build(target.data(), target.size() / 3, 3);
  }

  void angle_derivatives(const double p[6], bool compute_hessian = true) {  // NDT_I:286-398
    double cx, cy, cz, sx, sy, sz;
    if (std::fabs(p[3]) < 10e-5) { cx = 1.0; sx = 0.0; } else { cx = std::cos(p[3]); sx = std::sin(p[3]); }
    if (std::fabs(p[4]) < 10e-5) { cy = 1.0; sy = 0.0; } else { cy = std::cos(p[4]); sy = std::sin(p[4]); }
    if (std::fabs(p[5]) < 10e-5) { cz = 1.0; sz = 0.0; } else { cz = std::cos(p[5]); sz = std::sin(p[5]); }
    const double J[8][3] = {
        {(-sx * sz + cx * sy * cz), (-sx * cz - cx * sy * sz), (-cx * cy)},
        {(cx * sz + sx * sy * cz), (cx * cz - sx * sy * sz), (-sx * cy)},
        {(-sy * cz), sy * sz, cy},
        {sx * cy * cz, (-sx * cy * sz), sx * sy},
        {(-cx * cy * cz), cx * cy * sz, (-cx * sy)},
        {(-cy * sz), (-cy * cz), 0},
        {(cx * cz - sx * sy * sz), (-cx * sz - sx * sy * cz), 0},
        {(sx * cz + cx * sy * sz), (cx * sy * cz - sx * sz), 0}};
    for (int r = 0; r < 8; ++r) {
      for (int c = 0; c < 3; ++c) { j_ang_d[r][c] = J[r][c]; j_ang[r][c] = (float)J[r][c]; }
      j_ang[r][3] = 0.f;
    }
    if (!compute_hessian) return;
    const double Hh[15][3] = {
        {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), sx * cy},      // a2
        {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), (-cx * cy)},   // a3
        {(cx * cy * cz), (-cx * cy * sz), (cx * sy)},                         // b2
        {(sx * cy * cz), (-sx * cy * sz), (sx * sy)},                         // b3
        {(-sx * cz - cx * sy * sz), (sx * sz - cx * sy * cz), 0},             // c2
        {(cx * cz - sx * sy * sz), (-sx * sy * cz - cx * sz), 0},             // c3
        {(-cy * cz), (cy * sz), (sy)},                                        // d1
        {(-sx * sy * cz), (sx * sy * sz), (sx * cy)},                         // d2
        {(cx * sy * cz), (-cx * sy * sz), (-cx * cy)},                        // d3
        {(sy * sz), (sy * cz), 0},                                            // e1
        {(-sx * cy * sz), (-sx * cy * cz), 0},                                // e2
        {(cx * cy * sz), (cx * cy * cz), 0},                                  // e3
        {(-cy * cz), (cy * sz), 0},                                           // f1
        {(-cx * sz - sx * sy * cz), (-cx * cz + sx * sy * sz), 0},            // f2
        {(-sx * sz + cx * sy * cz), (-cx * sy * sz - sx * cz), 0}};           // f3
    for (int r = 0; r < 15; ++r) {
      for (int c = 0; c < 3; ++c) { h_ang_d[r][c] = Hh[r][c]; h_ang[r][c] = (float)Hh[r][c]; }
      h_ang[r][3] = 0.f;
    }
    for (int c = 0; c < 4; ++c) h_ang[15][c] = 0.f;
  }

  // NDT_I:401-445 (float overload).  pg: 4x6, ph: 24x6, both row-major here.
  void point_derivatives_f(const double x[3], float pg[4][6], float ph[24][6], bool compute_hessian = true) const {
    const float x4[4] = {(float)x[0], (float)x[1], (float)x[2], 0.0f};
    float xj[8];
    for (int r = 0; r < 8; ++r)
      xj[r] = ((j_ang[r][0] * x4[0] + j_ang[r][1] * x4[1]) + j_ang[r][2] * x4[2]) + j_ang[r][3] * x4[3];
    pg[1][3] = xj[0]; pg[2][3] = xj[1];
    pg[0][4] = xj[2]; pg[1][4] = xj[3]; pg[2][4] = xj[4];
    pg[0][5] = xj[5]; pg[1][5] = xj[6]; pg[2][5] = xj[7];
    if (!compute_hessian) return;
    float xh[16];
    for (int r = 0; r < 16; ++r)
      xh[r] = ((h_ang[r][0] * x4[0] + h_ang[r][1] * x4[1]) + h_ang[r][2] * x4[2]) + h_ang[r][3] * x4[3];
    const float a[4] = {0, xh[0], xh[1], 0.0f}, b[4] = {0, xh[2], xh[3], 0.0f}, c[4] = {0, xh[4], xh[5], 0.0f};
    const float d[4] = {xh[6], xh[7], xh[8], 0.0f}, e[4] = {xh[9], xh[10], xh[11], 0.0f}, f[4] = {xh[12], xh[13], xh[14], 0.0f};
    for (int k = 0; k < 4; ++k) {
      ph[12 + k][3] = a[k]; ph[16 + k][3] = b[k]; ph[20 + k][3] = c[k];
      ph[12 + k][4] = b[k]; ph[16 + k][4] = d[k]; ph[20 + k][4] = e[k];
      ph[12 + k][5] = c[k]; ph[16 + k][5] = e[k]; ph[20 + k][5] = f[k];
    }
  }
  // NDT_I:448-488 (double overload).  pg: 3x6, ph: 18x6.
  void point_derivatives_d(const double x[3], double pg[3][6], double ph[18][6]) const {
    auto dot = [&](const double* v) { return (x[0] * v[0] + x[1] * v[1]) + x[2] * v[2]; };
    pg[1][3] = dot(j_ang_d[0]); pg[2][3] = dot(j_ang_d[1]);
    pg[0][4] = dot(j_ang_d[2]); pg[1][4] = dot(j_ang_d[3]); pg[2][4] = dot(j_ang_d[4]);
    pg[0][5] = dot(j_ang_d[5]); pg[1][5] = dot(j_ang_d[6]); pg[2][5] = dot(j_ang_d[7]);
    const double a[3] = {0, dot(h_ang_d[0]), dot(h_ang_d[1])}, b[3] = {0, dot(h_ang_d[2]), dot(h_ang_d[3])};
    const double c[3] = {0, dot(h_ang_d[4]), dot(h_ang_d[5])};
    const double d[3] = {dot(h_ang_d[6]), dot(h_ang_d[7]), dot(h_ang_d[8])};
    const double e[3] = {dot(h_ang_d[9]), dot(h_ang_d[10]), dot(h_ang_d[11])};
    const double f[3] = {dot(h_ang_d[12]), dot(h_ang_d[13]), dot(h_ang_d[14])};
    for (int k = 0; k < 3; ++k) {
      ph[9 + k][3] = a[k]; ph[12 + k][3] = b[k]; ph[15 + k][3] = c[k];
      ph[9 + k][4] = b[k]; ph[12 + k][4] = d[k]; ph[15 + k][4] = e[k];
      ph[9 + k][5] = c[k]; ph[12 + k][5] = e[k]; ph[15 + k][5] = f[k];
    }
  }

  // NDT_I:491-548.  All per-term arithmetic in float, accumulated into double.
  double update_derivatives(double g[6], double H[36], const float pg[4][6], const float ph[24][6],
                            const double x_trans[3], const M3& c_inv, bool compute_hessian) const {
    const float x4[4] = {(float)x_trans[0], (float)x_trans[1], (float)x_trans[2], 0.0f};
    float C[4][4];
    for (int i = 0; i < 4; ++i)
      for (int j = 0; j < 4; ++j) C[i][j] = (i < 3 && j < 3) ? (float)c_inv.a[i][j] : 0.0f;
    const float gauss_d2 = (float)d2;
    float xC[4];
    for (int j = 0; j < 4; ++j) xC[j] = ((x4[0] * C[0][j] + x4[1] * C[1][j]) + x4[2] * C[2][j]) + x4[3] * C[3][j];
    const float q = ((x4[0] * xC[0] + x4[1] * xC[1]) + x4[2] * xC[2]) + x4[3] * xC[3];
    float e = (float)std::exp((double)(-gauss_d2 * q * 0.5f));
    const float score_inc = (float)(-d1 * (double)e);
    e = gauss_d2 * e;
    if (e > 1 || e < 0 || e != e) return 0;
    e = (float)((double)e * d1);
    float CJ[4][6];
    for (int r = 0; r < 4; ++r)
      for (int c = 0; c < 6; ++c) CJ[r][c] = ((C[r][0] * pg[0][c] + C[r][1] * pg[1][c]) + C[r][2] * pg[2][c]) + C[r][3] * pg[3][c];
    float xCJ[6];
    for (int c = 0; c < 6; ++c) xCJ[c] = ((x4[0] * CJ[0][c] + x4[1] * CJ[1][c]) + x4[2] * CJ[2][c]) + x4[3] * CJ[3][c];
    for (int c = 0; c < 6; ++c) g[c] += (double)(e * xCJ[c]);
    if (compute_hessian) {
      float JtCJ[6][6];
      for (int a = 0; a < 6; ++a)
        for (int b = 0; b < 6; ++b)
          JtCJ[a][b] = ((pg[0][a] * CJ[0][b] + pg[1][a] * CJ[1][b]) + pg[2][a] * CJ[2][b]) + pg[3][a] * CJ[3][b];
      for (int i = 0; i < 6; ++i) {
        float xCH[6];
        for (int j = 0; j < 6; ++j)
          xCH[j] = ((xC[0] * ph[i * 4 + 0][j] + xC[1] * ph[i * 4 + 1][j]) + xC[2] * ph[i * 4 + 2][j]) + xC[3] * ph[i * 4 + 3][j];
        for (int j = 0; j < 6; ++j)
          H[i * 6 + j] += (double)(e * (-gauss_d2 * xCJ[i] * xCJ[j] + xCH[j] + JtCJ[j][i]));
      }
    }
    return (double)score_inc;
  }

  int find_neighbors(int method, const float* pt, const Leaf** nb) const {
    switch (method) {
      case KDTREE: return grid.radius_search(pt, (double)resolution, nb, nullptr);
      case DIRECT26: return grid.neighborhood(OFF26, 26, pt, nb);
      case DIRECT1: return grid.neighborhood(OFF1, 1, pt, nb);
      default: return grid.neighborhood(OFF7, 7, pt, nb);
    }
  }

  // NDT_I:175-283
  double compute_derivatives(double g[6], double H[36], const std::vector<float>& trans_cloud, const double p[6],
                             bool compute_hessian = true) {
    const int nt = std::max(1, num_threads);
    std::vector<double> scores(nt, 0.0), gs(nt * 6, 0.0), Hs(nt * 36, 0.0);
    std::vector<long long> pairs(nt, 0);
    angle_derivatives(p);
    const long n = (long)(source.size() / 3);
#pragma omp parallel for num_threads(nt) schedule(guided, 8)
    for (long idx = 0; idx < n; ++idx) {
#ifdef _OPENMP
      const int t = omp_get_thread_num();
#else
      const int t = 0;
#endif
      float pg[4][6];
      float ph[24][6];
      std::memset(pg, 0, sizeof(pg));
      pg[0][0] = pg[1][1] = pg[2][2] = 1.f;
      std::memset(ph, 0, sizeof(ph));
      const float* xt = &trans_cloud[idx * 3];
      const Leaf* nb[27];
      const int cnt = find_neighbors(search_method, xt, nb);
      double score_pt = 0, g_pt[6] = {0, 0, 0, 0, 0, 0}, H_pt[36];
      for (int k = 0; k < 36; ++k) H_pt[k] = 0;
      for (int k = 0; k < cnt; ++k) {
        const Leaf* cell = nb[k];
        const double x[3] = {(double)source[idx * 3], (double)source[idx * 3 + 1], (double)source[idx * 3 + 2]};
        const double x_trans[3] = {(double)xt[0] - cell->mean[0], (double)xt[1] - cell->mean[1], (double)xt[2] - cell->mean[2]};
        point_derivatives_f(x, pg, ph);
        score_pt += update_derivatives(g_pt, H_pt, pg, ph, x_trans, cell->icov, compute_hessian);
      }
      pairs[t] += cnt;
      scores[t] += score_pt;
      for (int k = 0; k < 6; ++k) gs[t * 6 + k] += g_pt[k];
      for (int k = 0; k < 36; ++k) Hs[t * 36 + k] += H_pt[k];
    }
    double score = 0;
    for (int k = 0; k < 6; ++k) g[k] = 0;
    for (int k = 0; k < 36; ++k) H[k] = 0;
    n_pairs = 0;
    for (int t = 0; t < nt; ++t) {
      score += scores[t];
      n_pairs += pairs[t];
      for (int k = 0; k < 6; ++k) g[k] += gs[t * 6 + k];
      for (int k = 0; k < 36; ++k) H[k] += Hs[t * 36 + k];
    }
    return score;
  }

  // NDT_I:551-641: Hessian only, all double, neighbours ALWAYS by radius search, serial.
  void compute_hessian(double H[36], const std::vector<float>& trans_cloud) {
    for (int k = 0; k < 36; ++k) H[k] = 0;
    double pg[3][6], ph[18][6];
    std::memset(pg, 0, sizeof(pg));
    pg[0][0] = pg[1][1] = pg[2][2] = 1.0;
    std::memset(ph, 0, sizeof(ph));
    const long n = (long)(source.size() / 3);
    for (long idx = 0; idx < n; ++idx) {
      const float* xt = &trans_cloud[idx * 3];
      const Leaf* nb[27];
      const int cnt = grid.radius_search(xt, (double)resolution, nb, nullptr);
      for (int k = 0; k < cnt; ++k) {
        const Leaf* cell = nb[k];
        const double x[3] = {(double)source[idx * 3], (double)source[idx * 3 + 1], (double)source[idx * 3 + 2]};
        const double xtr[3] = {(double)xt[0] - cell->mean[0], (double)xt[1] - cell->mean[1], (double)xt[2] - cell->mean[2]};
        const M3& C = cell->icov;
        point_derivatives_d(x, pg, ph);
        // updateHessian, NDT_I:612-641
        auto Cv = [&](const double v[3], double o[3]) {
          for (int r = 0; r < 3; ++r) o[r] = (C.a[r][0] * v[0] + C.a[r][1] * v[1]) + C.a[r][2] * v[2];
        };
        auto dot3 = [](const double* a, const double* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; };
        double Cx[3];
        Cv(xtr, Cx);
        double e = d2 * std::exp(-d2 * dot3(xtr, Cx) / 2);
        if (e > 1 || e < 0 || e != e) continue;
        e *= d1;
        for (int i = 0; i < 6; ++i) {
          const double pgi[3] = {pg[0][i], pg[1][i], pg[2][i]};
          double cov_dxd_pi[3];
          Cv(pgi, cov_dxd_pi);
          for (int j = 0; j < 6; ++j) {
            const double pgj[3] = {pg[0][j], pg[1][j], pg[2][j]};
            double Cpgj[3], Cph[3];
            Cv(pgj, Cpgj);
            const double phb[3] = {ph[3 * i][j], ph[3 * i + 1][j], ph[3 * i + 2][j]};
            Cv(phb, Cph);
            H[i * 6 + j] += e * (-d2 * dot3(xtr, cov_dxd_pi) * dot3(xtr, Cpgj) + dot3(xtr, Cph) + dot3(pgj, cov_dxd_pi));
          }
        }
      }
    }
  }

  void transform_cloud(const float T[16], std::vector<float>& out) const {
    const size_t n = source.size() / 3;
    out.resize(n * 3);
    for (size_t i = 0; i < n; ++i) transform_point(T, &source[i * 3], &out[i * 3]);
  }

  // NDT_I:644-677
  static bool update_interval_mt(double& a_l, double& f_l, double& g_l, double& a_u, double& f_u, double& g_u,
                                 double a_t, double f_t, double g_t) {
    if (f_t > f_l) { a_u = a_t; f_u = f_t; g_u = g_t; return false; }
    else if (g_t * (a_l - a_t) > 0) { a_l = a_t; f_l = f_t; g_l = g_t; return false; }
    else if (g_t * (a_l - a_t) < 0) { a_u = a_l; f_u = f_l; g_u = g_l; a_l = a_t; f_l = f_t; g_l = g_t; return false; }
    else return true;
  }
  // NDT_I:680-757
  static double trial_value_selection_mt(double a_l, double f_l, double g_l, double a_u, double f_u, double g_u,
                                         double a_t, double f_t, double g_t) {
    if (f_t > f_l) {
      const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
      const double w = std::sqrt(z * z - g_t * g_l);
      const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
      const double a_q = a_l - 0.5 * (a_l - a_t) * g_l / (g_l - (f_l - f_t) / (a_l - a_t));
      if (std::fabs(a_c - a_l) < std::fabs(a_q - a_l)) return a_c;
      return 0.5 * (a_q + a_c);
    } else if (g_t * g_l < 0) {
      const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
      const double w = std::sqrt(z * z - g_t * g_l);
      const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
      const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
      if (std::fabs(a_c - a_t) >= std::fabs(a_s - a_t)) return a_c;
      return a_s;
    } else if (std::fabs(g_t) <= std::fabs(g_l)) {
      const double z = 3 * (f_t - f_l) / (a_t - a_l) - g_t - g_l;
      const double w = std::sqrt(z * z - g_t * g_l);
      const double a_c = a_l + (a_t - a_l) * (w - g_l - z) / (g_t - g_l + 2 * w);
      const double a_s = a_l - (a_l - a_t) / (g_l - g_t) * g_l;
      const double a_t_next = (std::fabs(a_c - a_t) < std::fabs(a_s - a_t)) ? a_c : a_s;
      if (a_t > a_l) return std::min(a_t + 0.66 * (a_u - a_t), a_t_next);
      return std::max(a_t + 0.66 * (a_u - a_t), a_t_next);
    } else {
      const double z = 3 * (f_t - f_u) / (a_t - a_u) - g_t - g_u;
      const double w = std::sqrt(z * z - g_t * g_u);
      return a_u + (a_t - a_u) * (w - g_u - z) / (g_t - g_u + 2 * w);
    }
  }

  void record(const double p[6], double score, const double g[6], const double H[36], double step, int ch, int kind) {
    TraceRec r;
    std::memcpy(r.p, p, sizeof(r.p));
    r.score = score;
    std::memcpy(r.g, g, sizeof(r.g));
    std::memcpy(r.H, H, sizeof(r.H));
    r.step = step; r.compute_hessian = ch; r.kind = kind;
    trace.push_back(r);
  }

  // NDT_I:760-916
  double step_length_mt(const double x[6], double step_dir[6], double step_init, double step_max, double step_min,
                        double& score, double g[6], double H[36], std::vector<float>& trans_cloud) {
    const double phi_0 = -score;
    double d_phi_0 = 0;
    for (int k = 0; k < 6; ++k) d_phi_0 += g[k] * step_dir[k];
    d_phi_0 = -d_phi_0;
    double x_t[6];
    if (d_phi_0 >= 0) {
      if (d_phi_0 == 0) return 0;
      d_phi_0 *= -1;
      for (int k = 0; k < 6; ++k) step_dir[k] *= -1;
    }
    const int max_step_iterations = 10;
    int step_iterations = 0;
    const double mu = 1.e-4, nu = 0.9;
    double a_l = 0, a_u = 0;
    double f_l = phi_0 - phi_0 - mu * d_phi_0 * a_l;  // auxilaryFunction_PsiMT, NDT_H:425-429
    double g_l = d_phi_0 - mu * d_phi_0;              // auxilaryFunction_dPsiMT, NDT_H:437-441
    double f_u = phi_0 - phi_0 - mu * d_phi_0 * a_u;
    double g_u = d_phi_0 - mu * d_phi_0;
    bool interval_converged = (step_max - step_min) > 0, open_interval = true;  // NDT_I:807 (sic)
    double a_t = step_init;
    a_t = std::min(a_t, step_max);
    a_t = std::max(a_t, step_min);
    for (int k = 0; k < 6; ++k) x_t[k] = x[k] + step_dir[k] * a_t;
    pose_to_matrix(x_t, final_T);
    transform_cloud(final_T, trans_cloud);
    score = compute_derivatives(g, H, trans_cloud, x_t, true);
    record(x_t, score, g, H, a_t, 1, 1);
    double phi_t = -score;
    double d_phi_t = 0;
    for (int k = 0; k < 6; ++k) d_phi_t += g[k] * step_dir[k];
    d_phi_t = -d_phi_t;
    double psi_t = phi_t - phi_0 - mu * d_phi_0 * a_t;
    double d_psi_t = d_phi_t - mu * d_phi_0;
    while (!interval_converged && step_iterations < max_step_iterations && !(psi_t <= 0 && d_phi_t <= -nu * d_phi_0)) {
      if (open_interval) a_t = trial_value_selection_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, psi_t, d_psi_t);
      else a_t = trial_value_selection_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, phi_t, d_phi_t);
      a_t = std::min(a_t, step_max);
      a_t = std::max(a_t, step_min);
      for (int k = 0; k < 6; ++k) x_t[k] = x[k] + step_dir[k] * a_t;
      pose_to_matrix(x_t, final_T);
      transform_cloud(final_T, trans_cloud);
      score = compute_derivatives(g, H, trans_cloud, x_t, false);
      record(x_t, score, g, H, a_t, 0, 2);
      phi_t = -score;
      d_phi_t = 0;
      for (int k = 0; k < 6; ++k) d_phi_t += g[k] * step_dir[k];
      d_phi_t = -d_phi_t;
      psi_t = phi_t - phi_0 - mu * d_phi_0 * a_t;
      d_psi_t = d_phi_t - mu * d_phi_0;
      if (open_interval && (psi_t <= 0 && d_psi_t >= 0)) {
        open_interval = false;
        f_l = f_l + phi_0 - mu * d_phi_0 * a_l;
        g_l = g_l + mu * d_phi_0;
        f_u = f_u + phi_0 - mu * d_phi_0 * a_u;
        g_u = g_u + mu * d_phi_0;
      }
      if (open_interval) interval_converged = update_interval_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, psi_t, d_psi_t);
      else interval_converged = update_interval_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, phi_t, d_phi_t);
      step_iterations++;
    }
    if (step_iterations) {
      compute_hessian(H, trans_cloud);
      record(x_t, score, g, H, a_t, 1, 3);
    }
    return a_t;
  }

  // pcl::Registration::align + NDT_I:73-164
  void align(const float guess[16]) {
    trace.clear();
    const size_t n = source.size() / 3;
    output = source;
    static const float I16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    std::memcpy(final_T, I16, sizeof(I16));
    std::memcpy(transformation, I16, sizeof(I16));
    nr_iterations = 0;
    converged = false;
    gauss();
    if (std::memcmp(guess, I16, sizeof(I16)) != 0) {
      bool differs = false;
      for (int k = 0; k < 16; ++k) differs |= (guess[k] != I16[k]);
      if (differs) {
        std::memcpy(final_T, guess, sizeof(I16));
        transform_cloud(guess, output);
      }
    }
    double p[6], delta_p[6], g[6], H[36];
    float eul[3];
    euler_012(final_T, eul);
    p[0] = final_T[12]; p[1] = final_T[13]; p[2] = final_T[14];
    p[3] = eul[0]; p[4] = eul[1]; p[5] = eul[2];
    double score = compute_derivatives(g, H, output, p);
    record(p, score, g, H, 0.0, 1, 0);
    while (!converged) {
      double neg_g[6];
      for (int k = 0; k < 6; ++k) neg_g[k] = -g[k];
      svd_solve6(H, neg_g, delta_p);
      double nrm2 = 0;
      for (int k = 0; k < 6; ++k) nrm2 += delta_p[k] * delta_p[k];
      double delta_p_norm = std::sqrt(nrm2);
      if (delta_p_norm == 0 || delta_p_norm != delta_p_norm) {
        trans_probability = score / (double)n;
        converged = delta_p_norm == delta_p_norm;
        return;
      }
      for (int k = 0; k < 6; ++k) delta_p[k] /= delta_p_norm;  // normalize()
      delta_p_norm = step_length_mt(p, delta_p, delta_p_norm, step_size, trans_eps / 2, score, g, H, output);
      for (int k = 0; k < 6; ++k) delta_p[k] *= delta_p_norm;
      pose_to_matrix(delta_p, transformation);
      for (int k = 0; k < 6; ++k) p[k] = p[k] + delta_p[k];
      if (nr_iterations > max_iterations || (nr_iterations && (std::fabs(delta_p_norm) < trans_eps))) converged = true;
      nr_iterations++;
    }
    trans_probability = score / (double)n;
  }

  // NDT_I:919-952
  double calculate_score(const std::vector<float>& cloud) const {
    double score = 0;
    const size_t n = cloud.size() / 3;
    for (size_t idx = 0; idx < n; ++idx) {
      const float* xt = &cloud[idx * 3];
      const Leaf* nb[27];
      const int cnt = grid.radius_search(xt, (double)resolution, nb, nullptr);
      for (int k = 0; k < cnt; ++k) {
        const Leaf* cell = nb[k];
        const double x[3] = {(double)xt[0] - cell->mean[0], (double)xt[1] - cell->mean[1], (double)xt[2] - cell->mean[2]};
        const M3& C = cell->icov;
        double Cx[3];
        for (int r = 0; r < 3; ++r) Cx[r] = (C.a[r][0] * x[0] + C.a[r][1] * x[1]) + C.a[r][2] * x[2];
        const double e = std::exp(-d2 * ((x[0] * Cx[0] + x[1] * Cx[1]) + x[2] * Cx[2]) / 2);
        const double score_inc = -d1 * e - d3;
        score += score_inc / cnt;
      }
    }
    return score / (double)n;
  }

  // pcl::Registration::getFitnessScore(max_range): mean squared 1-NN distance of the source
  // transformed by final_transformation_ against the RAW target cloud.
  struct NNGrid {
    float cell = 1.f;
    std::unordered_map<uint64_t, std::vector<uint32_t>> cells;
    int mn[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
    static uint64_t key(int i, int j, int k) {
      return ((uint64_t)(uint32_t)(i + (1 << 20)) & 0x1FFFFF) | (((uint64_t)(uint32_t)(j + (1 << 20)) & 0x1FFFFF) << 21) |
             (((uint64_t)(uint32_t)(k + (1 << 20)) & 0x1FFFFF) << 42);
    }
  } nn;
  bool nn_valid = false;
  void build_nn() {
    nn.cells.clear();
    nn.cell = std::max(resolution, 0.25f);
    const size_t m = target.size() / 3;
    for (int a = 0; a < 3; ++a) { nn.mn[a] = INT32_MAX; nn.mx[a] = INT32_MIN; }
    for (size_t i = 0; i < m; ++i) {
      int c[3];
      for (int a = 0; a < 3; ++a) {
        c[a] = (int)std::floor(target[i * 3 + a] / nn.cell);
        nn.mn[a] = std::min(nn.mn[a], c[a]);
        nn.mx[a] = std::max(nn.mx[a], c[a]);
      }
      nn.cells[NNGrid::key(c[0], c[1], c[2])].push_back((uint32_t)i);
    }
    nn_valid = true;
  }
  float nn_dist2(const float* q) const {
    int c[3];
    for (int a = 0; a < 3; ++a) c[a] = (int)std::floor(q[a] / nn.cell);
    float best = std::numeric_limits<float>::max();
    int maxring = 0;
    for (int a = 0; a < 3; ++a) maxring = std::max(maxring, std::max(std::abs(c[a] - nn.mn[a]), std::abs(c[a] - nn.mx[a])));
    for (int ring = 0; ring <= maxring + 1; ++ring) {
      if (ring >= 1) {
        // everything in rings >= `ring` is at least (ring-1)*cell away
        const float lim = (float)(ring - 1) * nn.cell;
        if (best <= lim * lim) break;
      }
      if (ring > 4) {
        // a query far outside the target (a 60 m scan against odom_node's 5 m localmap window): walking ever larger empty
        // shells costs ring^2 cells each; the minimum over ALL target points is the same number
        const size_t m = target.size() / 3;
        for (size_t i = 0; i < m; ++i) {
          const float ex = q[0] - target[i * 3], ey = q[1] - target[i * 3 + 1], ez = q[2] - target[i * 3 + 2];
          const float d = (ex * ex + ey * ey) + ez * ez;
          if (d < best) best = d;
        }
        break;
      }
      for (int dz = -ring; dz <= ring; ++dz)
        for (int dy = -ring; dy <= ring; ++dy)
          for (int dx = -ring; dx <= ring; ++dx) {
            if (std::max(std::abs(dx), std::max(std::abs(dy), std::abs(dz))) != ring) continue;
            auto it = nn.cells.find(NNGrid::key(c[0] + dx, c[1] + dy, c[2] + dz));
            if (it == nn.cells.end()) continue;
            for (uint32_t i : it->second) {
              const float ex = q[0] - target[i * 3], ey = q[1] - target[i * 3 + 1], ez = q[2] - target[i * 3 + 2];
              const float d = (ex * ex + ey * ey) + ez * ez;
              if (d < best) best = d;
            }
          }
    }
    return best;
  }
  double fitness(double max_range) {
    if (!nn_valid) build_nn();
    std::vector<float> tc;
    transform_cloud(final_T, tc);
    const long n = (long)(tc.size() / 3);
    std::vector<float> d(n);
#pragma omp parallel for num_threads(std::max(1, num_threads)) schedule(static)
    for (long i = 0; i < n; ++i) d[i] = nn_dist2(&tc[i * 3]);
    double s = 0;
    long nr = 0;
    for (long i = 0; i < n; ++i)
      if ((double)d[i] <= max_range) { s += d[i]; ++nr; }
    return nr > 0 ? s / nr : std::numeric_limits<double>::max();
  }
};

static void load_xyz(std::vector<float>& dst, const float* pts, size_t n, size_t stride_floats, bool append) {
  const size_t old = append ? dst.size() / 3 : 0;
  dst.resize((old + n) * 3);
  for (size_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) dst[(old + i) * 3 + a] = pts[i * stride_floats + a];
}

}  // namespace

// ----------------------------------------------------------------------------------------
// C API (ctypes)
// ----------------------------------------------------------------------------------------
extern "C" {

void* oracle_create() { return new Ndt(); }
void oracle_destroy(void* h) { delete (Ndt*)h; }

void oracle_set_params(void* h, float resolution, double step_size, double trans_eps, int max_iter, double outlier_ratio,
                       int search_method, int num_threads) {
  Ndt* o = (Ndt*)h;
  const bool rebuild = o->resolution != resolution && !o->target.empty();
  o->resolution = resolution;
  o->step_size = step_size; o->trans_eps = trans_eps; o->max_iterations = max_iter;
  o->outlier_ratio = outlier_ratio; o->search_method = search_method;
  o->num_threads = num_threads > 0 ? num_threads : 1;
  if (rebuild) { o->init_grid(); o->nn_valid = false; }
}
void oracle_set_voxel_params(void* h, int min_points, double eig_mult) {
  Ndt* o = (Ndt*)h;
  o->grid.min_points = min_points;
  o->grid.eig_mult = eig_mult;
}
void oracle_set_target(void* h, const float* pts, size_t n, size_t stride_floats) {
  Ndt* o = (Ndt*)h;
  load_xyz(o->target, pts, n, stride_floats, false);
  o->init_grid();
  o->nn_valid = false;
}
// cpu::VoxelGrid::update contract: result == full rebuild over the concatenated cloud.
void oracle_update_target(void* h, const float* pts, size_t n, size_t stride_floats) {
  Ndt* o = (Ndt*)h;
  load_xyz(o->target, pts, n, stride_floats, true);
  o->init_grid();
  o->nn_valid = false;
}
void oracle_set_source(void* h, const float* pts, size_t n, size_t stride_floats) {
  Ndt* o = (Ndt*)h;
  load_xyz(o->source, pts, n, stride_floats, false);
}
int oracle_grid_valid(void* h) { return ((Ndt*)h)->grid.valid ? 1 : 0; }
void oracle_grid_bounds(void* h, int* min_b, int* max_b) {
  Ndt* o = (Ndt*)h;
  for (int a = 0; a < 3; ++a) { min_b[a] = o->grid.min_b[a]; max_b[a] = o->grid.max_b[a]; }
}
size_t oracle_num_leaves(void* h) { return ((Ndt*)h)->grid.leaves.size(); }
// Dump all leaves in ascending linear-index order.  Any pointer may be null.
void oracle_dump_leaves(void* h, int* ijk, int* nr_points, int* n_raw, double* mean, double* cov, double* icov,
                        float* centroid, double* evals) {
  Ndt* o = (Ndt*)h;
  size_t i = 0;
  for (auto& kv : o->grid.leaves) {
    const Leaf& l = kv.second;
    if (ijk) for (int a = 0; a < 3; ++a) ijk[i * 3 + a] = l.ijk_abs[a];
    if (nr_points) nr_points[i] = l.nr_points;
    if (n_raw) n_raw[i] = l.n_raw;
    if (mean) for (int a = 0; a < 3; ++a) mean[i * 3 + a] = l.mean[a];
    if (cov) for (int a = 0; a < 9; ++a) cov[i * 9 + a] = l.cov.a[a / 3][a % 3];
    if (icov) for (int a = 0; a < 9; ++a) icov[i * 9 + a] = l.icov.a[a / 3][a % 3];
    if (centroid) for (int a = 0; a < 3; ++a) centroid[i * 3 + a] = l.centroid[a];
    if (evals) for (int a = 0; a < 3; ++a) evals[i * 3 + a] = l.evals[a];
    ++i;
  }
}
// Neighbour sets for n already-transformed points.  out_ijk: n x 27 x 3 (absolute cell coords,
// in the order the reference visits them), out_cnt: n.
void oracle_neighbors(void* h, const float* pts, size_t n, int method, int* out_cnt, int* out_ijk) {
  Ndt* o = (Ndt*)h;
  for (size_t i = 0; i < n; ++i) {
    const Leaf* nb[27];
    const int c = o->find_neighbors(method, pts + i * 3, nb);
    out_cnt[i] = c;
    for (int k = 0; k < c; ++k)
      for (int a = 0; a < 3; ++a) out_ijk[(i * 27 + k) * 3 + a] = nb[k]->ijk_abs[a];
  }
}
void oracle_gauss(void* h, double* d1, double* d2, double* d3) {
  Ndt* o = (Ndt*)h;
  o->gauss();
  *d1 = o->d1; *d2 = o->d2; *d3 = o->d3;
}
// One computeDerivatives at transform T (column-major float 4x4) and pose vector p.
double oracle_eval_derivatives(void* h, const float* T, const double* p, int compute_hessian, double* g, double* H,
                               long long* n_pairs) {
  Ndt* o = (Ndt*)h;
  o->gauss();
  std::vector<float> tc;
  o->transform_cloud(T, tc);
  const double s = o->compute_derivatives(g, H, tc, p, compute_hessian != 0);
  if (n_pairs) *n_pairs = o->n_pairs;
  return s;
}
void oracle_compute_hessian(void* h, const float* T, const double* p, double* H) {
  Ndt* o = (Ndt*)h;
  o->gauss();
  std::vector<float> tc;
  o->transform_cloud(T, tc);
  o->angle_derivatives(p);
  o->compute_hessian(H, tc);
}
void oracle_transform_source(void* h, const float* T, float* out) {
  Ndt* o = (Ndt*)h;
  std::vector<float> tc;
  o->transform_cloud(T, tc);
  std::memcpy(out, tc.data(), tc.size() * sizeof(float));
}
struct oracle_result { float T[16]; int iterations; int converged; double score; double trans_probability; int n_evals; };
void oracle_align(void* h, const float* guess, oracle_result* r) {
  Ndt* o = (Ndt*)h;
  o->align(guess);
  std::memcpy(r->T, o->final_T, sizeof(r->T));
  r->iterations = o->nr_iterations;
  r->converged = o->converged ? 1 : 0;
  r->trans_probability = o->trans_probability;
  r->score = o->trans_probability * (double)(o->source.size() / 3);
  r->n_evals = (int)o->trace.size();
}
// trace record = 6 p + score + 6 g + 36 H + step + compute_hessian + kind = 52 doubles
int oracle_get_trace(void* h, double* out, int max_recs) {
  Ndt* o = (Ndt*)h;
  const int n = std::min<int>(max_recs, (int)o->trace.size());
  for (int i = 0; i < n; ++i) {
    const TraceRec& t = o->trace[i];
    double* d = out + (size_t)i * 52;
    std::memcpy(d, t.p, 48);
    d[6] = t.score;
    std::memcpy(d + 7, t.g, 48);
    std::memcpy(d + 13, t.H, 288);
    d[49] = t.step; d[50] = t.compute_hessian; d[51] = t.kind;
  }
  return (int)o->trace.size();
}
void oracle_get_output(void* h, float* out) {
  Ndt* o = (Ndt*)h;
  std::memcpy(out, o->output.data(), o->output.size() * sizeof(float));
}
double oracle_fitness(void* h, double max_range) { return ((Ndt*)h)->fitness(max_range); }
void oracle_build_nn(void* h) { ((Ndt*)h)->build_nn(); }
void oracle_set_final_transform(void* h, const float* T) { std::memcpy(((Ndt*)h)->final_T, T, 64); }
double oracle_calculate_score(void* h, const float* T) {
  Ndt* o = (Ndt*)h;
  o->gauss();
  std::vector<float> tc;
  o->transform_cloud(T, tc);
  return o->calculate_score(tc);
}
void oracle_pose_to_matrix(const double* p, float* T) { pose_to_matrix(p, T); }
void oracle_matrix_to_pose(const float* T, double* p) {
  float e[3];
  euler_012(T, e);
  p[0] = T[12]; p[1] = T[13]; p[2] = T[14]; p[3] = e[0]; p[4] = e[1]; p[5] = e[2];
}
void oracle_svd_solve6(const double* H, const double* b, double* x) { svd_solve6(H, b, x); }
double oracle_mt_trial(double a_l, double f_l, double g_l, double a_u, double f_u, double g_u, double a_t, double f_t, double g_t) {
  return Ndt::trial_value_selection_mt(a_l, f_l, g_l, a_u, f_u, g_u, a_t, f_t, g_t);
}
int oracle_mt_update(double* s /*a_l f_l g_l a_u f_u g_u*/, double a_t, double f_t, double g_t) {
  return Ndt::update_interval_mt(s[0], s[1], s[2], s[3], s[4], s[5], a_t, f_t, g_t) ? 1 : 0;
}
// pcl::VoxelGrid centroid downsample as odom_node applies it to every keyframe (ODOM:334-335); the restatement lives
// in thirdparty_math.h.  in: n x stride floats (x y z, intensity at ioff); out: up to n x 4 floats; returns the count.
size_t oracle_voxelgrid_downsample(const float* in, size_t n, size_t stride, size_t ioff, float leaf, float* out) {
  std::vector<float> o;
  voxelgrid_centroids(in, n, stride, ioff, leaf, o);
  std::memcpy(out, o.data(), o.size() * sizeof(float));
  return o.size() / 4;
}
// filter_node's front end (src/filter_node.cpp:236-273): NaN removal + range gate, 0.5 m pcl::VoxelGrid, statistical outlier
// removal -- the restatements of thirdparty_math.h chained.  in: n x 4 floats (x y z intensity); out: up to n x 4;
// counts[0..2] = points after the gate, after the downsample, after the outlier removal; dist (may be null): the per-point
// mean neighbour distances of the downsampled cloud.  Returns counts[2].
size_t oracle_filter_scan(const float* in, size_t n, double r_min, double r_max, float leaf, int mean_k, double std_mul, float* out,
                          size_t* counts, float* dist) {
  std::vector<uint32_t> keep;
  range_gate(in, n, 4, r_min, r_max, keep);
  std::vector<float> gated(keep.size() * 4);
  for (size_t i = 0; i < keep.size(); ++i) std::memcpy(&gated[i * 4], in + (size_t)keep[i] * 4, 16);
  counts[0] = keep.size();
  std::vector<float> ds;
  voxelgrid_centroids(gated.data(), keep.size(), 4, 3, leaf, ds);
  counts[1] = ds.size() / 4;
  std::vector<float> d;
  sor_filter(ds.data(), ds.size() / 4, 4, mean_k, std_mul, keep, &d);
  if (dist && !d.empty()) std::memcpy(dist, d.data(), d.size() * sizeof(float));
  for (size_t i = 0; i < keep.size(); ++i) std::memcpy(out + i * 4, &ds[(size_t)keep[i] * 4], 16);
  counts[2] = keep.size();
  return keep.size();
}
int oracle_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
