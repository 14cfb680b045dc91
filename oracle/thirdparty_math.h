// thirdparty_math.h -- the ONE restatement of the arithmetic that the reference path takes from
// un-vendored third-party libraries (Eigen 3.3 / PCL 1.7, versions inferred -- SURVEY.md appendix B).
//
// TEST INFRASTRUCTURE ONLY (oracle/).  Shared by
//   * oracle/ndt_oracle.cpp          -- the CPU restatement of the reference path, and
//   * oracle/ref_stubs/Eigen/Dense   -- the stand-in headers against which the reference's OWN
//     sources (/root/reference/xchu_mapping/include/pclomp/*) are compiled into oracle/_ref/,
// so that the two differ only in who wrote the NDT code (the reference's authors vs the
// restatement), never in the linear-algebra primitives underneath.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <vector>

namespace tpm {

// ----------------------------------------------------------------------------------------
// small dense helpers (double)
// ----------------------------------------------------------------------------------------
struct M3 { double a[3][3]; };

// Eigen compute_inverse_size3: inv(j,i) = cofactor<i,j> * (1/det); det from column 0.
inline double cof3(const M3& m, int i, int j) {
  const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return m.a[i1][j1] * m.a[i2][j2] - m.a[i1][j2] * m.a[i2][j1];
}
inline M3 inverse3(const M3& m) {
  const double c00 = cof3(m, 0, 0), c10 = cof3(m, 1, 0), c20 = cof3(m, 2, 0);
  const double det = (c00 * m.a[0][0] + c10 * m.a[1][0]) + c20 * m.a[2][0];
  const double invdet = 1.0 / det;
  M3 r;
  r.a[0][0] = c00 * invdet; r.a[0][1] = c10 * invdet; r.a[0][2] = c20 * invdet;
  for (int j = 1; j < 3; ++j)
    for (int i = 0; i < 3; ++i) r.a[j][i] = cof3(m, i, j) * invdet;
  return r;
}
inline M3 mul3(const M3& x, const M3& y) {
  M3 r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      r.a[i][j] = (x.a[i][0] * y.a[0][j] + x.a[i][1] * y.a[1][j]) + x.a[i][2] * y.a[2][j];
  return r;
}

// Symmetric 3x3 eigen-decomposition (reads the LOWER triangle like Eigen's
// SelfAdjointEigenSolver, VGC_I:333-335).  Cyclic Jacobi; eigenvalues ascending,
// eigenvectors in the columns of V.
inline void eigh3(const M3& in, double w[3], M3& V) {
  double A[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j <= i; ++j) A[i][j] = A[j][i] = in.a[i][j];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) V.a[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 64; ++sweep) {
    const double off = A[0][1] * A[0][1] + A[0][2] * A[0][2] + A[1][2] * A[1][2];
    const double dia = A[0][0] * A[0][0] + A[1][1] * A[1][1] + A[2][2] * A[2][2];
    if (off <= 1e-34 * dia || off == 0.0) break;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        if (A[p][q] == 0.0) continue;
        const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
        const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
        const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 3; ++k) {  // A <- A J
          const double akp = A[k][p], akq = A[k][q];
          A[k][p] = c * akp - s * akq;
          A[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; ++k) {  // A <- J^T A
          const double apk = A[p][k], aqk = A[q][k];
          A[p][k] = c * apk - s * aqk;
          A[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; ++k) {  // V <- V J
          const double vkp = V.a[k][p], vkq = V.a[k][q];
          V.a[k][p] = c * vkp - s * vkq;
          V.a[k][q] = s * vkp + c * vkq;
        }
      }
  }
  int ord[3] = {0, 1, 2};
  double d[3] = {A[0][0], A[1][1], A[2][2]};
  std::sort(ord, ord + 3, [&](int x, int y) { return d[x] < d[y]; });
  M3 Vs;
  for (int k = 0; k < 3; ++k) {
    w[k] = d[ord[k]];
    for (int i = 0; i < 3; ++i) Vs.a[i][k] = V.a[i][ord[k]];
  }
  V = Vs;
}

// x = pinv(H) * b through a one-sided (Hestenes) Jacobi SVD, standing in for
// Eigen::JacobiSVD<Matrix6d>(H, FullU|FullV).solve(b)  (NDT_I:119-121).
inline void svd_solve6(const double H[36] /*row-major*/, const double b[6], double x[6]) {
  double U[6][6], V[6][6];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) { U[i][j] = H[i * 6 + j]; V[i][j] = (i == j) ? 1.0 : 0.0; }
  for (int sweep = 0; sweep < 60; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < 5; ++p)
      for (int q = p + 1; q < 6; ++q) {
        double alpha = 0, beta = 0, gamma = 0;
        for (int k = 0; k < 6; ++k) {
          alpha += U[k][p] * U[k][p];
          beta += U[k][q] * U[k][q];
          gamma += U[k][p] * U[k][q];
        }
        if (gamma == 0.0 || std::fabs(gamma) <= 1e-16 * std::sqrt(alpha * beta)) continue;
        rotated = true;
        const double zeta = (beta - alpha) / (2.0 * gamma);
        const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
        const double c = 1.0 / std::sqrt(1.0 + t * t), s = c * t;
        for (int k = 0; k < 6; ++k) {
          const double up = U[k][p], uq = U[k][q];
          U[k][p] = c * up - s * uq;
          U[k][q] = s * up + c * uq;
          const double vp = V[k][p], vq = V[k][q];
          V[k][p] = c * vp - s * vq;
          V[k][q] = s * vp + c * vq;
        }
      }
    if (!rotated) break;
  }
  double sig[6], smax = 0;
  for (int j = 0; j < 6; ++j) {
    double n2 = 0;
    for (int k = 0; k < 6; ++k) n2 += U[k][j] * U[k][j];
    sig[j] = std::sqrt(n2);
    smax = std::max(smax, sig[j]);
  }
  const double thresh = 6.0 * std::numeric_limits<double>::epsilon() * smax;
  for (int i = 0; i < 6; ++i) x[i] = 0;
  for (int j = 0; j < 6; ++j) {
    if (!(sig[j] > thresh)) continue;
    double ub = 0;  // (u_j . b) with u_j = U[:,j]/sig
    for (int k = 0; k < 6; ++k) ub += U[k][j] * b[k];
    const double coef = ub / (sig[j] * sig[j]);
    for (int i = 0; i < 6; ++i) x[i] += V[i][j] * coef;
  }
}

// ----------------------------------------------------------------------------------------
// float transform helpers (Eigen::Translation * AngleAxis * AngleAxis * AngleAxis)
// ----------------------------------------------------------------------------------------
inline float sinf_def(float a) { return (float)std::sin((double)a); }
inline float cosf_def(float a) { return (float)std::cos((double)a); }

// Eigen AngleAxis<float>::toRotationMatrix() (Rodrigues form, Eigen's operation order).
inline void angle_axis_matrix_v(float angle, const float ax[3], float R[3][3]) {
  const float s = sinf_def(angle), c = cosf_def(angle);
  const float sin_axis[3] = {s * ax[0], s * ax[1], s * ax[2]};
  const float one_c = 1.f - c;
  const float cos1_axis[3] = {one_c * ax[0], one_c * ax[1], one_c * ax[2]};
  float tmp;
  tmp = cos1_axis[0] * ax[1]; R[0][1] = tmp - sin_axis[2]; R[1][0] = tmp + sin_axis[2];
  tmp = cos1_axis[0] * ax[2]; R[0][2] = tmp + sin_axis[1]; R[2][0] = tmp - sin_axis[1];
  tmp = cos1_axis[1] * ax[2]; R[1][2] = tmp - sin_axis[0]; R[2][1] = tmp + sin_axis[0];
  for (int i = 0; i < 3; ++i) R[i][i] = cos1_axis[i] * ax[i] + c;
}
// ... for a unit coordinate axis.
inline void angle_axis_matrix(float angle, int axis_id, float R[3][3]) {
  float ax[3] = {0.f, 0.f, 0.f};
  ax[axis_id] = 1.f;
  angle_axis_matrix_v(angle, ax, R);
}
inline void mul3f(const float A[3][3], const float B[3][3], float C[3][3]) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C[i][j] = (A[i][0] * B[0][j] + A[i][1] * B[1][j]) + A[i][2] * B[2][j];
}
// NDT_H:210-217 / NDT_I:138-143, 815-819.  T column-major 4x4 float.
inline void pose_to_matrix(const double p[6], float T[16]) {
  float Rx[3][3], Ry[3][3], Rz[3][3], Rxy[3][3], R[3][3];
  angle_axis_matrix((float)p[3], 0, Rx);
  angle_axis_matrix((float)p[4], 1, Ry);
  angle_axis_matrix((float)p[5], 2, Rz);
  mul3f(Rx, Ry, Rxy);
  mul3f(Rxy, Rz, R);
  for (int c = 0; c < 3; ++c) {
    for (int r = 0; r < 3; ++r) T[c * 4 + r] = R[r][c];
    T[c * 4 + 3] = 0.f;
  }
  T[12] = (float)p[0]; T[13] = (float)p[1]; T[14] = (float)p[2]; T[15] = 1.f;
}
// Eigen 3.3 MatrixBase::eulerAngles(0,1,2) on the float rotation block (NDT_I:102).
inline void euler_012(const float T[16], float out[3]) {
  auto m = [&](int r, int c) { return T[c * 4 + r]; };
  const float pi = (float)M_PI;
  float r0 = std::atan2(m(1, 2), m(2, 2));
  // c2 = Vector2(coeff(i,i), coeff(i,j)).norm()
  const float c2 = std::sqrt(m(0, 0) * m(0, 0) + m(0, 1) * m(0, 1));
  float r1;
  if (r0 > 0.f) {
    if (r0 > 0.f) r0 -= pi; else r0 += pi;
    r1 = std::atan2(-m(0, 2), -c2);
  } else {
    r1 = std::atan2(-m(0, 2), c2);
  }
  const float s1 = std::sin(r0), c1 = std::cos(r0);
  const float r2 = std::atan2(s1 * m(2, 0) - c1 * m(1, 0), c1 * m(1, 1) - s1 * m(2, 1));
  out[0] = -r0; out[1] = -r1; out[2] = -r2;
}
// pcl::VoxelGrid<PointXYZI>::applyFilter (PCL 1.7 filters/impl/voxel_grid.hpp) with downsample_all_data_ = true
// (the default), as odom_node uses it for downSizeFilterLocalmap (src/odom_node.cpp:98, 334-335):
//   min/max over the cloud; min_b = floor(min * inv_leaf), max_b likewise, inv_leaf = 1.0f / leaf;
//   if dx*dy*dz (from (max - min) * inv_leaf + 1) overflows int32: warn and return the input unchanged;
//   leaf index of a point: (floor(x*inv) - min_b.x) + (floor(y*inv) - min_b.y) * dx + (floor(z*inv) - min_b.z) * dx * dy;
//   points sorted by leaf index (std::sort on the index alone; DECISION: input order among equals, i.e. a stable sort);
//   one output point per occupied leaf in ascending index: float sums of x, y, z, intensity in that order, each divided by
//   float(count) (DECISION: true division, Eigen >= 3.3 `/=`).
// in: n records of `stride` floats (x y z at 0..2, intensity at `ioff`); out: packed x y z intensity.
inline void voxelgrid_centroids(const float* in, size_t n, size_t stride, size_t ioff, float leaf, std::vector<float>& out) {
  out.clear();
  if (n == 0) return;
  const float inv = 1.0f / leaf;
  float mn[3] = {3.402823466e38f, 3.402823466e38f, 3.402823466e38f}, mx[3] = {-3.402823466e38f, -3.402823466e38f, -3.402823466e38f};
  for (size_t i = 0; i < n; ++i)
    for (int a = 0; a < 3; ++a) {
      const float v = in[i * stride + a];
      mn[a] = v < mn[a] ? v : mn[a];
      mx[a] = v > mx[a] ? v : mx[a];
    }
  const int64_t dx = (int64_t)((mx[0] - mn[0]) * inv) + 1, dy = (int64_t)((mx[1] - mn[1]) * inv) + 1, dz = (int64_t)((mx[2] - mn[2]) * inv) + 1;
  if (dx * dy * dz > (int64_t)std::numeric_limits<int32_t>::max()) {
    for (size_t i = 0; i < n; ++i) { out.push_back(in[i * stride]); out.push_back(in[i * stride + 1]); out.push_back(in[i * stride + 2]); out.push_back(in[i * stride + ioff]); }
    return;
  }
  int min_b[3], max_b[3];
  for (int a = 0; a < 3; ++a) { min_b[a] = (int)std::floor(mn[a] * inv); max_b[a] = (int)std::floor(mx[a] * inv); }
  const int div0 = max_b[0] - min_b[0] + 1, div1 = max_b[1] - min_b[1] + 1;
  const int mul[3] = {1, div0, div0 * div1};
  std::vector<std::pair<unsigned, unsigned>> iv(n);   // (leaf index, point index)
  for (size_t i = 0; i < n; ++i) {
    const int ijk0 = (int)(std::floor(in[i * stride] * inv) - (float)min_b[0]);
    const int ijk1 = (int)(std::floor(in[i * stride + 1] * inv) - (float)min_b[1]);
    const int ijk2 = (int)(std::floor(in[i * stride + 2] * inv) - (float)min_b[2]);
    iv[i] = std::make_pair((unsigned)(ijk0 * mul[0] + ijk1 * mul[1] + ijk2 * mul[2]), (unsigned)i);
  }
  std::stable_sort(iv.begin(), iv.end(), [](const std::pair<unsigned, unsigned>& a, const std::pair<unsigned, unsigned>& b) { return a.first < b.first; });
  size_t b = 0;
  while (b < n) {
    size_t e = b + 1;
    while (e < n && iv[e].first == iv[b].first) ++e;
    const float* p0 = in + (size_t)iv[b].second * stride;
    float c[4] = {p0[0], p0[1], p0[2], p0[ioff]};
    for (size_t q = b + 1; q < e; ++q) {
      const float* p = in + (size_t)iv[q].second * stride;
      c[0] += p[0]; c[1] += p[1]; c[2] += p[2]; c[3] += p[ioff];
    }
    const float cnt = (float)(e - b);
    for (int a = 0; a < 4; ++a) out.push_back(c[a] / cnt);
    b = e;
  }
}

// ---- filter_node front end (reference src/filter_node.cpp:236-273) --------------------------------------------------
// pcl::removeNaNFromPointCloud + the range gate of filter_node.cpp:242-247:
//   r = std::sqrt(pow(point.x, 2.0) + pow(point.y, 2.0)) in double; keep iff 1.0 < r && r < 60 (bounds passed in).
inline void range_gate(const float* in, size_t n, size_t stride, double r_min, double r_max, std::vector<uint32_t>& keep) {
  keep.clear();
  for (size_t i = 0; i < n; ++i) {
    const float x = in[i * stride], y = in[i * stride + 1], z = in[i * stride + 2];
    if (!std::isfinite(x) || !std::isfinite(y) || !std::isfinite(z)) continue;
    const double r = std::sqrt((double)x * (double)x + (double)y * (double)y);
    if (r_min < r && r < r_max) keep.push_back((uint32_t)i);
  }
}
// pcl::StatisticalOutlierRemoval<PointXYZI>::applyFilterIndices (PCL 1.7 filters/impl/statistical_outlier_removal.hpp),
// setMeanK(mean_k), setStddevMulThresh(std_mul), negative_ = false (filter_node.cpp:254-262):
//   per point: the mean_k + 1 nearest neighbours (the point itself is neighbour 0), squared float distances
//   ((dx^2 + dy^2) + dz^2) ascending; dist_sum (double) += sqrt(nn_dists[k]) for k = 1..mean_k
//   (DECISION: the unqualified sqrt binds ::sqrt(double), like the exp of NDT_I:507); distances[i] = float(dist_sum / mean_k);
//   sum += distances[i]; sq_sum += distances[i] * distances[i] (a FLOAT product added to a double), in point order;
//   mean = sum / n; variance = (sq_sum - sum * sum / n) / (n - 1); threshold = mean + std_mul * sqrt(variance);
//   a point is removed iff distances[i] > threshold.
// DECISION: a cloud with fewer than mean_k + 1 points is returned unchanged (PCL reads stale distances there).
// xyz: n records of `stride` floats.  Exact k-NN by exhaustive search (test sizes only).
inline void sor_mean_distances(const float* xyz, size_t n, size_t stride, int mean_k, std::vector<float>& distances) {
  distances.assign(n, 0.f);
  if (n < (size_t)mean_k + 1) return;
  const long nn = (long)n;
#pragma omp parallel
  {
    std::vector<float> d2(n);
#pragma omp for schedule(static)
    for (long i = 0; i < nn; ++i) {
      const float qx = xyz[i * stride], qy = xyz[i * stride + 1], qz = xyz[i * stride + 2];
      for (size_t j = 0; j < n; ++j) {
        const float ex = qx - xyz[j * stride], ey = qy - xyz[j * stride + 1], ez = qz - xyz[j * stride + 2];
        d2[j] = (ex * ex + ey * ey) + ez * ez;
      }
      std::partial_sort(d2.begin(), d2.begin() + mean_k + 1, d2.end());
      double dist_sum = 0.0;
      for (int k = 1; k < mean_k + 1; ++k) dist_sum += std::sqrt((double)d2[k]);
      distances[i] = (float)(dist_sum / mean_k);
    }
  }
}
inline double sor_threshold(const std::vector<float>& distances, double std_mul) {
  double sum = 0, sq_sum = 0;
  for (size_t i = 0; i < distances.size(); ++i) {
    sum += distances[i];
    sq_sum += distances[i] * distances[i];
  }
  const double nv = (double)distances.size();
  const double mean = sum / nv;
  const double variance = (sq_sum - sum * sum / nv) / (nv - 1);
  return mean + std_mul * std::sqrt(variance);
}
inline void sor_filter(const float* xyz, size_t n, size_t stride, int mean_k, double std_mul, std::vector<uint32_t>& keep,
                       std::vector<float>* distances_out = nullptr) {
  keep.clear();
  if (n < (size_t)mean_k + 1) {
    for (size_t i = 0; i < n; ++i) keep.push_back((uint32_t)i);
    return;
  }
  std::vector<float> dist;
  sor_mean_distances(xyz, n, stride, mean_k, dist);
  const double thr = sor_threshold(dist, std_mul);
  for (size_t i = 0; i < n; ++i)
    if (!(dist[i] > thr)) keep.push_back((uint32_t)i);
  if (distances_out) *distances_out = dist;
}

}  // namespace tpm
