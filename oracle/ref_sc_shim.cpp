// ref_sc_shim.cpp -- C entry points around the reference's OWN ScanContext sources, compiled unmodified from
// /root/reference/xchu_mapping/include/scancontext/{Scancontext.h, Scancontext.cpp, KDTreeVectorOfVectorsAdaptor.h,
// nanoflann.hpp, tic_toc.h} into oracle/_ref/libsc_ref.so (recipe: oracle/Makefile target `ref`).
//
// TEST INFRASTRUCTURE ONLY: tests/ use it to pin oracle/sc_oracle.cpp against the reference's code (SURVEY.md section 8(f)
// rank 4).  Eigen / PCL / OpenCV / cv_bridge are not in this image: Eigen and PCL are the stand-in headers of
// oracle/ref_stubs/ (eager, sequential reductions), the OpenCV / cv_bridge / pcl_conversions headers Scancontext.h names
// are empty stand-ins (nothing of them is used on this path).  nanoflann and the kd-tree adaptor are the reference's own.
#include <cstring>

#include "scancontext/Scancontext.cpp"

namespace {
typedef pcl::PointCloud<SCPointType> Cloud;
void fill(Cloud& c, const float* pts, size_t n, size_t stride_floats) {
  c.points.resize(n);
  for (size_t i = 0; i < n; ++i) {
    c.points[i].x = pts[i * stride_floats];
    c.points[i].y = pts[i * stride_floats + 1];
    c.points[i].z = pts[i * stride_floats + 2];
  }
  c.width = (uint32_t)n;
  c.height = 1;
}
MatrixXd from_buf(const double* d, int rows, int cols) {
  MatrixXd m = MatrixXd::Zero(rows, cols);
  for (int c = 0; c < cols; ++c)
    for (int r = 0; r < rows; ++r) m(r, c) = d[c * rows + r];
  return m;
}
void to_buf(const MatrixXd& m, double* d) {
  for (int c = 0; c < m.cols(); ++c)
    for (int r = 0; r < m.rows(); ++r) d[c * m.rows() + r] = m(r, c);
}
}  // namespace

extern "C" {

void* ref_sc_create() { return new SCManager(); }
void ref_sc_destroy(void* h) { delete static_cast<SCManager*>(h); }
void ref_sc_constants(void* h, int* ring, int* sector, double* max_radius, double* lidar_height, double* search_ratio, int* exclude_recent,
                      int* candidates, double* dist_thres, int* tree_period) {
  SCManager* m = static_cast<SCManager*>(h);
  *ring = m->PC_NUM_RING; *sector = m->PC_NUM_SECTOR; *max_radius = m->PC_MAX_RADIUS; *lidar_height = m->LIDAR_HEIGHT;
  *search_ratio = m->SEARCH_RATIO; *exclude_recent = m->NUM_EXCLUDE_RECENT; *candidates = m->NUM_CANDIDATES_FROM_TREE;
  *dist_thres = m->SC_DIST_THRES; *tree_period = m->TREE_MAKING_PERIOD_;
}
// desc: PC_NUM_RING x PC_NUM_SECTOR doubles, column-major (Eigen's default)
void ref_sc_make(void* h, const float* pts, size_t n, size_t stride_floats, double* desc) {
  Cloud c;
  fill(c, pts, n, stride_floats);
  MatrixXd d = static_cast<SCManager*>(h)->makeScancontext(c);
  to_buf(d, desc);
}
void ref_sc_keys(void* h, const double* desc, int rows, int cols, double* ring, double* sector) {
  SCManager* m = static_cast<SCManager*>(h);
  MatrixXd d = from_buf(desc, rows, cols);
  MatrixXd rk = m->makeRingkeyFromScancontext(d), sk = m->makeSectorkeyFromScancontext(d);
  for (int r = 0; r < rows; ++r) ring[r] = rk(r, 0);
  for (int c = 0; c < cols; ++c) sector[c] = sk(0, c);
}
double ref_sc_dist_direct(void* h, const double* d1, const double* d2, int rows, int cols) {
  MatrixXd a = from_buf(d1, rows, cols), b = from_buf(d2, rows, cols);
  return static_cast<SCManager*>(h)->distDirectSC(a, b);
}
int ref_sc_fast_align(void* h, const double* k1, const double* k2, int cols) {
  MatrixXd a = from_buf(k1, 1, cols), b = from_buf(k2, 1, cols);
  return static_cast<SCManager*>(h)->fastAlignUsingVkey(a, b);
}
void ref_sc_distance(void* h, const double* d1, const double* d2, int rows, int cols, double* dist, int* shift) {
  MatrixXd a = from_buf(d1, rows, cols), b = from_buf(d2, rows, cols);
  std::pair<double, int> r = static_cast<SCManager*>(h)->distanceBtnScanContext(a, b);
  *dist = r.first;
  *shift = r.second;
}
void ref_sc_add(void* h, const float* pts, size_t n, size_t stride_floats) {
  Cloud c;
  fill(c, pts, n, stride_floats);
  static_cast<SCManager*>(h)->makeAndSaveScancontextAndKeys(c);
}
void ref_sc_detect(void* h, int* loop_id, float* yaw_diff_rad) {
  std::pair<int, float> r = static_cast<SCManager*>(h)->detectLoopClosureID();
  *loop_id = r.first;
  *yaw_diff_rad = r.second;
}
size_t ref_sc_size(void* h) { return static_cast<SCManager*>(h)->polarcontexts_.size(); }
void ref_sc_get(void* h, size_t i, double* desc, double* ring, double* sector, float* invkey) {
  SCManager* m = static_cast<SCManager*>(h);
  to_buf(m->polarcontexts_[i], desc);
  for (int r = 0; r < m->PC_NUM_RING; ++r) { ring[r] = m->polarcontext_invkeys_[i](r, 0); invkey[r] = m->polarcontext_invkeys_mat_[i][r]; }
  for (int c = 0; c < m->PC_NUM_SECTOR; ++c) sector[c] = m->polarcontext_vkeys_[i](0, c);
}
float ref_sc_xy2theta(float x, float y) { return xy2theta(x, y); }

}  // extern "C"
