// sc_oracle.cpp -- CPU restatement of the reference's ScanContext path:
// xchu_mapping/include/scancontext/Scancontext.cpp (= SC) :23-37 xy2theta, :40-61 circshift, :71-93 distDirectSC,
// :96-115 fastAlignUsingVkey, :118-148 distanceBtnScanContext, :151-196 makeScancontext, :199-228 ring / sector keys,
// :331-423 detectLoopClosureID; constants Scancontext.h:83-103.
//
// TEST INFRASTRUCTURE ONLY -- imported by tests/ (and scripts timing it as the CPU baseline of this row), never by the
// product.  PARITY PIN: tests/test_scancontext.py requires every function here to equal, bit for bit, the reference's own
// Scancontext.cpp compiled unmodified over the stand-in Eigen / PCL headers (oracle/_ref/libsc_ref.so, oracle/ref_sc_shim.cpp).
// Third-party arithmetic as restated there: Eigen reductions (norm, dot, mean) sequential with ascending index; the
// unqualified sqrt / atan of SC:28-36, 172 bind to the double functions (what g++ resolves with <cmath> alone); the ring-key
// kNN is nanoflann's L2_Simple_Adaptor distance (float, sequential) with an exact search -- the reference's kd-tree is exact
// too (eps 0), so only the order among equidistant keys could differ.
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <utility>
#include <vector>

namespace {

struct Params {
  int num_ring = 20, num_sector = 60;
  double max_radius = 80.0, lidar_height = 2.0, search_ratio = 0.1;
  int exclude_recent = 30, candidates = 3, tree_period = 30;
  double dist_thres = 0.2;
};

// SC:23-37.  (180/M_PI) is a double constant expression; atan takes the float quotient widened to double; the double
// result narrows to float on return.
float xy2theta(float x, float y) {
  if ((x >= 0) & (y >= 0)) return (float)((180 / M_PI) * std::atan((double)(y / x)));
  if ((x < 0) & (y >= 0)) return (float)(180 - ((180 / M_PI) * std::atan((double)(y / (-x)))));
  if ((x < 0) & (y < 0)) return (float)(180 + ((180 / M_PI) * std::atan((double)(y / x))));
  if ((x >= 0) & (y < 0)) return (float)(360 - ((180 / M_PI) * std::atan((double)((-y) / x))));
  return 0.f;   // NaN input: the reference falls off the end of the function (undefined); not exercised
}

// SC:151-196.  desc column-major rows x cols (Eigen default).
void make_desc(const Params& P, const float* pts, size_t n, size_t stride, double* desc) {
  const int R = P.num_ring, S = P.num_sector;
  const double NO_POINT = -1000;
  for (int k = 0; k < R * S; ++k) desc[k] = NO_POINT;
  for (size_t i = 0; i < n; ++i) {
    const float px = pts[i * stride], py = pts[i * stride + 1];
    const float pz = (float)((double)pts[i * stride + 2] + P.lidar_height);      // SC:168: float = float + double
    const float azim_range = (float)std::sqrt((double)(px * px + py * py));     // SC:171
    const float azim_angle = xy2theta(px, py);
    if ((double)azim_range > P.max_radius) continue;                             // SC:175
    const int ring_idx = std::max(std::min(R, int(std::ceil(((double)azim_range / P.max_radius) * R))), 1);
    const int sctor_idx = std::max(std::min(S, int(std::ceil(((double)azim_angle / 360.0) * S))), 1);
    double& d = desc[(size_t)(sctor_idx - 1) * R + (ring_idx - 1)];
    if (d < (double)pz) d = (double)pz;                                          // SC:182-183
  }
  for (int k = 0; k < R * S; ++k)
    if (desc[k] == NO_POINT) desc[k] = 0;
}

// SC:199-228: Eigen mean() = sum() / size(), sequential
void make_keys(const Params& P, const double* desc, double* ring, double* sector) {
  const int R = P.num_ring, S = P.num_sector;
  for (int r = 0; r < R; ++r) {
    double s = desc[r];
    for (int c = 1; c < S; ++c) s += desc[(size_t)c * R + r];
    ring[r] = s / (double)S;
  }
  for (int c = 0; c < S; ++c) {
    double s = desc[(size_t)c * R];
    for (int r = 1; r < R; ++r) s += desc[(size_t)c * R + r];
    sector[c] = s / (double)R;
  }
}

// column `c` of circshift(m, shift) is column (c - shift) mod S of m (SC:53-57)
inline int src_col(int c, int shift, int S) { return ((c - shift) % S + S) % S; }

// SC:71-93 on sc1 and circshift(sc2, shift)
double dist_direct(const Params& P, const double* sc1, const double* sc2, int shift) {
  const int R = P.num_ring, S = P.num_sector;
  int num_eff_cols = 0;
  double sum_sector_similarity = 0;
  for (int c = 0; c < S; ++c) {
    const double* a = sc1 + (size_t)c * R;
    const double* b = sc2 + (size_t)src_col(c, shift, S) * R;
    double na = a[0] * a[0], nb = b[0] * b[0], dot = a[0] * b[0];
    for (int r = 1; r < R; ++r) { na += a[r] * a[r]; nb += b[r] * b[r]; dot += a[r] * b[r]; }
    const double n1 = std::sqrt(na), n2 = std::sqrt(nb);
    if ((n1 == 0) | (n2 == 0)) continue;
    const double sector_similarity = dot / (n1 * n2);
    sum_sector_similarity = sum_sector_similarity + sector_similarity;
    num_eff_cols = num_eff_cols + 1;
  }
  const double sc_sim = sum_sector_similarity / num_eff_cols;
  return 1.0 - sc_sim;
}

// SC:96-115
int fast_align(const Params& P, const double* vkey1, const double* vkey2) {
  const int S = P.num_sector;
  int argmin = 0;
  double min_norm = 10000000;
  for (int shift = 0; shift < S; ++shift) {
    double d0 = vkey1[0] - vkey2[src_col(0, shift, S)];
    double ss = d0 * d0;
    for (int c = 1; c < S; ++c) { const double d = vkey1[c] - vkey2[src_col(c, shift, S)]; ss += d * d; }
    const double cur = std::sqrt(ss);
    if (cur < min_norm) { argmin = shift; min_norm = cur; }
  }
  return argmin;
}

// SC:118-148
std::pair<double, int> distance(const Params& P, const double* sc1, const double* sc2) {
  const int S = P.num_sector;
  std::vector<double> r1(P.num_ring), r2(P.num_ring), v1(S), v2(S);
  make_keys(P, sc1, r1.data(), v1.data());
  make_keys(P, sc2, r2.data(), v2.data());
  const int a = fast_align(P, v1.data(), v2.data());
  const int SEARCH_RADIUS = (int)std::round(0.5 * P.search_ratio * S);
  std::vector<int> space{a};
  for (int ii = 1; ii < SEARCH_RADIUS + 1; ++ii) {
    space.push_back((a + ii + S) % S);
    space.push_back((a - ii + S) % S);
  }
  std::sort(space.begin(), space.end());
  int argmin_shift = 0;
  double min_sc_dist = 10000000;
  for (int num_shift : space) {
    const double cur = dist_direct(P, sc1, sc2, num_shift);
    if (cur < min_sc_dist) { argmin_shift = num_shift; min_sc_dist = cur; }
  }
  return std::make_pair(min_sc_dist, argmin_shift);
}

// nanoflann L2_Simple_Adaptor (metric_L2_Simple is what KDTreeVectorOfVectorsAdaptor's default metric_L2 ... ) -- see below
struct Manager {
  Params P;
  std::vector<std::vector<double>> desc, ring, sector;
  std::vector<std::vector<float>> invkey;
  int tree_counter = 0;
  size_t tree_size = 0;       // number of ring keys in the tree built last (SC:358-361)
  bool tree_built = false;
};

// squared L2 between float keys the way nanoflann's L2_Adaptor::evalMetric accumulates it (nanoflann.hpp: groups of four
// differences summed into the float result, then the remainder one by one)
float key_dist2(const float* a, const float* b, int dim) {
  float result = 0.f;
  int k = 0;
  for (; k + 4 <= dim; k += 4) {
    const float d0 = a[k] - b[k], d1 = a[k + 1] - b[k + 1], d2 = a[k + 2] - b[k + 2], d3 = a[k + 3] - b[k + 3];
    result += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
  }
  for (; k < dim; ++k) { const float d = a[k] - b[k]; result += d * d; }
  return result;
}

}  // namespace

extern "C" {

void* sco_create() { return new Manager(); }
void sco_destroy(void* h) { delete static_cast<Manager*>(h); }
void sco_set_dist_thres(void* h, double t) { static_cast<Manager*>(h)->P.dist_thres = t; }
float sco_xy2theta(float x, float y) { return xy2theta(x, y); }
void sco_make(void* h, const float* pts, size_t n, size_t stride_floats, double* desc) { make_desc(static_cast<Manager*>(h)->P, pts, n, stride_floats, desc); }
void sco_keys(void* h, const double* desc, double* ring, double* sector) { make_keys(static_cast<Manager*>(h)->P, desc, ring, sector); }
double sco_dist_direct(void* h, const double* d1, const double* d2) { return dist_direct(static_cast<Manager*>(h)->P, d1, d2, 0); }
int sco_fast_align(void* h, const double* k1, const double* k2) { return fast_align(static_cast<Manager*>(h)->P, k1, k2); }
void sco_distance(void* h, const double* d1, const double* d2, double* dist, int* shift) {
  std::pair<double, int> r = distance(static_cast<Manager*>(h)->P, d1, d2);
  *dist = r.first;
  *shift = r.second;
}
// SC:249-260 makeAndSaveScancontextAndKeys
void sco_add(void* h, const float* pts, size_t n, size_t stride_floats) {
  Manager* m = static_cast<Manager*>(h);
  const int R = m->P.num_ring, S = m->P.num_sector;
  std::vector<double> d((size_t)R * S), rk(R), sk(S);
  make_desc(m->P, pts, n, stride_floats, d.data());
  make_keys(m->P, d.data(), rk.data(), sk.data());
  std::vector<float> ik(R);
  for (int r = 0; r < R; ++r) ik[r] = (float)rk[r];     // eig2stdvec (SC:64-68): double -> float
  m->desc.push_back(d); m->ring.push_back(rk); m->sector.push_back(sk); m->invkey.push_back(ik);
}
size_t sco_size(void* h) { return static_cast<Manager*>(h)->desc.size(); }
void sco_get(void* h, size_t i, double* desc, double* ring, double* sector, float* invkey) {
  Manager* m = static_cast<Manager*>(h);
  std::copy(m->desc[i].begin(), m->desc[i].end(), desc);
  std::copy(m->ring[i].begin(), m->ring[i].end(), ring);
  std::copy(m->sector[i].begin(), m->sector[i].end(), sector);
  std::copy(m->invkey[i].begin(), m->invkey[i].end(), invkey);
}
// SC:331-423 detectLoopClosureID.  Also returns the minimum distance, its candidate and the size of the searched set.
void sco_detect(void* h, int* loop_id, float* yaw_diff_rad, double* min_dist_out, int* nn_idx_out, int* tree_size_out) {
  Manager* m = static_cast<Manager*>(h);
  const Params& P = m->P;
  *loop_id = -1; *yaw_diff_rad = 0.f; *min_dist_out = 10000000; *nn_idx_out = 0; *tree_size_out = (int)m->tree_size;
  if ((int)m->invkey.size() < P.exclude_recent + 1) return;                        // SC:343-347
  if (m->tree_counter % P.tree_period == 0) {                                      // SC:350-361
    m->tree_size = m->invkey.size() - P.exclude_recent;
    m->tree_built = true;
  }
  m->tree_counter = m->tree_counter + 1;
  *tree_size_out = (int)m->tree_size;
  const std::vector<float>& q = m->invkey.back();
  const std::vector<double>& qd = m->desc.back();
  // exact k nearest ring keys, ascending distance (nanoflann KNNResultSet keeps them sorted; ties: lower index first here)
  std::vector<std::pair<float, int>> all(m->tree_size);
  for (size_t i = 0; i < m->tree_size; ++i) all[i] = std::make_pair(key_dist2(q.data(), m->invkey[i].data(), P.num_ring), (int)i);
  const size_t k = std::min((size_t)P.candidates, all.size());
  std::partial_sort(all.begin(), all.begin() + k, all.end());
  double min_dist = 10000000;
  int nn_align = 0, nn_idx = 0;
  for (int c = 0; c < P.candidates; ++c) {
    // fewer keys in the tree than candidates: nanoflann leaves the remaining result slots at their initial value (index 0
    // of a zero-initialised vector, SC:372-373), so descriptor 0 is compared again
    const int idx = (size_t)c < k ? all[c].second : 0;
    std::pair<double, int> r = distance(P, qd.data(), m->desc[idx].data());
    if (r.first < min_dist) { min_dist = r.first; nn_align = r.second; nn_idx = idx; }
  }
  if (min_dist < P.dist_thres) *loop_id = nn_idx;
  const double unit = 360.0 / double(P.num_sector);
  const float deg = (float)(nn_align * unit);                                       // deg2rad(float): SC:17-20
  *yaw_diff_rad = (float)((double)deg * M_PI / 180.0);
  *min_dist_out = min_dist; *nn_idx_out = nn_idx;
}

}  // extern "C"
