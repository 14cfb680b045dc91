// scancontext/Scancontext.h -- drop-in for the reference's header of the same path
// (xchu_mapping/include/scancontext/Scancontext.h:56-121) over the device implementation (ndt_b200/sc_c_api.h, libndt_b200.so).
// Same class name, same member names and return types for what pgo_node calls on its `SCManager scManager`:
//   setSCdistThres (src/pgo_node.cpp:58), makeAndSaveScancontextAndKeys (:236), detectLoopClosureID (:344),
// so those lines compile unchanged; examples/sc_sequence.cpp is compiled both over this header and over the reference's own
// Scancontext.cpp and must print the same loop ids and yaw differences (tests/test_cpp_shim.py).
// The descriptors live in HBM; the host keeps the reference's tree bookkeeping (Scancontext.cpp:350-363).
#pragma once
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include <ndt_b200/sc_c_api.h>
#include <pclomp/pcl_compat.h>

using SCPointType = pcl::PointXYZI;

class SCManager {
 public:
  explicit SCManager(int device = 0) {
    ndtb_sc_default_params(&p_);
    if (ndtb_sc_create(device, &p_, &s_)) throw std::runtime_error(std::string("ndtb_sc_create: ") + ndtb_sc_last_error(nullptr));
  }
  ~SCManager() { ndtb_sc_destroy(s_); }
  SCManager(const SCManager&) = delete;
  SCManager& operator=(const SCManager&) = delete;

  // Scancontext.cpp:262-265
  void setSCdistThres(double _new_thres) { p_.dist_thres = _new_thres; check(ndtb_sc_set_dist_thres(s_, _new_thres)); }

  // Scancontext.cpp:249-260
  void makeAndSaveScancontextAndKeys(pcl::PointCloud<SCPointType>& _scan_down) {
    const void* ptr = _scan_down.points.data();
    const size_t n = _scan_down.points.size();
    int id = 0;
    check(ndtb_sc_add_scans(s_, 1, &ptr, &n, sizeof(SCPointType), &id));
  }
  // many keyframes in one launch (what the device is for): re-describing a session, batched map loading
  void makeAndSaveScancontextAndKeys(std::vector<pcl::PointCloud<SCPointType>*>& scans) {
    std::vector<const void*> ptrs;
    std::vector<size_t> ns;
    for (auto* c : scans) { ptrs.push_back(c->points.data()); ns.push_back(c->points.size()); }
    int id = 0;
    check(ndtb_sc_add_scans(s_, (int)scans.size(), ptrs.data(), ns.data(), sizeof(SCPointType), &id));
  }

  // Scancontext.cpp:331-423: (nearest node index or -1, relative yaw in radians)
  std::pair<int, float> detectLoopClosureID(void) {
    const int n = ndtb_sc_size(s_);
    if (n < p_.num_exclude_recent + 1) return std::pair<int, float>(-1, 0.0f);            // :343-347
    if (tree_making_period_conter % p_.tree_making_period == 0) tree_size_ = n - p_.num_exclude_recent;   // :350-361
    tree_making_period_conter = tree_making_period_conter + 1;
    int q = n - 1, loop_id = -1;
    float yaw = 0.f;
    check(ndtb_sc_detect(s_, 1, &q, &tree_size_, &loop_id, &yaw, nullptr, nullptr));
    return std::pair<int, float>(loop_id, yaw);
  }

  size_t size() const { return (size_t)ndtb_sc_size(s_); }     // polarcontexts_.size()
  ndtb_sc* handle() { return s_; }

  int tree_making_period_conter = 0;   // sic, Scancontext.h:83

 private:
  void check(int rc) {
    if (rc) throw std::runtime_error(ndtb_sc_last_error(s_));
  }
  ndtb_sc* s_ = nullptr;
  ndtb_sc_params p_;
  int tree_size_ = 0;
};
