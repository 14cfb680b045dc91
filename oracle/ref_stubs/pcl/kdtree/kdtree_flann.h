// Stand-in for <pcl/kdtree/kdtree_flann.h>.  TEST INFRASTRUCTURE ONLY -- see pcl/point_types.h.
// FLANN is not in this image; the exact kd-tree underneath is the reference's OWN vendored
// nanoflann (/root/reference/xchu_mapping/include/scancontext/nanoflann.hpp, on the include path of
// the oracle/_ref build).  Semantics: squared float L2 ((dx^2+dy^2)+dz^2),
// radius search includes d^2 < r^2 with r^2 = float(r*r), unlimited neighbours, ascending distance.
#pragma once
#include <pcl/point_types.h>
#include <scancontext/nanoflann.hpp>

namespace pcl {

template <class PointT>
class KdTreeFLANN {
  struct Adaptor {
    const std::vector<PointT>* pts = nullptr;
    size_t kdtree_get_point_count() const { return pts->size(); }
    float kdtree_get_pt(size_t i, size_t d) const { return (*pts)[i].data[d]; }
    template <class BBOX> bool kdtree_get_bbox(BBOX&) const { return false; }
  };
  typedef nanoflann::KDTreeSingleIndexAdaptor<nanoflann::L2_Simple_Adaptor<float, Adaptor>, Adaptor, 3, int> Tree;
  typename pcl::PointCloud<PointT>::ConstPtr cloud_;
  Adaptor ad_;
  std::unique_ptr<Tree> tree_;

 public:
  typedef boost::shared_ptr<KdTreeFLANN<PointT>> Ptr;
  void setInputCloud(const typename pcl::PointCloud<PointT>::ConstPtr& cloud, const IndicesConstPtr& = IndicesConstPtr()) {
    cloud_ = cloud;
    ad_.pts = &cloud_->points;
    tree_.reset(new Tree(3, ad_, nanoflann::KDTreeSingleIndexAdaptorParams(15)));
    tree_->buildIndex();
  }
  int nearestKSearch(const PointT& p, int k, std::vector<int>& idx, std::vector<float>& d2) const {
    idx.assign(k, 0); d2.assign(k, 0.f);
    if (!tree_ || cloud_->points.empty()) { idx.clear(); d2.clear(); return 0; }
    const float q[3] = {p.x, p.y, p.z};
    const size_t n = tree_->knnSearch(q, (size_t)k, idx.data(), d2.data());
    idx.resize(n); d2.resize(n);
    return (int)n;
  }
  int radiusSearch(const PointT& p, double radius, std::vector<int>& idx, std::vector<float>& d2, unsigned int max_nn = 0) const {
    idx.clear(); d2.clear();
    if (!tree_ || cloud_->points.empty()) return 0;
    const float q[3] = {p.x, p.y, p.z};
    std::vector<std::pair<int, float>> hits;
    nanoflann::SearchParams sp;
    sp.sorted = true;
    tree_->radiusSearch(q, static_cast<float>(radius * radius), hits, sp);
    if (max_nn && hits.size() > max_nn) hits.resize(max_nn);
    for (auto& h : hits) { idx.push_back(h.first); d2.push_back(h.second); }
    return (int)idx.size();
  }
};

}  // namespace pcl
