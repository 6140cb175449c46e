// Stand-in for pcl/kdtree/kdtree_flann.h: exact nearest neighbours by a pass over all points (the reference only
// asks for the single nearest position among a few hundred waypoints, sspp/src/pathplanner.cpp:628-645).
#pragma once
#include <algorithm>
#include <vector>
#include "../point_types.h"
namespace pcl {
template <class PointT> class KdTreeFLANN {
 public:
  void setInputCloud(const typename PointCloud<PointT>::ConstPtr& c) { cloud_ = c; }
  // <FLANN L2_Simple>: squared distance accumulated in float, x then y then z
  static float sq(const PointT& a, const PointT& b) {
    float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
    return (dx * dx + dy * dy) + dz * dz;
  }
  int nearestKSearch(const PointT& q, int k, std::vector<int>& idx, std::vector<float>& d2) const {
    std::vector<std::pair<float, int> > all;
    for (size_t i = 0; i < cloud_->points.size(); ++i) all.push_back(std::make_pair(sq(cloud_->points[i], q), (int)i));
    std::stable_sort(all.begin(), all.end());
    const int n = std::min<int>(k, (int)all.size());
    idx.resize(n); d2.resize(n);
    for (int i = 0; i < n; ++i) { idx[i] = all[i].second; d2[i] = all[i].first; }
    return n;
  }
  int radiusSearch(const PointT& q, double radius, std::vector<int>& idx, std::vector<float>& d2, unsigned max_nn = 0) const {
    std::vector<std::pair<float, int> > all;
    const float r2 = (float)(radius * radius);
    for (size_t i = 0; i < cloud_->points.size(); ++i) {
      float d = sq(cloud_->points[i], q);
      if (d < r2) all.push_back(std::make_pair(d, (int)i));
    }
    std::stable_sort(all.begin(), all.end());
    if (max_nn && all.size() > max_nn) all.resize(max_nn);
    idx.resize(all.size()); d2.resize(all.size());
    for (size_t i = 0; i < all.size(); ++i) { idx[i] = all[i].second; d2[i] = all[i].first; }
    return (int)all.size();
  }

 private:
  typename PointCloud<PointT>::ConstPtr cloud_;
};
}  // namespace pcl
