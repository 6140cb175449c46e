// Stand-in for pcl/surface/concave_hull.h.  findClusterOuterPoints (coverage_path_planning_heuristic.cpp:192-216) is
// never called by the planners (its call sites are commented out, sspp/src/pathplanner.cpp:746,817): the hull of the
// stand-in is the input itself.
#pragma once
#include "../point_types.h"
namespace pcl {
template <class PointT> class ConcaveHull {
 public:
  void setInputCloud(const typename PointCloud<PointT>::ConstPtr& c) { cloud_ = c; }
  void setAlpha(double) {}
  void reconstruct(PointCloud<PointT>& hull) { if (cloud_) hull = *cloud_; }

 private:
  typename PointCloud<PointT>::ConstPtr cloud_;
};
template <class PointT> class ConvexHull : public ConcaveHull<PointT> {};
}  // namespace pcl
