// Stand-in for pcl/search/kdtree.h: the search object handed to EuclideanClusterExtraction.
#pragma once
#include "../kdtree/kdtree_flann.h"
namespace pcl {
namespace search {
template <class PointT> class KdTree : public pcl::KdTreeFLANN<PointT> {
 public:
  typedef std::shared_ptr<KdTree<PointT> > Ptr;
};
}  // namespace search
}  // namespace pcl
