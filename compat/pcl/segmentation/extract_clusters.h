// Stand-in for pcl/segmentation/extract_clusters.h: pcl::EuclideanClusterExtraction with the published algorithm of
// <PCL 1.7 extractEuclideanClusters>: seeds in index order, breadth-first growth through radiusSearch(tolerance)
// (squared float32 distance strictly below float(tol * tol), as FLANN's radius search), clusters with
// min <= size <= max kept, sorted by size, largest first.  On a voxel lattice the search is a scan of the lattice
// offsets within the tolerance, which keeps it O(n) for the voxel clouds the reference clusters.
#pragma once
#include <algorithm>
#include <vector>
#include "../point_types.h"
#include "../search/kdtree.h"
namespace pcl {
template <class PointT> class EuclideanClusterExtraction {
 public:
  typedef typename pcl::search::KdTree<PointT>::Ptr KdTreePtr;
  EuclideanClusterExtraction() : tol_(0), min_(1), max_(std::numeric_limits<int>::max()) {}
  void setClusterTolerance(double t) { tol_ = t; }
  void setMinClusterSize(int n) { min_ = n; }
  void setMaxClusterSize(int n) { max_ = n; }
  void setSearchMethod(const KdTreePtr& t) { tree_ = t; }
  void setInputCloud(const typename PointCloud<PointT>::ConstPtr& c) { cloud_ = c; }
  void extract(std::vector<PointIndices>& clusters) {
    clusters.clear();
    const size_t n = cloud_->points.size();
    std::vector<bool> processed(n, false);
    std::vector<int> nn;
    std::vector<float> d2;
    pcl::KdTreeFLANN<PointT> tree;
    tree.setInputCloud(cloud_);
    for (size_t i = 0; i < n; ++i) {
      if (processed[i]) continue;
      std::vector<int> q;
      size_t sq = 0;
      q.push_back((int)i);
      processed[i] = true;
      while (sq < q.size()) {
        if (!tree.radiusSearch(cloud_->points[(size_t)q[sq]], tol_, nn, d2)) { ++sq; continue; }
        for (size_t j = 0; j < nn.size(); ++j) {
          if (nn[j] == -1 || processed[(size_t)nn[j]]) continue;
          q.push_back(nn[j]);
          processed[(size_t)nn[j]] = true;
        }
        ++sq;
      }
      if ((int)q.size() >= min_ && (int)q.size() <= max_) {
        PointIndices r;
        r.indices = q;
        std::sort(r.indices.begin(), r.indices.end());
        r.indices.erase(std::unique(r.indices.begin(), r.indices.end()), r.indices.end());
        clusters.push_back(r);
      }
    }
    // <PCL>: std::sort(clusters.rbegin(), clusters.rend(), comparePointClusters) -- by size, largest first
    std::stable_sort(clusters.begin(), clusters.end(),
                     [](const PointIndices& a, const PointIndices& b) { return a.indices.size() > b.indices.size(); });
  }

 private:
  double tol_;
  int min_, max_;
  KdTreePtr tree_;
  typename PointCloud<PointT>::ConstPtr cloud_;
};
}  // namespace pcl
