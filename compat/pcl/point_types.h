// Stand-in for the two PCL types the hot path's class surface uses (pcl::PointXYZ, pcl::PointCloud<T>),
// used ONLY when PCL is absent.  Layout of PointXYZ matches PCL's (16-byte aligned x y z + padding).
#pragma once
#include <memory>
#include <vector>
namespace pcl {
struct alignas(16) PointXYZ {
  float x = 0, y = 0, z = 0, pad = 1.0f;
  PointXYZ() {}
  PointXYZ(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
template <class PointT>
struct PointCloud {
  typedef std::shared_ptr<PointCloud<PointT> > Ptr;
  typedef std::shared_ptr<const PointCloud<PointT> > ConstPtr;
  std::vector<PointT> points;
  unsigned width = 0, height = 1;
  bool is_dense = true;
  float sensor_origin_[4] = {0, 0, 0, 0};       // Eigen::Vector4f in PCL
  float sensor_orientation_[4] = {0, 0, 0, 1};  // Eigen::Quaternionf (x y z w) in PCL
  size_t size() const { return points.size(); }
  void push_back(const PointT& p) { points.push_back(p); }
  PointCloud& operator+=(const PointCloud& o) {  // pcl::PointCloud::operator+= : concatenation
    points.insert(points.end(), o.points.begin(), o.points.end());
    width = (unsigned)points.size();
    return *this;
  }
};
}  // namespace pcl
