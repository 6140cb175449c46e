// Stand-in for pcl/common/io.h: copyPointCloud.
#pragma once
#include <vector>
#include "../point_types.h"
namespace pcl {
template <class PointT> inline void copyPointCloud(const PointCloud<PointT>& in, PointCloud<PointT>& out) {
  out = in;
}
template <class PointT> inline void copyPointCloud(const PointCloud<PointT>& in, const std::vector<int>& indices,
                                                   PointCloud<PointT>& out) {
  if (indices.size() == in.points.size()) { out = in; return; }  // <PCL>: "do we want to copy everything?"
  out.points.resize(indices.size());
  out.header = in.header;
  out.width = (uint32_t)indices.size();
  out.height = 1;
  out.is_dense = in.is_dense;
  out.sensor_orientation_ = in.sensor_orientation_;
  out.sensor_origin_ = in.sensor_origin_;
  for (size_t i = 0; i < indices.size(); ++i) out.points[i] = in.points[(size_t)indices[i]];
}
}  // namespace pcl
