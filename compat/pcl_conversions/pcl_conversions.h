// Stand-in for pcl_conversions: clouds are only converted for RViz topics, nothing displays them here.
#pragma once
#include <pcl/point_types.h>
#include <sensor_msgs/PointCloud2.h>
namespace pcl {
template <class PointT> inline void toROSMsg(const PointCloud<PointT>&, sensor_msgs::PointCloud2&) {}
}  // namespace pcl
