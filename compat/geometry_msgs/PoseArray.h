// Stand-in for ROS geometry_msgs/PoseArray.h (see Pose.h).
#pragma once
#include <vector>
#include "Pose.h"
namespace geometry_msgs {
struct PoseArray { std_msgs::Header header; std::vector<Pose> poses; };
}  // namespace geometry_msgs
