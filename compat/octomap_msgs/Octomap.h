// Stand-in for ROS octomap_msgs/Octomap.h (only published for RViz by the reference).
#pragma once
#include <string>
#include <std_msgs/Header.h>
namespace octomap_msgs {
struct Octomap { std_msgs::Header header; bool binary = false; std::string id; double resolution = 0; };
}  // namespace octomap_msgs
