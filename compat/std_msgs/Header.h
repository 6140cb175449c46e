// Stand-in for ROS std_msgs/Header.h.
#pragma once
#include <string>
#include <ros/ros.h>
namespace std_msgs {
struct Header { unsigned seq = 0; ros::Time stamp; std::string frame_id; };
struct ColorRGBA { float r = 0, g = 0, b = 0, a = 0; };
struct Bool { bool data = false; };
}  // namespace std_msgs
