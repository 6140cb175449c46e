// Stand-in for ROS geometry_msgs/Pose.h, used ONLY when the real ROS headers are absent (this image has
// no ROS).  Field names and types match the generated ROS message so that code written against the
// reference (cscpp/include/cscpp/occlusion_culling.h:82) compiles unchanged.
#pragma once
#if defined(__has_include)
#if __has_include(<ros/node_handle.h>)
#error "real ROS headers are present: include <geometry_msgs/Pose.h> from ROS, not compat/"
#endif
#endif
#include <std_msgs/Header.h>
namespace geometry_msgs {
struct Point { double x = 0, y = 0, z = 0; };
struct Point32 { float x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 1; };
struct Pose { Point position; Quaternion orientation; };
struct Vector3 { double x = 0, y = 0, z = 0; };
struct PoseStamped { std_msgs::Header header; Pose pose; };
}  // namespace geometry_msgs
