// Stand-in for ROS nav_msgs/Odometry.h.
#pragma once
#include <std_msgs/Header.h>
namespace nav_msgs { struct Odometry { std_msgs::Header header; }; }
