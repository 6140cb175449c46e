// Stand-in for ROS geometry_msgs/Quaternion.h (see Pose.h).
#pragma once
#include "Pose.h"
