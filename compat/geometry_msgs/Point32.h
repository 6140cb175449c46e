// Stand-in for ROS geometry_msgs/Point32.h (see Pose.h).
#pragma once
#include "Pose.h"
