// Stand-in for ROS geometry_msgs/Point.h (see Pose.h).
#pragma once
#include "Pose.h"
