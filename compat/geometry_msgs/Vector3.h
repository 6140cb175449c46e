// Stand-in for ROS geometry_msgs/Vector3.h (see Pose.h).
#pragma once
#include "Pose.h"
