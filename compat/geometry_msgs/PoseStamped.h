// Stand-in for ROS geometry_msgs/PoseStamped.h (see Pose.h).
#pragma once
#include "Pose.h"
