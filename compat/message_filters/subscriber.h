// Stand-in for ROS message_filters/subscriber.h (unused by the path).
#pragma once
