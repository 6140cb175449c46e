// Stand-in for pcl/sample_consensus/method_types.h: included by the reference, nothing of it is used on the path.
#pragma once
#include <pcl/point_types.h>
