// Stand-in for pcl/features/normal_3d.h: included by the reference, nothing of it is used on the path.
#pragma once
#include <pcl/point_types.h>
