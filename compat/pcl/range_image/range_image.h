// Stand-in for pcl/range_image/range_image.h: included by the reference, nothing of it is used on the path.
#pragma once
#include <pcl/point_types.h>
