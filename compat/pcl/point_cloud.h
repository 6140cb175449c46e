// Stand-in for pcl/point_cloud.h.
#pragma once
#include "point_types.h"
