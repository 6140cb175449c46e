// Stand-in for pcl/kdtree/kdtree.h.
#pragma once
#include "../search/kdtree.h"
