// Stand-in for pcl/common/transforms.h (nothing of it is used on the path).
#pragma once
#include <Eigen/Core>
