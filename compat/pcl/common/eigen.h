// Stand-in for pcl/common/eigen.h.
#pragma once
#include <Eigen/Core>
