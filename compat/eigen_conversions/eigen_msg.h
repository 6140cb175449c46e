// Stand-in for eigen_conversions/eigen_msg.h (nothing of it is used on the path).
#pragma once
#include <Eigen/Core>
