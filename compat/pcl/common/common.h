// Stand-in for pcl/common/common.h: getMinMax3D over an index list (<PCL 1.7 common/impl/common.hpp>).
#pragma once
#include <cfloat>
#include <vector>
#include "../point_types.h"
namespace pcl {
template <class PointT> inline void getMinMax3D(const PointCloud<PointT>& cloud, const std::vector<int>& indices,
                                                Eigen::Vector4f& min_pt, Eigen::Vector4f& max_pt) {
  Eigen::Vector4f mn = Eigen::Vector4f::Constant(FLT_MAX), mx = Eigen::Vector4f::Constant(-FLT_MAX);
  for (size_t i = 0; i < indices.size(); ++i) {
    const PointT& p = cloud.points[(size_t)indices[i]];
    if (!cloud.is_dense && (!pcl_isfinite(p.x) || !pcl_isfinite(p.y) || !pcl_isfinite(p.z))) continue;
    const float v[4] = {p.x, p.y, p.z, 1.0f};
    for (int c = 0; c < 4; ++c) { if (v[c] < mn[c]) mn[c] = v[c]; if (v[c] > mx[c]) mx[c] = v[c]; }
  }
  min_pt = mn;
  max_pt = mx;
}
}  // namespace pcl
