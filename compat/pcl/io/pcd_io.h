// Stand-in for pcl::io::loadPCDFile<pcl::PointXYZ> (occlusion_culling.cpp:39): PCD v0.7, FIELDS x y z,
// float32, DATA ascii | binary.  Sets sensor_origin_ from VIEWPOINT like PCL does.
#pragma once
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include "../point_types.h"
namespace pcl {
namespace io {
template <class PointT> inline int loadPCDFile(const std::string& path, PointCloud<PointT>& cloud) {
  std::ifstream f(path.c_str(), std::ios::binary);
  if (!f) return -1;
  std::string line, data;
  size_t npts = 0;
  while (std::getline(f, line)) {
    if (!line.empty() && line[line.size() - 1] == '\r') line.erase(line.size() - 1);
    std::istringstream ss(line);
    std::string key;
    ss >> key;
    if (key == "FIELDS") {
      std::string a, b, c;
      ss >> a >> b >> c;
      if (a != "x" || b != "y" || c != "z") return -2;
    } else if (key == "POINTS") {
      ss >> npts;
    } else if (key == "VIEWPOINT") {
      ss >> cloud.sensor_origin_[0] >> cloud.sensor_origin_[1] >> cloud.sensor_origin_[2];
      cloud.sensor_origin_[3] = 0;
    } else if (key == "DATA") {
      ss >> data;
      break;
    }
  }
  cloud.points.resize(npts);
  if (data == "binary") {
    std::vector<float> raw(3 * npts);
    f.read(reinterpret_cast<char*>(raw.data()), (std::streamsize)(raw.size() * sizeof(float)));
    if ((size_t)f.gcount() != raw.size() * sizeof(float)) return -3;
    for (size_t i = 0; i < npts; ++i) { PointT q; q.x = raw[3 * i]; q.y = raw[3 * i + 1]; q.z = raw[3 * i + 2]; cloud.points[i] = q; }
  } else if (data == "ascii") {
    for (size_t i = 0; i < npts; ++i) {
      float x, y, z;
      if (!(f >> x >> y >> z)) return -3;
      { PointT q; q.x = x; q.y = y; q.z = z; cloud.points[i] = q; }
    }
  } else {
    return -4;
  }
  cloud.width = (unsigned)npts;
  return 0;
}
}  // namespace io
}  // namespace pcl
