// Stand-in for ros/package.h: ros::package::getPath("cscpp") = $ASSCPP_PACKAGE_PATH/cscpp (default: ".").
#pragma once
#include <cstdlib>
#include <string>
namespace ros {
namespace package {
inline std::string getPath(const std::string& name) {
  const char* base = std::getenv("ASSCPP_PACKAGE_PATH");
  return std::string(base ? base : ".") + "/" + name;
}
}  // namespace package
}  // namespace ros
