// Stand-in for ros/ros.h, used ONLY when ROS is absent (this image has none).  Enough of roscpp for the
// reference's planners and the coverage heuristic to compile unchanged: a node handle that hands out publishers
// which drop their messages (RViz visualisation is out of scope: no graphics stack), wall-clock Time / Duration,
// and the logging macros.
#pragma once
#include <unistd.h>

#include <chrono>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <set>
#include <sstream>
#include <cstdio>
#include <map>
#include <string>
#include <thread>
#include <vector>
namespace ros {
struct Duration {
  double sec_;
  Duration(double s = 0) : sec_(s) {}
  double toSec() const { return sec_; }
  bool sleep() const { return true; }  // visualisation pauses of the reference are skipped
};
struct Time {
  double sec_;
  Time(double s = 0) : sec_(s) {}
  static Time now() {
    return Time(std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count());
  }
  double toSec() const { return sec_; }
  Duration operator-(const Time& o) const { return Duration(sec_ - o.sec_); }
};
struct Rate {
  Rate(double) {}
  bool sleep() { return true; }
};
struct Publisher {
  template <class M> void publish(const M&) const {}
  int getNumSubscribers() const { return 0; }
};
struct Subscriber {};
class NodeHandle {
 public:
  NodeHandle() {}
  explicit NodeHandle(const std::string&) {}
  template <class M> Publisher advertise(const std::string&, int, bool = false) { return Publisher(); }
  template <class T> bool getParam(const std::string&, T&) const { return false; }
  template <class T> bool param(const std::string&, T& v, const T& d) const { v = d; return false; }
  bool hasParam(const std::string&) const { return false; }
};
inline void init(int&, char**, const std::string&) {}
inline bool ok() { return true; }
inline void spinOnce() {}
inline void shutdown() {}
}  // namespace ros
#define ROS_INFO(...) do { std::printf(__VA_ARGS__); std::printf("\n"); } while (0)
#define ROS_WARN(...) do { std::fprintf(stderr, __VA_ARGS__); std::fprintf(stderr, "\n"); } while (0)
#define ROS_ERROR(...) do { std::fprintf(stderr, __VA_ARGS__); std::fprintf(stderr, "\n"); } while (0)
#define ROS_DEBUG(...) do { } while (0)
#define ROS_INFO_STREAM(x) do { } while (0)
#define ROS_WARN_STREAM(x) do { } while (0)
#define ROS_ERROR_STREAM(x) do { } while (0)
