// Stand-in for the part of sspp/include/sspp/utils.h the plugin interface needs (Pose :392-413, Tree :493-497,
// Dist :351-359), used ONLY when the sspp package headers are not on the include path.  With the reference's
// headers in front (-I<reference>/sspp/include) these files are never reached.
#ifndef UTILS_H
#define UTILS_H
#include <cmath>
#include <vector>
#include <geometry_msgs/Pose.h>
#ifndef DTOR
#define DTOR(d) ((d) * M_PI / 180)
#endif
inline double Dist(geometry_msgs::Point a, geometry_msgs::Point b) {
  return sqrt((a.x - b.x) * (a.x - b.x) + (a.y - b.y) * (a.y - b.y) + (a.z - b.z) * (a.z - b.z));
}
inline double Dist(geometry_msgs::Pose a, geometry_msgs::Pose b) { return Dist(a.position, b.position); }
class Pose {
 public:
  geometry_msgs::Pose p;
  double phi;
  Pose() {}
  Pose(double x, double y, double z, double theta) { p.position.x = x; p.position.y = y; p.position.z = z; phi = theta; }
};
typedef struct _tree {
  geometry_msgs::Pose location;
  std::vector<geometry_msgs::Pose> children;
} Tree;
#endif
