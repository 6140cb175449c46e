// Stand-in for the part of CGAL the reference uses (cscpp/include/cscpp/coverage_path_planning_heuristic.h:63-74,
// cscpp/src/coverage_path_planning_heuristic.cpp:285-335,1355-1451): Simple_cartesian<double> points, segments, rays,
// triangles and an AABB tree answering number_of_intersected_primitives(segment | ray) and squared_distance(point).
// Used ONLY when CGAL is absent (this image has none).  The tree is replaced by a pass over all triangles with a
// bounding-box reject; the predicates are the textbook double-precision ones (division-free Moeller-Trumbore with
// inclusive borders; closest point by Voronoi region), the same formulas as the repo's oracle (orc_mesh_*) and
// mesh.cu.  CGAL's own filtered predicates can differ from these only for queries that graze an edge or a vertex.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <limits>
#include <list>
#include <vector>
namespace CGAL {
template <class FT_> struct Simple_cartesian {
  typedef FT_ FT;
  struct Point_3 {
    FT c[3];
    Point_3() { c[0] = c[1] = c[2] = 0; }
    Point_3(FT x, FT y, FT z) { c[0] = x; c[1] = y; c[2] = z; }
    FT x() const { return c[0]; } FT y() const { return c[1]; } FT z() const { return c[2]; }
    FT operator[](int i) const { return c[i]; }
  };
  struct Segment_3 {
    Point_3 a, b;
    Segment_3() {}
    Segment_3(const Point_3& p, const Point_3& q) : a(p), b(q) {}
    const Point_3& source() const { return a; }
    const Point_3& target() const { return b; }
  };
  struct Ray_3 {  // from a through b
    Point_3 a, b;
    Ray_3() {}
    Ray_3(const Point_3& p, const Point_3& q) : a(p), b(q) {}
  };
  struct Line_3 { Point_3 a, b; Line_3() {} Line_3(const Point_3& p, const Point_3& q) : a(p), b(q) {} };
  struct Triangle_3 {
    Point_3 v[3];
    Triangle_3() {}
    Triangle_3(const Point_3& p, const Point_3& q, const Point_3& r) { v[0] = p; v[1] = q; v[2] = r; }
    const Point_3& vertex(int i) const { return v[i]; }
    const Point_3& operator[](int i) const { return v[i]; }
  };
};
template <class P> inline bool collinear(const P& p, const P& q, const P& r) {
  const double u[3] = {q[0] - p[0], q[1] - p[1], q[2] - p[2]}, v[3] = {r[0] - p[0], r[1] - p[1], r[2] - p[2]};
  return u[1] * v[2] - u[2] * v[1] == 0.0 && u[2] * v[0] - u[0] * v[2] == 0.0 && u[0] * v[1] - u[1] * v[0] == 0.0;
}
template <class K1, class K2> struct Cartesian_converter {};
}  // namespace CGAL
