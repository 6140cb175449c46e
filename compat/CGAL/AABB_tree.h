// Stand-in for CGAL::AABB_tree over triangles: see Simple_cartesian.h in this directory.
#pragma once
#include "Simple_cartesian.h"
namespace CGAL {
template <class K, class Iterator> struct AABB_triangle_primitive { typedef K Kernel; typedef Iterator Id; };
template <class K, class Primitive> struct AABB_traits { typedef K Kernel; };
template <class Traits> class AABB_tree {
 public:
  typedef typename Traits::Kernel K;
  typedef typename K::FT FT;
  template <class It> AABB_tree(It first, It last) {
    for (It it = first; it != last; ++it) {
      Tri t;
      for (int c = 0; c < 3; ++c)
        for (int a = 0; a < 3; ++a) t.p[3 * c + a] = (double)(*it).vertex(c)[a];
      for (int a = 0; a < 3; ++a) {
        t.lo[a] = std::min(t.p[a], std::min(t.p[3 + a], t.p[6 + a]));
        t.hi[a] = std::max(t.p[a], std::max(t.p[3 + a], t.p[6 + a]));
      }
      tris_.push_back(t);
    }
  }
  size_t size() const { return tris_.size(); }
  size_t number_of_intersected_primitives(const typename K::Segment_3& s) const {
    const double o[3] = {s.a[0], s.a[1], s.a[2]};
    const double d[3] = {s.b[0] - s.a[0], s.b[1] - s.a[1], s.b[2] - s.a[2]};
    double lo[3], hi[3];
    for (int a = 0; a < 3; ++a) { lo[a] = std::min(s.a[a], s.b[a]); hi[a] = std::max(s.a[a], s.b[a]); }
    size_t n = 0;
    for (size_t k = 0; k < tris_.size(); ++k) {
      const Tri& t = tris_[k];
      if (t.hi[0] < lo[0] || t.lo[0] > hi[0] || t.hi[1] < lo[1] || t.lo[1] > hi[1] || t.hi[2] < lo[2] || t.lo[2] > hi[2]) continue;
      if (hit(o, d, t.p, true)) ++n;
    }
    return n;
  }
  size_t number_of_intersected_primitives(const typename K::Ray_3& r) const {
    const double o[3] = {r.a[0], r.a[1], r.a[2]};
    const double d[3] = {r.b[0] - r.a[0], r.b[1] - r.a[1], r.b[2] - r.a[2]};
    size_t n = 0;
    for (size_t k = 0; k < tris_.size(); ++k)
      if (hit(o, d, tris_[k].p, false)) ++n;
    return n;
  }
  FT squared_distance(const typename K::Point_3& q) const {
    const double p[3] = {q[0], q[1], q[2]};
    double best = std::numeric_limits<double>::infinity();
    for (size_t k = 0; k < tris_.size(); ++k) best = std::min(best, sqdist(p, tris_[k].p));
    return (FT)best;
  }

 private:
  struct Tri { double p[9], lo[3], hi[3]; };
  std::vector<Tri> tris_;
  static double dot(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
  static bool hit(const double* o, const double* d, const double* t, bool segment) {
    const double e1[3] = {t[3] - t[0], t[4] - t[1], t[5] - t[2]}, e2[3] = {t[6] - t[0], t[7] - t[1], t[8] - t[2]};
    const double h[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
    double det = dot(e1, h);
    if (det == 0.0) return false;
    const double sv[3] = {o[0] - t[0], o[1] - t[1], o[2] - t[2]};
    double u = dot(sv, h);
    const double q[3] = {sv[1] * e1[2] - sv[2] * e1[1], sv[2] * e1[0] - sv[0] * e1[2], sv[0] * e1[1] - sv[1] * e1[0]};
    double v = dot(d, q), w = dot(e2, q);
    if (det < 0.0) { det = -det; u = -u; v = -v; w = -w; }
    if (u < 0.0 || v < 0.0 || u + v > det || w < 0.0) return false;
    return !segment || w <= det;
  }
  static double sqdist(const double* p, const double* t) {
    const double *a = t, *b = t + 3, *c = t + 6;
    const double ab[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, ac[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
    const double ap[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
    double cl[3];
    const double d1 = dot(ab, ap), d2 = dot(ac, ap);
#define ASSCPP_CGAL_FIN { const double e[3] = {p[0] - cl[0], p[1] - cl[1], p[2] - cl[2]}; return dot(e, e); }
    if (d1 <= 0.0 && d2 <= 0.0) { cl[0] = a[0]; cl[1] = a[1]; cl[2] = a[2]; ASSCPP_CGAL_FIN }
    const double bp[3] = {p[0] - b[0], p[1] - b[1], p[2] - b[2]};
    const double d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.0 && d4 <= d3) { cl[0] = b[0]; cl[1] = b[1]; cl[2] = b[2]; ASSCPP_CGAL_FIN }
    const double vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
      const double v = d1 / (d1 - d3);
      for (int k = 0; k < 3; ++k) cl[k] = a[k] + v * ab[k];
      ASSCPP_CGAL_FIN
    }
    const double cp[3] = {p[0] - c[0], p[1] - c[1], p[2] - c[2]};
    const double d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.0 && d5 <= d6) { cl[0] = c[0]; cl[1] = c[1]; cl[2] = c[2]; ASSCPP_CGAL_FIN }
    const double vb = d5 * d2 - d1 * d6;
    if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
      const double w = d2 / (d2 - d6);
      for (int k = 0; k < 3; ++k) cl[k] = a[k] + w * ac[k];
      ASSCPP_CGAL_FIN
    }
    const double va = d3 * d6 - d5 * d4;
    if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
      const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
      for (int k = 0; k < 3; ++k) cl[k] = b[k] + w * (c[k] - b[k]);
      ASSCPP_CGAL_FIN
    }
    const double denom = 1.0 / (va + vb + vc);
    const double v = vb * denom, w = vc * denom;
    for (int k = 0; k < 3; ++k) cl[k] = a[k] + ab[k] * v + ac[k] * w;
    ASSCPP_CGAL_FIN
#undef ASSCPP_CGAL_FIN
  }
};
}  // namespace CGAL
