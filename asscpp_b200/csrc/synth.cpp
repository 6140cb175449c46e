// synth.cpp -- host-side input generators and pose helpers of the C ABI.
//
//  * asscpp_synth_structure_v1 / asscpp_synth_views_v1: the synthetic structures and candidate
//    viewpoints of BASELINE.json configs 2/3/5.  Deterministic PCG32 streams, so the
//    GPU path and the CPU oracle receive bit-identical inputs.
//  * asscpp_robot2sensor: sspp/src/sensors.cpp:47-90 (robot waypoint -> sensor viewpoint), the
//    batcher's input transform, in double precision like the reference.
#include <cmath>
#include <cstdint>
#include <vector>

#include "../../include/asscpp_gpu.h"

namespace {

struct Pcg32 {
  uint64_t state, inc;
  Pcg32(uint64_t seed, uint64_t seq) {
    state = 0;
    inc = (seq << 1) | 1u;
    next();
    state += seed;
    next();
  }
  uint32_t next() {
    uint64_t old = state;
    state = old * 6364136223846793005ull + inc;
    uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
    uint32_t rot = (uint32_t)(old >> 59u);
    return (xs >> rot) | (xs << ((32 - rot) & 31));
  }
  double uni() { return (double)next() * (1.0 / 4294967296.0); }  // [0, 1)
  double range(double a, double b) { return a + (b - a) * uni(); }
};

struct Shape {
  int kind;       // 0 box, 1 vertical cylinder
  double p[6];    // box: x0 y0 z0 ex ey ez; cylinder: cx cy z0 r h -
  double area;
  double face[6]; // box: cumulative face areas; cylinder: lateral, top, bottom
};

const double kDomX = 40.0, kDomY = 40.0, kDomZ = 10.0;

std::vector<Shape> make_shapes(uint64_t seed) {
  Pcg32 r(seed, 1);
  std::vector<Shape> s;
  for (int i = 0; i < 64; ++i) {
    Shape b;
    b.kind = 0;
    double ex = r.range(1, 8), ey = r.range(1, 8), ez = r.range(1, 8);
    double x0 = r.range(0, kDomX - ex), y0 = r.range(0, kDomY - ey);
    double zfree = r.range(0, kDomZ - ez);
    double z0 = (i & 1) ? zfree : 0.0;  // every other box stands on the ground
    b.p[0] = x0; b.p[1] = y0; b.p[2] = z0; b.p[3] = ex; b.p[4] = ey; b.p[5] = ez;
    double fa[6] = {ey * ez, ey * ez, ex * ez, ex * ez, ex * ey, ex * ey};  // -x +x -y +y -z +z
    double acc = 0;
    for (int k = 0; k < 6; ++k) { acc += fa[k]; b.face[k] = acc; }
    b.area = acc;
    s.push_back(b);
  }
  for (int i = 0; i < 16; ++i) {
    Shape c;
    c.kind = 1;
    double rad = r.range(0.5, 2), h = r.range(2, 9);
    double cx = r.range(rad, kDomX - rad), cy = r.range(rad, kDomY - rad);
    c.p[0] = cx; c.p[1] = cy; c.p[2] = 0.0; c.p[3] = rad; c.p[4] = h; c.p[5] = 0;
    double lat = 2 * M_PI * rad * h, cap = M_PI * rad * rad;
    c.face[0] = lat; c.face[1] = lat + cap; c.face[2] = lat + 2 * cap;
    c.face[3] = c.face[4] = c.face[5] = c.face[2];
    c.area = c.face[2];
    s.push_back(c);
  }
  return s;
}

// uniform-by-area surface sample: position + outward normal
void sample_surface(const std::vector<Shape>& shapes, const std::vector<double>& cum, Pcg32& r, double pos[3],
                    double nrm[3]) {
  double t = r.uni() * cum.back();
  size_t lo = 0, hi = cum.size() - 1;
  while (lo < hi) {
    size_t mid = (lo + hi) / 2;
    if (cum[mid] <= t) lo = mid + 1; else hi = mid;
  }
  const Shape& s = shapes[lo];
  double f = r.uni() * s.area, a = r.uni(), b = r.uni();
  nrm[0] = nrm[1] = nrm[2] = 0;
  if (s.kind == 0) {
    int face = 0;
    while (face < 5 && s.face[face] <= f) ++face;
    double x0 = s.p[0], y0 = s.p[1], z0 = s.p[2], ex = s.p[3], ey = s.p[4], ez = s.p[5];
    switch (face) {
      case 0: pos[0] = x0;      pos[1] = y0 + a * ey; pos[2] = z0 + b * ez; nrm[0] = -1; break;
      case 1: pos[0] = x0 + ex; pos[1] = y0 + a * ey; pos[2] = z0 + b * ez; nrm[0] = 1; break;
      case 2: pos[0] = x0 + a * ex; pos[1] = y0;      pos[2] = z0 + b * ez; nrm[1] = -1; break;
      case 3: pos[0] = x0 + a * ex; pos[1] = y0 + ey; pos[2] = z0 + b * ez; nrm[1] = 1; break;
      case 4: pos[0] = x0 + a * ex; pos[1] = y0 + b * ey; pos[2] = z0;      nrm[2] = -1; break;
      default: pos[0] = x0 + a * ex; pos[1] = y0 + b * ey; pos[2] = z0 + ez; nrm[2] = 1; break;
    }
  } else {
    double cx = s.p[0], cy = s.p[1], z0 = s.p[2], rad = s.p[3], h = s.p[4];
    if (f < s.face[0]) {
      double th = 2 * M_PI * a;
      nrm[0] = std::cos(th); nrm[1] = std::sin(th);
      pos[0] = cx + rad * nrm[0]; pos[1] = cy + rad * nrm[1]; pos[2] = z0 + b * h;
    } else {
      double th = 2 * M_PI * a, rr = rad * std::sqrt(b);
      pos[0] = cx + rr * std::cos(th); pos[1] = cy + rr * std::sin(th);
      if (f < s.face[1]) { pos[2] = z0 + h; nrm[2] = 1; } else { pos[2] = z0; nrm[2] = -1; }
    }
  }
}

std::vector<double> cumulative(const std::vector<Shape>& shapes) {
  std::vector<double> cum;
  double acc = 0;
  for (const Shape& s : shapes) { acc += s.area; cum.push_back(acc); }
  return cum;
}

void quat_to_mat(const double q[4], double R[3][3]) {  // <tf Matrix3x3::setRotation>
  double x = q[0], y = q[1], z = q[2], w = q[3];
  double d = x * x + y * y + z * z + w * w, s = 2.0 / d;
  double xs = x * s, ys = y * s, zs = z * s, wx = w * xs, wy = w * ys, wz = w * zs;
  double xx = x * xs, xy = x * ys, xz = x * zs, yy = y * ys, yz = y * zs, zz = z * zs;
  R[0][0] = 1.0 - (yy + zz); R[0][1] = xy - wz; R[0][2] = xz + wy;
  R[1][0] = xy + wz; R[1][1] = 1.0 - (xx + zz); R[1][2] = yz - wx;
  R[2][0] = xz - wy; R[2][1] = yz + wx; R[2][2] = 1.0 - (xx + yy);
}

void mat_to_quat(const double m[3][3], double q[4]) {  // <tf Matrix3x3::getRotation>
  double trace = m[0][0] + m[1][1] + m[2][2];
  double t[4];
  if (trace > 0.0) {
    double s = std::sqrt(trace + 1.0);
    t[3] = s * 0.5;
    s = 0.5 / s;
    t[0] = (m[2][1] - m[1][2]) * s;
    t[1] = (m[0][2] - m[2][0]) * s;
    t[2] = (m[1][0] - m[0][1]) * s;
  } else {
    int i = m[0][0] < m[1][1] ? (m[1][1] < m[2][2] ? 2 : 1) : (m[0][0] < m[2][2] ? 2 : 0);
    int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0);
    t[i] = s * 0.5;
    s = 0.5 / s;
    t[3] = (m[k][j] - m[j][k]) * s;
    t[j] = (m[j][i] + m[i][j]) * s;
    t[k] = (m[k][i] + m[i][k]) * s;
  }
  for (int c = 0; c < 4; ++c) q[c] = t[c];
}

void rpy_to_quat(const double rpy[3], double q[4]) {  // <tf createQuaternionFromRPY>
  double hy = rpy[2] * 0.5, hp = rpy[1] * 0.5, hr = rpy[0] * 0.5;
  double cy = std::cos(hy), sy = std::sin(hy), cp = std::cos(hp), sp = std::sin(hp), cr = std::cos(hr), sr = std::sin(hr);
  q[0] = sr * cp * cy - cr * sp * sy;
  q[1] = cr * sp * cy + sr * cp * sy;
  q[2] = cr * cp * sy - sr * sp * cy;
  q[3] = cr * cp * cy + sr * sp * sy;
}

}  // namespace

extern "C" {

int asscpp_robot2sensor(const double robot_pose[7], const double offset[3], const double rpy[3],
                        double sensor_pose[7]) {
  if (!robot_pose || !offset || !rpy || !sensor_pose) return ASSCPP_EINVAL;
  double R1[3][3], R2[3][3], q2[4];
  quat_to_mat(robot_pose + 3, R1);             // sensors.cpp:54-60
  rpy_to_quat(rpy, q2);                        // :63
  quat_to_mat(q2, R2);                         // :64-69
  double M[3][3], T[3];
  for (int i = 0; i < 3; ++i) {                // :72 robotPoseMat * robot2sensorMat
    for (int j = 0; j < 3; ++j) M[i][j] = R1[i][0] * R2[0][j] + R1[i][1] * R2[1][j] + R1[i][2] * R2[2][j];
    T[i] = R1[i][0] * offset[0] + R1[i][1] * offset[1] + R1[i][2] * offset[2] + robot_pose[i];
  }
  const double F[3][3] = {{1, 0, 0}, {0, 0, -1}, {0, 1, 0}};  // :76-79 (+90 deg about the sensor X axis)
  double N[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) N[i][j] = M[i][0] * F[0][j] + M[i][1] * F[1][j] + M[i][2] * F[2][j];
  mat_to_quat(N, sensor_pose + 3);             // :87
  sensor_pose[0] = T[0]; sensor_pose[1] = T[1]; sensor_pose[2] = T[2];
  return ASSCPP_OK;
}

int asscpp_robot2sensor_batch(const double* robot_poses, int64_t n, const double offset[3], const double rpy[3],
                              double* sensor_poses) {
  if (n < 0 || ((!robot_poses || !sensor_poses) && n > 0) || !offset || !rpy) return ASSCPP_EINVAL;
  // same operation order per pose as the one-pose entry point (the products are not re-associated), so the
  // two agree bit for bit
  for (int64_t k = 0; k < n; ++k) {
    int rc = asscpp_robot2sensor(robot_poses + 7 * k, offset, rpy, sensor_poses + 7 * k);
    if (rc) return rc;
  }
  return ASSCPP_OK;
}

int asscpp_synth_structure_v1(uint64_t structure_seed, uint64_t point_seed, int64_t n, float* xyz_out) {
  if (n < 0 || (!xyz_out && n > 0)) return ASSCPP_EINVAL;
  std::vector<Shape> shapes = make_shapes(structure_seed);
  std::vector<double> cum = cumulative(shapes);
  Pcg32 r(point_seed, 2);
  for (int64_t i = 0; i < n; ++i) {
    double pos[3], nrm[3];
    sample_surface(shapes, cum, r, pos, nrm);
    xyz_out[3 * i] = (float)pos[0];
    xyz_out[3 * i + 1] = (float)pos[1];
    xyz_out[3 * i + 2] = (float)pos[2];
  }
  return ASSCPP_OK;
}

int asscpp_synth_views_v1(uint64_t structure_seed, uint64_t view_seed, int64_t nviews, double* poses_out) {
  if (nviews < 0 || (!poses_out && nviews > 0)) return ASSCPP_EINVAL;
  std::vector<Shape> shapes = make_shapes(structure_seed);
  std::vector<double> cum = cumulative(shapes);
  Pcg32 r(view_seed, 3);
  int64_t made = 0;
  while (made < nviews) {
    double pos[3], nrm[3];
    sample_surface(shapes, cum, r, pos, nrm);
    double dist = r.range(1.0, 4.0);
    uint32_t yaw_pick = r.next() & 7u;
    bool pitch_up = (r.next() & 1u) != 0;
    double eye[3] = {pos[0] + dist * nrm[0], pos[1] + dist * nrm[1], pos[2] + dist * nrm[2]};
    if (eye[2] < 0.3) continue;
    int k;
    if (std::fabs(nrm[0]) + std::fabs(nrm[1]) > 1e-9) {
      // nearest multiple of 45 degrees to the direction facing the surface point
      double yaw = std::atan2(-nrm[1], -nrm[0]);
      k = (int)std::floor(yaw / (M_PI / 4) + 0.5);
      k = ((k % 8) + 8) % 8;
    } else {
      k = (int)yaw_pick;
    }
    // robot yaw about Z, sensor pitch about Y, then the sensor fix of sensors.cpp:76-79
    double robot[7] = {eye[0], eye[1], eye[2], 0, 0, std::sin(k * (M_PI / 4) * 0.5), std::cos(k * (M_PI / 4) * 0.5)};
    double offset[3] = {0, 0, 0};
    double rpy[3] = {0, pitch_up ? -0.349 : 0.349, 0};
    asscpp_robot2sensor(robot, offset, rpy, poses_out + 7 * made);
    ++made;
  }
  return ASSCPP_OK;
}

}  // extern "C"
