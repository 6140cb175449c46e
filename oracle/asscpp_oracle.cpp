/*
 * asscpp_oracle.cpp -- CPU restatement of the kucars/asscpp coverage-evaluation hot path.
 * TEST INFRASTRUCTURE ONLY (see asscpp_oracle.h).  PARITY UNPINNED (no reference golden data).
 *
 * Build: g++ -O3 -std=c++17 -ffp-contract=off -fno-fast-math -fPIC -shared (oracle/Makefile).
 * Every float32 expression is written out operation by operation; no FMA contraction, no
 * library dot/normalize.  Citations are to /root/reference/cscpp/... unless prefixed sspp/.
 */
#include "asscpp_oracle.h"

#include <algorithm>
#include <map>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace {

typedef float f32;

struct Ctx {
  int64_t n = 0;
  std::vector<f32> xyz;  // 3*n
  OrcParams p;
  f32 inv_leaf = 0;      // <PCL setLeafSize>: inverse_leaf_size_ = 1 / leaf
  f32 inv_cov = 0;
  std::vector<int32_t> cov_id;  // per point compact coverage-voxel id
  int64_t n_cov = 0;
  int64_t n_occ = 0;
  int32_t min_b[3] = {0, 0, 0}, max_b[3] = {0, 0, 0};  // occlusion-lattice model box
  int32_t cov_min_b[3] = {0, 0, 0};
  std::vector<int32_t> cov_ijk;  // 3*n_cov lattice coordinates of the coverage voxels, compact-id order
};

inline int32_t ifloor(f32 v) { return static_cast<int32_t>(std::floor(v)); }

// <Eigen> Vector4f dot / squaredNorm reduction order (SSE2 predux): (a0+a2)+(a1+a3)
inline f32 dot4(f32 a0, f32 a1, f32 a2, f32 a3) {
  f32 s02 = a0 + a2;
  f32 s13 = a1 + a3;
  return s02 + s13;
}
// <Eigen> 3-vector dot: a0b0 + (a1b1 + a2b2)
inline f32 dot3(const f32 a[3], const f32 b[3]) {
  f32 p0 = a[0] * b[0];
  f32 p1 = a[1] * b[1];
  f32 p2 = a[2] * b[2];
  f32 s12 = p1 + p2;
  return p0 + s12;
}
inline void cross3(const f32 a[3], const f32 b[3], f32 o[3]) {
  f32 t0 = a[1] * b[2], t1 = a[2] * b[1];
  f32 t2 = a[2] * b[0], t3 = a[0] * b[2];
  f32 t4 = a[0] * b[1], t5 = a[1] * b[0];
  o[0] = t0 - t1;
  o[1] = t2 - t3;
  o[2] = t4 - t5;
}

// occlusion_culling.cpp:271-290 -- <tf Matrix3x3::setRotation> in double, narrowed to f32.
void pose_to_camera(const double pose[7], f32 R[3][3], f32 T[3]) {
  double x = pose[3], y = pose[4], z = pose[5], w = pose[6];
  double d = x * x + y * y + z * z + w * w;
  double s = 2.0 / d;
  double xs = x * s, ys = y * s, zs = z * s;
  double wx = w * xs, wy = w * ys, wz = w * zs;
  double xx = x * xs, xy = x * ys, xz = x * zs;
  double yy = y * ys, yz = y * zs, zz = z * zs;
  R[0][0] = (f32)(1.0 - (yy + zz)); R[0][1] = (f32)(xy - wz);         R[0][2] = (f32)(xz + wy);
  R[1][0] = (f32)(xy + wz);         R[1][1] = (f32)(1.0 - (xx + zz)); R[1][2] = (f32)(yz - wx);
  R[2][0] = (f32)(xz - wy);         R[2][1] = (f32)(yz + wx);         R[2][2] = (f32)(1.0 - (xx + yy));
  T[0] = (f32)pose[0]; T[1] = (f32)pose[1]; T[2] = (f32)pose[2];
}

// frustum_culling.cpp:82-141.  planes in order l r t b f n (the order of the test at :157-162).
void frustum(const OrcParams& p, const double pose[7], f32 planes[24], f32 T[3], f32 corners[24]) {
  f32 R[3][3];
  pose_to_camera(pose, R, T);
  f32 view[3] = {R[0][0], R[1][0], R[2][0]};   // :82 first column
  f32 up[3] = {R[0][1], R[1][1], R[2][1]};     // :83
  f32 right[3] = {R[0][2], R[1][2], R[2][2]};  // :84
  f32 vr = (f32)((double)p.vfov_deg * M_PI / 180);  // :88
  f32 hr = (f32)((double)p.hfov_deg * M_PI / 180);  // :89
  f32 vr2 = vr / 2, hr2 = hr / 2;
  f32 np_h = (f32)(2 * std::tan((double)vr2) * (double)p.near_dist);  // :91
  f32 np_w = (f32)(2 * std::tan((double)hr2) * (double)p.near_dist);  // :92
  f32 fp_h = (f32)(2 * std::tan((double)vr2) * (double)p.far_dist);   // :94
  f32 fp_w = (f32)(2 * std::tan((double)hr2) * (double)p.far_dist);   // :95
  f32 fp_c[3], np_c[3], fp_tl[3], fp_tr[3], fp_bl[3], fp_br[3], np_tl[3], np_tr[3], np_bl[3], np_br[3];
  for (int c = 0; c < 3; ++c) {
    f32 vf = view[c] * p.far_dist;
    fp_c[c] = T[c] + vf;  // :97
    f32 uh = (up[c] * fp_h) / 2;
    f32 rw = (right[c] * fp_w) / 2;
    fp_tl[c] = (fp_c[c] + uh) - rw;  // :98
    fp_tr[c] = (fp_c[c] + uh) + rw;  // :99
    fp_bl[c] = (fp_c[c] - uh) - rw;  // :100
    fp_br[c] = (fp_c[c] - uh) + rw;  // :101
    f32 vn = view[c] * p.near_dist;
    np_c[c] = T[c] + vn;  // :103
    f32 uhn = (up[c] * np_h) / 2;
    f32 rwn = (right[c] * np_w) / 2;
    np_tl[c] = (np_c[c] + uhn) - rwn;  // :104
    np_tr[c] = (np_c[c] + uhn) + rwn;  // :105
    np_bl[c] = (np_c[c] - uhn) - rwn;  // :106
    np_br[c] = (np_c[c] - uhn) + rwn;  // :107
  }
  if (corners) {
    const f32* cs[8] = {fp_tl, fp_tr, fp_bl, fp_br, np_tl, np_tr, np_bl, np_br};
    for (int k = 0; k < 8; ++k)
      for (int c = 0; c < 3; ++c) corners[3 * k + c] = cs[k][c];
  }
  f32 *pl_l = planes, *pl_r = planes + 4, *pl_t = planes + 8, *pl_b = planes + 12, *pl_f = planes + 16,
      *pl_n = planes + 20;
  f32 e1[3], e2[3];
  for (int c = 0; c < 3; ++c) { e1[c] = fp_bl[c] - fp_br[c]; e2[c] = fp_tr[c] - fp_br[c]; }
  cross3(e1, e2, pl_f);          // :109
  pl_f[3] = -dot3(fp_c, pl_f);   // :110
  for (int c = 0; c < 3; ++c) { e1[c] = np_tr[c] - np_br[c]; e2[c] = np_bl[c] - np_br[c]; }
  cross3(e1, e2, pl_n);          // :112
  pl_n[3] = -dot3(np_c, pl_n);   // :113
  f32 a[3], b[3], cc[3], d[3];
  for (int c = 0; c < 3; ++c) {
    a[c] = fp_bl[c] - T[c];   // :115
    b[c] = fp_br[c] - T[c];   // :116
    cc[c] = fp_tr[c] - T[c];  // :117
    d[c] = fp_tl[c] - T[c];   // :118
  }
  cross3(b, cc, pl_r);  // :133
  cross3(d, a, pl_l);   // :134
  cross3(cc, d, pl_t);  // :135
  cross3(a, b, pl_b);   // :136
  pl_r[3] = -dot3(T, pl_r);  // :138
  pl_l[3] = -dot3(T, pl_l);
  pl_t[3] = -dot3(T, pl_t);
  pl_b[3] = -dot3(T, pl_b);
}

// frustum_culling.cpp:153-162: pt=(x,y,z,1); inside iff every pt.dot(pl) <= 0
inline f32 plane_dot(const f32* pl, f32 x, f32 y, f32 z) {
  return dot4(x * pl[0], y * pl[1], z * pl[2], 1.0f * pl[3]);
}
inline bool in_frustum(const f32* planes, f32 x, f32 y, f32 z) {
  return plane_dot(planes, x, y, z) <= 0 && plane_dot(planes + 4, x, y, z) <= 0 &&
         plane_dot(planes + 8, x, y, z) <= 0 && plane_dot(planes + 12, x, y, z) <= 0 &&
         plane_dot(planes + 16, x, y, z) <= 0 && plane_dot(planes + 20, x, y, z) <= 0;
}

// Per-view voxel grid: voxel_grid_occlusion_estimation.cpp:43-65 + <PCL VoxelGrid::applyFilter>.
struct ViewGrid {
  f32 leaf, inv, half;
  int32_t min_b[3], max_b[3], div[3], mul[3];
  int64_t cells;
  f32 b_min[3], b_max[3];
  f32 origin[3];           // sensor_origin_ (xyz; w = 0)
  const uint8_t* occ;      // leaf_layout_ != -1
  int64_t steps;           // DDA loop iterations
  int64_t* eps;            // epsilon-class counters or NULL

  // <PCL getCentroidIndexAt>: linear index, only the linear range is checked (quirk Q3)
  inline int64_t lin(const int32_t ijk[3]) const {
    // PCL computes this dot product in int (32-bit); keep the same wrap-free range by contract
    int32_t idx = (ijk[0] - min_b[0]) * mul[0] + (ijk[1] - min_b[1]) * mul[1] + (ijk[2] - min_b[2]) * mul[2];
    return idx;
  }
  inline bool occupied(const int32_t ijk[3]) const {
    int64_t idx = lin(ijk);
    if (idx < 0 || idx >= cells) return false;
    return occ[idx] != 0;
  }
  // include/cscpp/voxel_grid_occlusion_estimation.h:143-157 (cell centre; both branches equal ijk-min_b)
  inline void centre(const int32_t ijk[3], f32 o[3]) const {
    for (int c = 0; c < 3; ++c) {
      f32 base = b_min[c] + half;
      f32 off = (f32)(ijk[c] - min_b[c]) * leaf;
      o[c] = base + off;
    }
  }
  // voxel_grid_occlusion_estimation.cpp:174-238
  f32 ray_box(const f32 o[3], const f32 d[3]) const {
    f32 tmin, tmax, tymin, tymax, tzmin, tzmax;
    if (d[0] >= 0) { tmin = (b_min[0] - o[0]) / d[0]; tmax = (b_max[0] - o[0]) / d[0]; }
    else           { tmin = (b_max[0] - o[0]) / d[0]; tmax = (b_min[0] - o[0]) / d[0]; }
    if (d[1] >= 0) { tymin = (b_min[1] - o[1]) / d[1]; tymax = (b_max[1] - o[1]) / d[1]; }
    else           { tymin = (b_max[1] - o[1]) / d[1]; tymax = (b_min[1] - o[1]) / d[1]; }
    if ((tmin > tymax) || (tymin > tmax)) return -1.0f;
    if (tymin > tmin) tmin = tymin;
    if (tymax < tmax) tmax = tymax;
    if (d[2] >= 0) { tzmin = (b_min[2] - o[2]) / d[2]; tzmax = (b_max[2] - o[2]) / d[2]; }
    else           { tzmin = (b_max[2] - o[2]) / d[2]; tzmax = (b_min[2] - o[2]) / d[2]; }
    if ((tmin > tzmax) || (tzmin > tmax)) return -1.0f;
    if (tzmin > tmin) tmin = tzmin;
    if (tzmax < tmax) tmax = tzmax;
    return tmin;
  }
  // direction = (centre - origin).normalize() on a Vector4f with w = 0 - 0
  // voxel_grid_occlusion_estimation.cpp:111-113, occlusion_culling.cpp:338-339
  inline void direction(const f32 ctr[3], const f32 o[3], f32 d[3]) const {
    f32 v0 = ctr[0] - o[0], v1 = ctr[1] - o[1], v2 = ctr[2] - o[2], v3 = 0.0f - 0.0f;
    f32 n2 = dot4(v0 * v0, v1 * v1, v2 * v2, v3 * v3);
    f32 nrm = std::sqrt(n2);
    d[0] = v0 / nrm; d[1] = v1 / nrm; d[2] = v2 / nrm;
  }
  // voxel_grid_occlusion_estimation.cpp:410-582.  early_exit stops at the first qualifying
  // occluder (the return value can only stay 1 from there on).
  int traverse(const int32_t target[3], const f32 o[3], const f32 d[3], f32 tmin, bool early_exit) {
    f32 start[3];
    for (int c = 0; c < 3; ++c) { f32 td = tmin * d[c]; start[c] = o[c] + td; }  // :422
    int32_t ijk[3];
    for (int c = 0; c < 3; ++c) {  // :425, header :217-235 round = floor(d + 0.5f)
      f32 q = start[c] * inv;
      ijk[c] = (int32_t)std::floor(q + 0.5f);
    }
    f32 vm[3];
    centre(ijk, vm);  // :432
    int32_t step[3];
    for (int c = 0; c < 3; ++c) {  // :436-465
      if (d[c] >= 0) { vm[c] += half; step[c] = 1; }
      else           { vm[c] -= half; step[c] = -1; }
    }
    f32 t_max[3], t_delta[3];
    for (int c = 0; c < 3; ++c) {
      f32 q = (vm[c] - start[c]) / d[c];
      t_max[c] = tmin + q;               // :513-515
      t_delta[c] = leaf / std::fabs(d[c]);  // :517-519
    }
    f32 tc[3];
    centre(target, tc);
    const double thr = (double)(leaf * 2);  // :561 leaf_size_[0]*2 (float) promoted
    int result = 0;
    while ((ijk[0] <= max_b[0] + 1) && (ijk[0] + 1 >= min_b[0]) && (ijk[1] <= max_b[1] + 1) &&
           (ijk[1] + 1 >= min_b[1]) && (ijk[2] <= max_b[2] + 1) && (ijk[2] + 1 >= min_b[2])) {  // :538-540
      ++steps;
      if (ijk[0] == target[0] && ijk[1] == target[1] && ijk[2] == target[2]) break;  // :551-552
      if (occupied(ijk)) {  // :555-556
        f32 h[3];
        centre(ijk, h);
        f32 dx = h[0] - tc[0], dy = h[1] - tc[1], dz = h[2] - tc[2];
        f32 sx = dx * dx, sy = dy * dy, sz = dz * dz;
        f32 sxy = sx + sy;
        f32 sum = sxy + sz;
        double dist = std::sqrt((double)sum);  // :560
        if (eps) {
          int di = ijk[0] - target[0], dj = ijk[1] - target[1], dk = ijk[2] - target[2];
          if (di * di + dj * dj + dk * dk == 4) ++eps[ORC_EPS_E4];
        }
        if (dist > thr) {  // :561-562
          result = 1;
          if (early_exit) break;
        }
      }
      if (eps) {
        f32 a = t_max[0], b = t_max[1], cz = t_max[2];
        f32 lo = std::min(a, std::min(b, cz));
        f32 mid = std::max(std::min(a, b), std::min(std::max(a, b), cz));
        if (std::fabs(mid - lo) <= 1e-6f * std::max(std::fabs(lo), std::fabs(mid))) ++eps[ORC_EPS_E3];
      }
      if (t_max[0] <= t_max[1] && t_max[0] <= t_max[2]) {        // :565-569
        t_max[0] += t_delta[0]; ijk[0] += step[0];
      } else if (t_max[1] <= t_max[2] && t_max[1] <= t_max[0]) { // :570-574
        t_max[1] += t_delta[1]; ijk[1] += step[1];
      } else {                                                   // :575-579
        t_max[2] += t_delta[2]; ijk[2] += step[2];
      }
    }
    return result;
  }
  // voxel_grid_occlusion_estimation.cpp:99-128; returns -1 (state untouched), else 0/1
  int occlusion_estimation(const int32_t target[3], bool early_exit) {
    f32 ctr[3], d[3];
    centre(target, ctr);
    direction(ctr, origin, d);
    f32 tmin = ray_box(origin, d);
    if (tmin == -1) return -1;  // :117-122
    return traverse(target, origin, d, tmin, early_exit);
  }
};

struct Workspace {
  std::vector<int32_t> surv;
  std::vector<uint8_t> occ;
  std::vector<int8_t> memo_state;  // -2 unknown, -1 fail, 0, 1
  std::vector<int8_t> memo_gate;   // -1 unknown, 0, 1
  std::vector<int64_t> touched;
  std::vector<uint64_t> vox_seen;  // per-view coverage-voxel scratch
};

uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

inline bool bit_get(const uint64_t* b, int64_t i) { return (b[i >> 6] >> (i & 63)) & 1ull; }
inline void bit_set(uint64_t* b, int64_t i) { b[i >> 6] |= (1ull << (i & 63)); }

// occlusion_culling.cpp:265-373 for one viewpoint.
int eval_view(const Ctx& c, Workspace& ws, const double pose[7], int memo, const uint64_t* cov_pts,
              const uint64_t* cov_vox, int32_t* surv_out, int32_t* vis_out, int64_t* st, int64_t* eps,
              uint64_t* acc_pts, uint64_t* acc_vox, int64_t acc_min_visible) {
  for (int k = 0; k < ORC_ST_COUNT; ++k) st[k] = 0;
  f32 planes[24], T[3];
  frustum(c.p, pose, planes, T, nullptr);
  const f32* P = c.xyz.data();
  // 1. frustum cull, ascending index order (frustum_culling.cpp:150-171)
  ws.surv.clear();
  for (int64_t i = 0; i < c.n; ++i) {
    f32 x = P[3 * i], y = P[3 * i + 1], z = P[3 * i + 2];
    if (in_frustum(planes, x, y, z)) ws.surv.push_back((int32_t)i);
  }
  if (eps) {
    for (int64_t i = 0; i < c.n; ++i) {
      f32 x = P[3 * i], y = P[3 * i + 1], z = P[3 * i + 2];
      double pn = std::sqrt((double)x * x + (double)y * y + (double)z * z);
      bool near = false;
      for (int k = 0; k < 6 && !near; ++k) {
        const f32* pl = planes + 4 * k;
        double nn = std::sqrt((double)pl[0] * pl[0] + (double)pl[1] * pl[1] + (double)pl[2] * pl[2]);
        double dv = std::fabs((double)plane_dot(pl, x, y, z));
        if (dv <= 1e-4 * nn * std::max(1.0, pn)) near = true;
      }
      if (near) ++eps[ORC_EPS_E1];
    }
  }
  const int64_t S = (int64_t)ws.surv.size();
  st[ORC_ST_SURVIVORS] = S;
  uint64_t sh = 0;
  for (int64_t s = 0; s < S; ++s) sh += mix64((uint64_t)ws.surv[s]);
  st[ORC_ST_SURV_HASH] = (int64_t)sh;
  if (surv_out) std::memcpy(surv_out, ws.surv.data(), S * sizeof(int32_t));
  if (S == 0) return 0;  // empty cloud: VoxelGrid::filter yields nothing, loop at :317 has no iterations

  // 2. per-view grid over the survivors (occlusion_culling.cpp:300-306; <PCL applyFilter>)
  ViewGrid g;
  g.leaf = c.p.leaf; g.inv = c.inv_leaf; g.half = c.p.leaf * 0.5f;
  g.steps = 0; g.eps = eps;
  f32 mn[3], mx[3];
  for (int k = 0; k < 3; ++k) { mn[k] = P[3 * (int64_t)ws.surv[0] + k]; mx[k] = mn[k]; }
  for (int64_t s = 1; s < S; ++s) {
    const f32* q = P + 3 * (int64_t)ws.surv[s];
    for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], q[k]); mx[k] = std::max(mx[k], q[k]); }
  }
  for (int k = 0; k < 3; ++k) {
    g.min_b[k] = ifloor(mn[k] * g.inv);
    g.max_b[k] = ifloor(mx[k] * g.inv);
    g.div[k] = g.max_b[k] - g.min_b[k] + 1;
  }
  g.mul[0] = 1; g.mul[1] = g.div[0]; g.mul[2] = g.div[0] * g.div[1];
  g.cells = (int64_t)g.div[0] * g.div[1] * g.div[2];
  if (g.cells > 0x7fffffffLL) return -2;  // <PCL> "Leaf size is too small": unsupported
  for (int k = 0; k < 3; ++k) {
    g.b_min[k] = (f32)g.min_b[k] * g.leaf;        // voxel_grid_occlusion_estimation.cpp:53-55
    g.b_max[k] = (f32)(g.max_b[k] + 1) * g.leaf;  // :56-58
    g.origin[k] = T[k];                            // occlusion_culling.cpp:301, :63
  }
  st[ORC_ST_CELLS] = g.cells;
  if ((int64_t)ws.occ.size() < g.cells) {
    ws.occ.assign(g.cells, 0);
    ws.memo_state.assign(g.cells, -2);
    ws.memo_gate.assign(g.cells, -1);
  }
  ws.touched.clear();
  for (int64_t s = 0; s < S; ++s) {
    const f32* q = P + 3 * (int64_t)ws.surv[s];
    int32_t ijk[3] = {ifloor(q[0] * g.inv), ifloor(q[1] * g.inv), ifloor(q[2] * g.inv)};
    int64_t idx = g.lin(ijk);
    if (!ws.occ[idx]) { ws.occ[idx] = 1; ws.touched.push_back(idx); }
    if (eps) {
      bool e2 = false;
      for (int k = 0; k < 3; ++k) {
        f32 v = q[k] * g.inv;
        f32 fr = v - std::floor(v);
        if (fr <= 1e-4f || fr >= 1.0f - 1e-4f) e2 = true;
      }
      if (e2) ++eps[ORC_EPS_E2];
    }
  }
  g.occ = ws.occ.data();
  st[ORC_ST_UNIQUE] = (int64_t)ws.touched.size();

  // 3. per-survivor loop (occlusion_culling.cpp:308-370)
  const f32 po[3] = {c.p.pcd_origin[0], c.p.pcd_origin[1], c.p.pcd_origin[2]};
  {
    size_t words = (size_t)((c.n_cov + 63) / 64);
    if (ws.vox_seen.size() < words) ws.vox_seen.assign(words, 0);
  }
  std::vector<int32_t> local_vis;
  bool need_list = (acc_pts || acc_vox);
  int state = 0;  // uninitialised in the reference (:308); canonical start value 0 (quirk Q5)
  int64_t nvis = 0, carry = 0, gain_p = 0, gain_v = 0;
  uint64_t vh = 0;
  std::vector<int32_t> seen_ids;
  for (int64_t s = 0; s < S; ++s) {
    int32_t i = ws.surv[s];
    const f32* q = P + 3 * (int64_t)i;
    int32_t ijk[3] = {ifloor(q[0] * g.inv), ifloor(q[1] * g.inv), ifloor(q[2] * g.inv)};  // :320
    if (!g.occupied(ijk)) continue;  // :322,:329 index != -1
    int64_t idx = g.lin(ijk);
    int r;
    if (memo) {
      if (ws.memo_state[idx] == -2) ws.memo_state[idx] = (int8_t)g.occlusion_estimation(ijk, true);
      r = ws.memo_state[idx];
    } else {
      r = g.occlusion_estimation(ijk, false);  // :332
    }
    if (r >= 0) state = r; else ++carry;  // ret = -1 leaves state untouched
    if (state != 1) {  // :335
      int gate;
      if (memo && ws.memo_gate[idx] >= 0) {
        gate = ws.memo_gate[idx];
      } else {
        f32 ctr[3], d[3];
        g.centre(ijk, ctr);
        g.direction(ctr, po, d);         // :338-339
        f32 tmin = g.ray_box(po, d);     // :341
        gate = (tmin != -1) ? 1 : 0;     // :342
        if (memo) ws.memo_gate[idx] = (int8_t)gate;
      }
      if (gate) {  // :362
        if (vis_out) vis_out[nvis] = i;
        if (need_list) local_vis.push_back(i);
        ++nvis;
        vh += mix64((uint64_t)i);
        if (cov_pts && !bit_get(cov_pts, i)) ++gain_p;
        {
          int32_t id = c.cov_id[i];
          if (!bit_get(ws.vox_seen.data(), id)) {
            bit_set(ws.vox_seen.data(), id);
            seen_ids.push_back(id);
            if (!cov_vox || !bit_get(cov_vox, id)) ++gain_v;
          }
        }
      }
    }
  }
  if (!cov_pts) gain_p = nvis;
  for (int32_t id : seen_ids) ws.vox_seen[id >> 6] &= ~(1ull << (id & 63));
  if (eps) eps[ORC_EPS_E5] += carry;
  st[ORC_ST_VISIBLE] = nvis;
  st[ORC_ST_CARRY] = carry;
  st[ORC_ST_STEPS] = g.steps;
  st[ORC_ST_GAIN_PTS] = gain_p;
  st[ORC_ST_GAIN_VOX] = gain_v;
  st[ORC_ST_HASH] = (int64_t)vh;
  if (need_list && nvis > acc_min_visible) {
    // set union: coverage_path_planning_heuristic.cpp:344-346 (globalCloud += pts if pts.size()>5)
    for (int32_t i : local_vis) {
      if (acc_pts) __atomic_fetch_or(&acc_pts[i >> 6], 1ull << (i & 63), __ATOMIC_RELAXED);
      if (acc_vox) {
        int32_t id = c.cov_id[i];
        __atomic_fetch_or(&acc_vox[id >> 6], 1ull << (id & 63), __ATOMIC_RELAXED);
      }
    }
  }
  // reset the touched cells for the next view handled by this workspace
  for (int64_t idx : ws.touched) { ws.occ[idx] = 0; ws.memo_state[idx] = -2; ws.memo_gate[idx] = -1; }
  return 0;
}

template <class F>
void parallel_views(int64_t nviews, int nthreads, F&& fn) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads == 1 || nviews < 2) {
    Workspace ws;
    for (int64_t v = 0; v < nviews; ++v) fn(ws, v);
    return;
  }
  std::atomic<int64_t> next(0);
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t)
    th.emplace_back([&]() {
      Workspace ws;
      for (;;) {
        int64_t v = next.fetch_add(1);
        if (v >= nviews) break;
        fn(ws, v);
      }
    });
  for (auto& t : th) t.join();
}

// <tf createQuaternionFromRPY> / <tf Matrix3x3::getRotation> helpers for sspp/src/sensors.cpp
void quat_to_mat(const double q[4], double R[3][3]) {
  double x = q[0], y = q[1], z = q[2], w = q[3];
  double d = x * x + y * y + z * z + w * w, s = 2.0 / d;
  double xs = x * s, ys = y * s, zs = z * s, wx = w * xs, wy = w * ys, wz = w * zs;
  double xx = x * xs, xy = x * ys, xz = x * zs, yy = y * ys, yz = y * zs, zz = z * zs;
  R[0][0] = 1.0 - (yy + zz); R[0][1] = xy - wz; R[0][2] = xz + wy;
  R[1][0] = xy + wz; R[1][1] = 1.0 - (xx + zz); R[1][2] = yz - wx;
  R[2][0] = xz - wy; R[2][1] = yz + wx; R[2][2] = 1.0 - (xx + yy);
}
void mat_to_quat(const double m[3][3], double q[4]) {
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
  q[0] = t[0]; q[1] = t[1]; q[2] = t[2]; q[3] = t[3];
}


// <tf Matrix3x3::getRPY> = getEulerYPR(yaw, pitch, roll, solution 1) on the matrix of <tf Matrix3x3::setRotation>;
// called at occlusion_culling.cpp:497-500 and :558-561.  Restated from the published tf algorithm (tf is not
// vendored under /root/reference).  All double, libm asin / atan2 / cos.
void pose_rpy(const double pose[7], double rpy[3]) {
  double R[3][3];
  quat_to_mat(pose + 3, R);
  const double pi = 3.1415926535897932384626433832795029;  // TFSIMD_PI
  double yaw, pitch, roll;
  if (std::fabs(R[2][0]) >= 1) {
    yaw = 0;
    if (R[2][0] < 0) {
      pitch = pi / 2.0;
      roll = std::atan2(R[0][1], R[0][2]);
    } else {
      pitch = -pi / 2.0;
      roll = std::atan2(-R[0][1], -R[0][2]);
    }
  } else {
    pitch = -std::asin(R[2][0]);
    roll = std::atan2(R[2][1] / std::cos(pitch), R[2][2] / std::cos(pitch));
    yaw = std::atan2(R[1][0] / std::cos(pitch), R[0][0] / std::cos(pitch));
  }
  rpy[0] = roll; rpy[1] = pitch; rpy[2] = yaw;
}

// OcclusionCulling::pointCloudViewportTransform (occlusion_culling.cpp:538-618) for one point: translate by
// -camera position, rotate about z by -yaw, about y by -pitch.  Every stage goes through transformPointMatVec
// (:521-531): double arithmetic on float32 inputs (geometry_msgs::Point32), result narrowed back to float32.
struct ViewportXf {
  double t[3];          // cameraPoseTrans (:545)
  double cz, sz, cy, sy;  // cos / sin of yaw' = -yaw (:562-569) and of pitch' = -pitch (:579-584)
};
ViewportXf viewport_xf(const double pose[7]) {
  ViewportXf x;
  x.t[0] = -1 * pose[0]; x.t[1] = -1 * pose[1]; x.t[2] = -1 * pose[2];
  double rpy[3];
  pose_rpy(pose, rpy);
  const double yaw = -1 * rpy[2], pitch = -1 * rpy[1];
  x.cz = std::cos(yaw); x.sz = std::sin(yaw);
  x.cy = std::cos(pitch); x.sy = std::sin(pitch);
  return x;
}
inline void viewport_point(const ViewportXf& x, const f32 in[3], f32 out[3]) {
  // stage 1 (:594): identity rotation + translation; 1*a + 0*b + 0*c + t evaluated left to right in double
  f32 a0 = (f32)(1.0 * in[0] + 0.0 * in[1] + 0.0 * in[2] + x.t[0]);
  f32 a1 = (f32)(0.0 * in[0] + 1.0 * in[1] + 0.0 * in[2] + x.t[1]);
  f32 a2 = (f32)(0.0 * in[0] + 0.0 * in[1] + 1.0 * in[2] + x.t[2]);
  // stage 2 (:598): rotZ = [c -s 0; s c 0; 0 0 1], no translation
  f32 b0 = (f32)(x.cz * a0 + (-1 * x.sz) * a1 + 0.0 * a2 + 0.0);
  f32 b1 = (f32)(x.sz * a0 + x.cz * a1 + 0.0 * a2 + 0.0);
  f32 b2 = (f32)(0.0 * a0 + 0.0 * a1 + 1.0 * a2 + 0.0);
  // stage 3 (:602): rotY = [c 0 s; 0 1 0; -s 0 c]
  out[0] = (f32)(x.cy * b0 + 0.0 * b1 + x.sy * b2 + 0.0);
  out[1] = (f32)(0.0 * b0 + 1.0 * b1 + 0.0 * b2 + 0.0);
  out[2] = (f32)((-1 * x.sy) * b0 + 0.0 * b1 + x.cy * b2 + 0.0);
}

}  // namespace

struct OrcCtx { Ctx c; };

extern "C" {

uint64_t orc_mix64(uint64_t x) { return mix64(x); }

int orc_create(const float* xyz, int64_t n, const OrcParams* p, OrcCtx** out) {
  if (!xyz || !p || !out || n < 0 || n > 0x7fffffffLL) return -1;
  if (!(p->leaf > 0) || !(p->cov_leaf > 0)) return -1;
  OrcCtx* o = new OrcCtx();
  Ctx& c = o->c;
  c.n = n;
  c.xyz.assign(xyz, xyz + 3 * n);
  c.p = *p;
  c.inv_leaf = 1.0f / p->leaf;
  c.inv_cov = 1.0f / p->cov_leaf;
  // model grids (occlusion_culling.cpp:41-65): only occupancy on a lattice matters; compact ids in
  // ascending (k, j, i) order of the linearised model grid.
  auto build = [&](f32 inv, std::vector<int32_t>* ids, int64_t* count, int32_t* mnb, int32_t* mxb,
                   std::vector<int32_t>* vox_ijk) -> int {
    if (n == 0) { *count = 0; return 0; }
    int32_t lo[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, hi[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
    for (int64_t i = 0; i < n; ++i)
      for (int k = 0; k < 3; ++k) {
        int32_t v = ifloor(xyz[3 * i + k] * inv);
        lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v);
      }
    int64_t dx = (int64_t)hi[0] - lo[0] + 1, dy = (int64_t)hi[1] - lo[1] + 1, dz = (int64_t)hi[2] - lo[2] + 1;
    if (dx * dy * dz > 0x7fffffffLL) return -2;  // <PCL> "Leaf size is too small"
    std::vector<int64_t> key(n);
    for (int64_t i = 0; i < n; ++i) {
      int64_t a = ifloor(xyz[3 * i] * inv) - lo[0], b = ifloor(xyz[3 * i + 1] * inv) - lo[1],
              cc = ifloor(xyz[3 * i + 2] * inv) - lo[2];
      key[i] = a + dx * (b + dy * cc);
    }
    std::vector<int64_t> uniq(key);
    std::sort(uniq.begin(), uniq.end());
    uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
    *count = (int64_t)uniq.size();
    if (ids) {
      ids->resize(n);
      for (int64_t i = 0; i < n; ++i)
        (*ids)[i] = (int32_t)(std::lower_bound(uniq.begin(), uniq.end(), key[i]) - uniq.begin());
    }
    if (mnb) for (int k = 0; k < 3; ++k) { mnb[k] = lo[k]; if (mxb) mxb[k] = hi[k]; }
    if (vox_ijk) {
      vox_ijk->resize(3 * uniq.size());
      for (size_t u = 0; u < uniq.size(); ++u) {
        (*vox_ijk)[3 * u] = (int32_t)(uniq[u] % dx) + lo[0];
        (*vox_ijk)[3 * u + 1] = (int32_t)((uniq[u] / dx) % dy) + lo[1];
        (*vox_ijk)[3 * u + 2] = (int32_t)(uniq[u] / (dx * dy)) + lo[2];
      }
    }
    return 0;
  };
  int rc = build(c.inv_cov, &c.cov_id, &c.n_cov, c.cov_min_b, nullptr, &c.cov_ijk);
  if (rc == 0) rc = build(c.inv_leaf, nullptr, &c.n_occ, c.min_b, c.max_b, nullptr);
  if (rc != 0) { delete o; return rc; }
  *out = o;
  return 0;
}

void orc_destroy(OrcCtx* c) { delete c; }
int64_t orc_num_points(const OrcCtx* c) { return c->c.n; }
int64_t orc_num_cov_voxels(const OrcCtx* c) { return c->c.n_cov; }
int64_t orc_num_occ_voxels(const OrcCtx* c) { return c->c.n_occ; }
void orc_model_box(const OrcCtx* c, int32_t min_b[3], int32_t max_b[3]) {
  for (int k = 0; k < 3; ++k) { min_b[k] = c->c.min_b[k]; max_b[k] = c->c.max_b[k]; }
}
void orc_point_cov_ids(const OrcCtx* c, int32_t* out) {
  std::memcpy(out, c->c.cov_id.data(), c->c.cov_id.size() * sizeof(int32_t));
}

void orc_planes(const OrcParams* p, const double pose[7], float planes[24], float T[3]) {
  frustum(*p, pose, planes, T, nullptr);
}
void orc_corners(const OrcParams* p, const double pose[7], float corners[24]) {
  float planes[24], T[3];
  frustum(*p, pose, planes, T, corners);
}

int orc_eval_view(OrcCtx* c, const double pose[7], int memo, const uint64_t* covered_pts,
                  const uint64_t* covered_vox, int32_t* surv_out, int32_t* vis_out,
                  int64_t stats[ORC_ST_COUNT], int64_t* eps) {
  Workspace ws;
  return eval_view(c->c, ws, pose, memo, covered_pts, covered_vox, surv_out, vis_out, stats, eps, nullptr,
                   nullptr, 0);
}

int orc_eval_batch(OrcCtx* c, const double* poses, int64_t nviews, int memo, int nthreads,
                   const uint64_t* covered_pts, const uint64_t* covered_vox, int64_t* stats) {
  std::atomic<int> rc(0);
  parallel_views(nviews, nthreads, [&](Workspace& ws, int64_t v) {
    int r = eval_view(c->c, ws, poses + 7 * v, memo, covered_pts, covered_vox, nullptr, nullptr,
                      stats + v * ORC_ST_COUNT, nullptr, nullptr, nullptr, 0);
    if (r != 0) rc = r;
  });
  return rc;
}

int orc_accumulate(OrcCtx* c, const double* poses, int64_t nviews, int nthreads, int64_t min_visible,
                   uint64_t* covered_pts, uint64_t* covered_vox) {
  std::atomic<int> rc(0);
  parallel_views(nviews, nthreads, [&](Workspace& ws, int64_t v) {
    int64_t st[ORC_ST_COUNT];
    int r = eval_view(c->c, ws, poses + 7 * v, 1, nullptr, nullptr, nullptr, nullptr, st, nullptr, covered_pts,
                      covered_vox, min_visible);
    if (r != 0) rc = r;
  });
  return rc;
}

float orc_coverage_percent_vox(const OrcCtx* c, const uint64_t* covered_vox) {
  // occlusion_culling.cpp:401,431-432: float counters, float ratio, float * 100
  float matched = (float)orc_popcount(covered_vox, c->c.n_cov);
  float orig = (float)c->c.n_cov;
  float ratio = matched / orig;
  return ratio * 100;
}

float orc_coverage_percent_idx(const OrcCtx* c, const int32_t* idx, int64_t n) {
  // occlusion_culling.cpp:387-428: voxelise the cloud, match each occupied cell against the model
  // grid.  Both grids share the lattice floor(x*inv), so Matched = |voxels(cloud) ^ voxels(model)|.
  std::vector<uint64_t> bits((size_t)((c->c.n_cov + 63) / 64), 0);
  for (int64_t k = 0; k < n; ++k) bit_set(bits.data(), c->c.cov_id[idx[k]]);
  return orc_coverage_percent_vox(c, bits.data());
}

int64_t orc_popcount(const uint64_t* bits, int64_t nbits) {
  int64_t total = 0, words = nbits / 64;
  for (int64_t w = 0; w < words; ++w) total += __builtin_popcountll(bits[w]);
  if (nbits % 64) total += __builtin_popcountll(bits[words] & ((1ull << (nbits % 64)) - 1));
  return total;
}

void orc_robot2sensor(const double robot_pose[7], const double offset[3], const double rpy[3],
                      double sensor_pose[7]) {
  // sspp/src/sensors.cpp:47-90, all double
  double R1[3][3], R2[3][3];
  quat_to_mat(robot_pose + 3, R1);
  // <tf createQuaternionFromRPY> -> setRPY(roll, pitch, yaw)
  double hy = rpy[2] * 0.5, hp = rpy[1] * 0.5, hr = rpy[0] * 0.5;
  double cy = std::cos(hy), sy = std::sin(hy), cp = std::cos(hp), sp = std::sin(hp), cr = std::cos(hr),
         sr = std::sin(hr);
  double q2[4] = {sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy,
                  cr * cp * cy + sr * sp * sy};
  quat_to_mat(q2, R2);
  double M[3][3], T[3];
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) M[i][j] = R1[i][0] * R2[0][j] + R1[i][1] * R2[1][j] + R1[i][2] * R2[2][j];
    T[i] = R1[i][0] * offset[0] + R1[i][1] * offset[1] + R1[i][2] * offset[2] + robot_pose[i];
  }
  // sensor2sensorMat (:76-79): columns (1,0,0), (0,0,1), (0,-1,0)
  double F[3][3] = {{1, 0, 0}, {0, 0, -1}, {0, 1, 0}};
  double N[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) N[i][j] = M[i][0] * F[0][j] + M[i][1] * F[1][j] + M[i][2] * F[2][j];
  double q[4];
  mat_to_quat(N, q);
  sensor_pose[0] = T[0]; sensor_pose[1] = T[1]; sensor_pose[2] = T[2];
  sensor_pose[3] = q[0]; sensor_pose[4] = q[1]; sensor_pose[5] = q[2]; sensor_pose[6] = q[3];
}

void orc_pose_rpy(const double pose[7], double rpy[3]) { pose_rpy(pose, rpy); }

void orc_viewport_transform(const double pose[7], const float* xyz, int64_t n, float* out) {
  const ViewportXf x = viewport_xf(pose);
  for (int64_t i = 0; i < n; ++i) viewport_point(x, xyz + 3 * i, out + 3 * i);
}

void orc_accuracy_idx(const OrcCtx* c, const double pose[7], const int32_t* idx, int64_t n, double* err_sum_depth,
                      double* err_sum_z, float* depth_min, float* depth_max) {
  // calcAvgAccuracy(cloud, pose) occlusion_culling.cpp:459-475 and calcAvgAccuracy(cloud) :445-458: sums run over
  // the cloud in its own order (extractVisibleSurface returns ascending original index); the depth extrema are the
  // quantities SSMaxMinAccuracy folds over the whole search space (:505-513)
  const ViewportXf x = viewport_xf(pose);
  double sd = 0, sz = 0;
  float lo = INFINITY, hi = -INFINITY;
  for (int64_t k = 0; k < n; ++k) {
    const f32* p = c->c.xyz.data() + 3 * (int64_t)idx[k];
    f32 q[3];
    viewport_point(x, p, q);
    double val = q[0];
    double e = 0.0000285 * val * val;  // :469
    sd += e;
    double vz = p[2];
    sz += 0.0000285 * vz * vz;         // :452
    lo = std::min(lo, q[0]);
    hi = std::max(hi, q[0]);
  }
  *err_sum_depth = sd; *err_sum_z = sz; *depth_min = lo; *depth_max = hi;
}

int64_t orc_uncovered_ex(const OrcCtx* oc, const uint64_t* covered_vox, int pcl_aliasing, int32_t* ids, float* centres) {
  // pointCloudDiff (coverage_path_planning_heuristic.cpp:88-145): scan the model's coverage grid in k, j, i order
  // (= ascending compact id) and keep the occupied cells whose centre falls in no covered voxel.  Both grids share
  // the lattice floor(x * inv_leaf), so "centre not in a covered voxel" is "bit not set" -- except for quirk Q11
  // (pcl_aliasing): the lookup into the covered cloud's grid is <PCL getCentroidIndexAt> (:122), which linearises
  // (ijk - min_b) . (1, div0, div0 div1) with THAT grid's box and checks only 0 <= idx < size, so a model cell
  // outside the box can read the entry of another, covered cell and is then not reported.
  // Centre = getCentroidCoordinate (voxel_grid_occlusion_estimation.h:143-157) on the model grid.
  const Ctx& c = oc->c;
  const f32 leaf = c.p.cov_leaf;
  auto covered = [&](int64_t u) { return covered_vox && ((covered_vox[u >> 6] >> (u & 63)) & 1ull); };
  int mn[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, mx[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
  std::map<std::array<int32_t, 3>, int64_t> index;
  if (pcl_aliasing)
    for (int64_t u = 0; u < c.n_cov; ++u) {
      index[{c.cov_ijk[3 * u], c.cov_ijk[3 * u + 1], c.cov_ijk[3 * u + 2]}] = u;
      if (!covered(u)) continue;
      for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], c.cov_ijk[3 * u + k]); mx[k] = std::max(mx[k], c.cov_ijk[3 * u + k]); }
    }
  const bool any = pcl_aliasing && mn[0] <= mx[0];
  const int d0 = any ? mx[0] - mn[0] + 1 : 0, d1 = any ? mx[1] - mn[1] + 1 : 0, d2 = any ? mx[2] - mn[2] + 1 : 0;
  int64_t n = 0;
  for (int64_t u = 0; u < c.n_cov; ++u) {
    if (covered(u)) continue;
    if (any) {
      const int lin = (c.cov_ijk[3 * u] - mn[0]) + (c.cov_ijk[3 * u + 1] - mn[1]) * d0 + (c.cov_ijk[3 * u + 2] - mn[2]) * d0 * d1;
      if (lin >= 0 && lin < d0 * d1 * d2) {
        const int a2 = lin / (d0 * d1), rem = lin - a2 * d0 * d1, a1 = rem / d0, a0 = rem - a1 * d0;
        auto it = index.find({a0 + mn[0], a1 + mn[1], a2 + mn[2]});
        if (it != index.end() && covered(it->second)) continue;  // leaf_layout_[lin] != -1
      }
    }
    if (ids) ids[n] = (int32_t)u;
    if (centres)
      for (int k = 0; k < 3; ++k) {
        f32 b_min = (f32)c.cov_min_b[k] * leaf;
        centres[3 * n + k] = (b_min + leaf * 0.5f) + (f32)(c.cov_ijk[3 * u + k] - c.cov_min_b[k]) * leaf;
      }
    ++n;
  }
  return n;
}

int64_t orc_uncovered(const OrcCtx* oc, const uint64_t* covered_vox, int32_t* ids, float* centres) {
  return orc_uncovered_ex(oc, covered_vox, 0, ids, centres);
}

int32_t orc_euclidean_clusters(const float* xyz, int64_t n, double tol, int32_t min_size, int32_t max_size,
                               int32_t* labels, int64_t* ties) {
  // clusteringPointCloud (coverage_path_planning_heuristic.cpp:147-190) = <PCL extractEuclideanClusters>: for every
  // unprocessed point in index order grow a region breadth-first through radius searches; keep it when
  // min <= size <= max; finally sort the clusters largest first.  <FLANN radiusSearch>: squared float32 distance
  // strictly below float(tol * tol).  Restated from the published algorithms (PCL / FLANN are not vendored).
  // The neighbour search is a uniform hash grid of cell size tol; the distance test is literal.
  const f32 r2 = (f32)(tol * tol);
  if (ties) *ties = 0;
  for (int64_t i = 0; i < n; ++i) labels[i] = -1;
  if (n == 0) return 0;
  f32 lo[3] = {xyz[0], xyz[1], xyz[2]};
  for (int64_t i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) lo[k] = std::min(lo[k], xyz[3 * i + k]);
  auto cell = [&](int64_t i, int k) { return (int64_t)std::floor(((double)xyz[3 * i + k] - lo[k]) / tol); };
  std::vector<std::pair<uint64_t, int32_t>> grid(n);
  auto pack = [](int64_t a, int64_t b, int64_t c) {
    return ((uint64_t)(a + 1) << 42) | ((uint64_t)(b + 1) << 21) | (uint64_t)(c + 1);
  };
  for (int64_t i = 0; i < n; ++i) grid[i] = std::make_pair(pack(cell(i, 0), cell(i, 1), cell(i, 2)), (int32_t)i);
  std::sort(grid.begin(), grid.end());
  std::vector<char> processed(n, 0);
  std::vector<std::vector<int32_t>> clusters;
  std::vector<int32_t> queue;
  for (int64_t s = 0; s < n; ++s) {
    if (processed[s]) continue;
    queue.clear();
    queue.push_back((int32_t)s);
    processed[s] = 1;
    for (size_t q = 0; q < queue.size(); ++q) {
      const int32_t i = queue[q];
      const int64_t ca = cell(i, 0), cb = cell(i, 1), cc = cell(i, 2);
      for (int64_t a = ca - 1; a <= ca + 1; ++a)
        for (int64_t b = cb - 1; b <= cb + 1; ++b)
          for (int64_t c = cc - 1; c <= cc + 1; ++c) {
            if (a < 0 || b < 0 || c < 0) continue;
            const uint64_t key = pack(a, b, c);
            auto it = std::lower_bound(grid.begin(), grid.end(), std::make_pair(key, (int32_t)INT32_MIN));
            for (; it != grid.end() && it->first == key; ++it) {
              const int32_t j = it->second;
              f32 d0 = xyz[3 * i] - xyz[3 * j], d1 = xyz[3 * i + 1] - xyz[3 * j + 1], d2 = xyz[3 * i + 2] - xyz[3 * j + 2];
              f32 dist = d0 * d0 + d1 * d1 + d2 * d2;  // <FLANN L2_Simple<float>>
              if (ties && dist == r2 && i < j) ++*ties;
              if (!(dist < r2) || processed[j]) continue;
              processed[j] = 1;
              queue.push_back(j);
            }
          }
    }
    if ((int64_t)queue.size() >= min_size && (int64_t)queue.size() <= max_size) clusters.push_back(queue);
  }
  // largest first (PCL's comparePointClusters); equal sizes: by first member, which PCL leaves unspecified
  std::stable_sort(clusters.begin(), clusters.end(),
                   [](const std::vector<int32_t>& a, const std::vector<int32_t>& b) { return a.size() > b.size(); });
  for (size_t k = 0; k < clusters.size(); ++k)
    for (int32_t i : clusters[k]) labels[i] = (int32_t)k;
  return (int32_t)clusters.size();
}

void orc_cluster_bb(const float* xyz, int64_t n, float leaf, double grid_start[3], double grid_size[3]) {
  // findClusterBB (coverage_path_planning_heuristic.cpp:218-246): voxel grid over the cluster at `leaf`, centres of
  // its min and max cells, box grown by 5 m in x, y and 3 m in z
  const f32 inv = 1.0f / leaf;
  int32_t mn[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, mx[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
  for (int64_t i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) {
      int32_t v = ifloor(xyz[3 * i + k] * inv);
      mn[k] = std::min(mn[k], v); mx[k] = std::max(mx[k], v);
    }
  f32 cmin[3], cmax[3];
  for (int k = 0; k < 3; ++k) {
    f32 b_min = (f32)mn[k] * leaf;
    cmin[k] = (b_min + leaf * 0.5f) + (f32)0 * leaf;
    cmax[k] = (b_min + leaf * 0.5f) + (f32)(mx[k] - mn[k]) * leaf;
  }
  const double maximizeSizeXY = 5, maximizeSizeZ = 3;
  grid_size[0] = std::abs(cmax[0] - cmin[0]) + maximizeSizeXY;
  grid_size[1] = std::abs(cmax[1] - cmin[1]) + maximizeSizeXY;
  grid_size[2] = std::abs(cmax[2] - cmin[2]) + maximizeSizeZ;
  grid_start[0] = cmin[0] - double(maximizeSizeXY / 2);
  grid_start[1] = cmin[1] - double(maximizeSizeXY / 2);
  grid_start[2] = cmin[2];
}

// ---- mesh filters (coverage_path_planning_heuristic.cpp:300-335) ------------------------------------------
// The reference asks a CGAL AABB tree (Simple_cartesian<double>): number_of_intersected_primitives(ray | segment)
// and squared_distance(point).  CGAL is not vendored; restated with textbook double-precision geometry: every
// triangle is visited, a ray / segment hits a triangle when the division-free Moeller-Trumbore parameters are in
// range with inclusive borders, the distance is to the closest point found by Voronoi regions.  Queries grazing
// an edge or a vertex may count differently in CGAL: PARITY UNPINNED for those (reported by orc_mesh_grazing).
struct OrcMesh { std::vector<double> tri; };  // 9 doubles per triangle

static inline double mdot(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

static bool mesh_hit(const double* o, const double* d, const double* t, bool segment, double* margin) {
  double e1[3] = {t[3] - t[0], t[4] - t[1], t[5] - t[2]}, e2[3] = {t[6] - t[0], t[7] - t[1], t[8] - t[2]};
  double h[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
  double det = mdot(e1, h);
  if (det == 0.0) return false;
  double sv[3] = {o[0] - t[0], o[1] - t[1], o[2] - t[2]};
  double u = mdot(sv, h);
  double q[3] = {sv[1] * e1[2] - sv[2] * e1[1], sv[2] * e1[0] - sv[0] * e1[2], sv[0] * e1[1] - sv[1] * e1[0]};
  double v = mdot(d, q), w = mdot(e2, q);
  if (det < 0.0) { det = -det; u = -u; v = -v; w = -w; }
  if (margin) {
    // normalised distance of the barycentric parameters from the triangle's border (for the grazing report)
    double m = std::min(std::min(u, v), det - (u + v)) / det;
    *margin = std::fabs(m);
  }
  if (u < 0.0 || v < 0.0 || u + v > det || w < 0.0) return false;
  return !segment || w <= det;
}

static double mesh_sqdist(const double* p, const double* t) {
  const double *a = t, *b = t + 3, *c = t + 6;
  double ab[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, ac[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
  double ap[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
  double cl[3];
  double d1 = mdot(ab, ap), d2 = mdot(ac, ap);
  auto fin = [&]() { double e[3] = {p[0] - cl[0], p[1] - cl[1], p[2] - cl[2]}; return mdot(e, e); };
  if (d1 <= 0.0 && d2 <= 0.0) { cl[0] = a[0]; cl[1] = a[1]; cl[2] = a[2]; return fin(); }
  double bp[3] = {p[0] - b[0], p[1] - b[1], p[2] - b[2]};
  double d3 = mdot(ab, bp), d4 = mdot(ac, bp);
  if (d3 >= 0.0 && d4 <= d3) { cl[0] = b[0]; cl[1] = b[1]; cl[2] = b[2]; return fin(); }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
    double v = d1 / (d1 - d3);
    for (int k = 0; k < 3; ++k) cl[k] = a[k] + v * ab[k];
    return fin();
  }
  double cp[3] = {p[0] - c[0], p[1] - c[1], p[2] - c[2]};
  double d5 = mdot(ab, cp), d6 = mdot(ac, cp);
  if (d6 >= 0.0 && d5 <= d6) { cl[0] = c[0]; cl[1] = c[1]; cl[2] = c[2]; return fin(); }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
    double w = d2 / (d2 - d6);
    for (int k = 0; k < 3; ++k) cl[k] = a[k] + w * ac[k];
    return fin();
  }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
    double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    for (int k = 0; k < 3; ++k) cl[k] = b[k] + w * (c[k] - b[k]);
    return fin();
  }
  double denom = 1.0 / (va + vb + vc);
  double v = vb * denom, w = vc * denom;
  for (int k = 0; k < 3; ++k) cl[k] = a[k] + ab[k] * v + ac[k] * w;
  return fin();
}

OrcMesh* orc_mesh_create(const float* vertices, int64_t nv, const int32_t* triangles, int64_t nt) {
  OrcMesh* m = new OrcMesh();
  for (int64_t k = 0; k < nt; ++k) {
    double t[9];
    bool ok = true;
    for (int c = 0; c < 3; ++c) {
      int32_t vi = triangles[3 * k + c];
      if (vi < 0 || vi >= nv) { ok = false; break; }
      for (int a = 0; a < 3; ++a) t[3 * c + a] = (double)vertices[3 * (int64_t)vi + a];
    }
    if (!ok) { delete m; return nullptr; }
    // loadOBJFile :1409 / :1430 skips collinear triangles
    double u[3] = {t[3] - t[0], t[4] - t[1], t[5] - t[2]}, v[3] = {t[6] - t[0], t[7] - t[1], t[8] - t[2]};
    if (u[1] * v[2] - u[2] * v[1] == 0.0 && u[2] * v[0] - u[0] * v[2] == 0.0 && u[0] * v[1] - u[1] * v[0] == 0.0) continue;
    m->tri.insert(m->tri.end(), t, t + 9);
  }
  return m;
}
void orc_mesh_destroy(OrcMesh* m) { delete m; }
int64_t orc_mesh_num_triangles(const OrcMesh* m) { return (int64_t)m->tri.size() / 9; }

void orc_mesh_filter(const OrcMesh* m, const double* pos, int64_t n, const double through[3], double min_dist,
                     double max_dist, int nthreads, int32_t* hits, double* sqd, uint8_t* pass, double* graze) {
  // isFilteringConditionSatisfied :319-333
  const int64_t nt = (int64_t)m->tri.size() / 9;
  parallel_views(n, nthreads, [&](Workspace&, int64_t i) {
    const double* a = pos + 3 * i;
    double d[3] = {through[0] - a[0], through[1] - a[1], through[2] - a[2]};
    int32_t cnt = 0;
    double best = INFINITY, gz = INFINITY;
    for (int64_t k = 0; k < nt; ++k) {
      double mg = INFINITY;
      if (mesh_hit(a, d, &m->tri[9 * k], false, &mg)) ++cnt;
      gz = std::min(gz, mg);
      best = std::min(best, mesh_sqdist(a, &m->tri[9 * k]));
    }
    if (hits) hits[i] = cnt;
    if (sqd) sqd[i] = best;
    if (graze) graze[i] = gz;
    if (pass) pass[i] = (cnt % 2 != 1) && best >= (min_dist * min_dist) && best <= (max_dist * max_dist);
  });
}

void orc_mesh_segments_free(const OrcMesh* m, const double* a, const double* b, int64_t n, int nthreads, uint8_t* out) {
  // isConnectionConditionSatisfied :300-315
  const int64_t nt = (int64_t)m->tri.size() / 9;
  parallel_views(n, nthreads, [&](Workspace&, int64_t i) {
    double d[3] = {b[3 * i] - a[3 * i], b[3 * i + 1] - a[3 * i + 1], b[3 * i + 2] - a[3 * i + 2]};
    int32_t cnt = 0;
    for (int64_t k = 0; k < nt; ++k)
      if (mesh_hit(a + 3 * i, d, &m->tri[9 * k], true, nullptr)) ++cnt;
    out[i] = cnt == 0;
  });
}

}  // extern "C"
