// preprocess.cpp -- host-side model layout for the B200 path.
//
// The reference builds a PCL voxel grid over the model once (occlusion_culling.cpp:41-65) and then
// scans all N points for every viewpoint (frustum_culling.cpp:150-171).  Here the cloud is laid out
// once so that a viewpoint only touches the points near its frustum:
//   * points sorted by the Morton code of their occlusion-lattice voxel (ties: original index),
//   * a voxel table {i, j, k, first point} in the same order,
//   * leaf buckets ("chunks") of whole voxels holding <= kChunkMaxPoints points with the tight
//     AABB of their points, and super-chunks of kSuperFanout chunks.
// Voxel coordinates use the reference's lattice: ijk = floor(x * inverse_leaf) in float32
// (<PCL VoxelGrid::getGridCoordinates>), so per-view occupancy is a subset of this table.
#include "preprocess.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>

namespace asscpp {

namespace {

inline uint64_t spread21(uint64_t v) {
  v &= 0x1fffffull;
  v = (v | v << 32) & 0x1f00000000ffffull;
  v = (v | v << 16) & 0x1f0000ff0000ffull;
  v = (v | v << 8) & 0x100f00f00f00f00full;
  v = (v | v << 4) & 0x10c30c30c30c30c3ull;
  v = (v | v << 2) & 0x1249249249249249ull;
  return v;
}
inline uint64_t morton3(uint32_t i, uint32_t j, uint32_t k) {
  return spread21(i) | (spread21(j) << 1) | (spread21(k) << 2);
}

struct KeyIdx {
  uint64_t key;
  uint32_t idx;
};

template <class It, class Cmp>
void parallel_sort(It begin, It end, Cmp cmp) {
  const int64_t n = end - begin;
  unsigned hw = std::thread::hardware_concurrency();
  int parts = (n < (1 << 18) || hw < 2) ? 1 : (int)std::min<unsigned>(hw, 16);
  if (parts == 1) { std::sort(begin, end, cmp); return; }
  std::vector<int64_t> cut(parts + 1);
  for (int p = 0; p <= parts; ++p) cut[p] = n * p / parts;
  std::vector<std::thread> th;
  for (int p = 0; p < parts; ++p) th.emplace_back([&, p]() { std::sort(begin + cut[p], begin + cut[p + 1], cmp); });
  for (auto& t : th) t.join();
  for (int width = 1; width < parts; width *= 2) {
    std::vector<std::thread> mt;
    for (int p = 0; p + width < parts; p += 2 * width) {
      int64_t a = cut[p], b = cut[p + width], c = cut[std::min(p + 2 * width, parts)];
      mt.emplace_back([=]() { std::inplace_merge(begin + a, begin + b, begin + c, cmp); });
    }
    for (auto& t : mt) t.join();
  }
}

}  // namespace

FrustumConst frustum_const(const AsscppParams& p) {
  // frustum_culling.cpp:88-95: degrees -> radians through double (M_PI), tan in double, narrowed
  FrustumConst f;
  float vr = (float)((double)p.vfov_deg * M_PI / 180);
  float hr = (float)((double)p.hfov_deg * M_PI / 180);
  float vr2 = vr / 2, hr2 = hr / 2;
  f.np_h = (float)(2 * std::tan((double)vr2) * (double)p.near_dist);
  f.np_w = (float)(2 * std::tan((double)hr2) * (double)p.near_dist);
  f.fp_h = (float)(2 * std::tan((double)vr2) * (double)p.far_dist);
  f.fp_w = (float)(2 * std::tan((double)hr2) * (double)p.far_dist);
  f.near_dist = p.near_dist;
  f.far_dist = p.far_dist;
  return f;
}

double frustum_max_extent(const AsscppParams& p) {
  FrustumConst f = frustum_const(p);
  // farthest pair of frustum points: far-plane diagonal, or near corner to opposite far corner
  double fd = std::sqrt((double)f.fp_h * f.fp_h + (double)f.fp_w * f.fp_w);
  double dh = 0.5 * ((double)f.fp_h + f.np_h), dw = 0.5 * ((double)f.fp_w + f.np_w);
  double depth = (double)p.far_dist - (double)p.near_dist;
  double cross = std::sqrt(dh * dh + dw * dw + depth * depth);
  return std::max(fd, cross);
}

// <tf Matrix3x3::getRPY> (getEulerYPR, solution 1) on the <tf Matrix3x3::setRotation> matrix of the pose's
// quaternion, as called at occlusion_culling.cpp:497-500 and :558-561.  Host libm, double.
void pose_rpy_host(const double pose[7], double rpy[3]) {
  const double x = pose[3], y = pose[4], z = pose[5], w = pose[6];
  const double d = x * x + y * y + z * z + w * w, s = 2.0 / d;
  const double xs = x * s, ys = y * s, zs = z * s, wx = w * xs, wy = w * ys, wz = w * zs;
  const double xx = x * xs, xy = x * ys, xz = x * zs, yy = y * ys, yz = y * zs, zz = z * zs;
  const double r00 = 1.0 - (yy + zz), r01 = xy - wz, r02 = xz + wy, r10 = xy + wz, r20 = xz - wy, r21 = yz + wx,
               r22 = 1.0 - (xx + yy);
  const double pi = 3.1415926535897932384626433832795029;
  double yaw, pitch, roll;
  if (std::fabs(r20) >= 1) {
    yaw = 0;
    if (r20 < 0) { pitch = pi / 2.0; roll = std::atan2(r01, r02); }
    else { pitch = -pi / 2.0; roll = std::atan2(-r01, -r02); }
  } else {
    pitch = -std::asin(r20);
    roll = std::atan2(r21 / std::cos(pitch), r22 / std::cos(pitch));
    yaw = std::atan2(r10 / std::cos(pitch), r00 / std::cos(pitch));
  }
  rpy[0] = roll; rpy[1] = pitch; rpy[2] = yaw;
}

void viewport_trig_host(const double pose[7], ViewTrig* out) {
  double rpy[3];
  pose_rpy_host(pose, rpy);
  const double yaw = -1 * rpy[2], pitch = -1 * rpy[1];  // occlusion_culling.cpp:562, 579
  out->t[0] = -1 * pose[0]; out->t[1] = -1 * pose[1]; out->t[2] = -1 * pose[2];  // :545
  out->cz = std::cos(yaw); out->sz = std::sin(yaw);
  out->cy = std::cos(pitch); out->sy = std::sin(pitch);
  out->pad = 0;
}

void frustum_planes_host(const AsscppParams& p, const double pose[7], float planes[24], float origin[3],
                         float corners[24]) {
  frustum_planes_host_fc(frustum_const(p), pose, planes, origin, corners);
}

void frustum_planes_host_fc(const FrustumConst& fc, const double pose[7], float planes[24], float origin[3],
                            float corners[24]) {
  // occlusion_culling.cpp:271-290 (tf::Matrix3x3(q) in double, cast to float) then
  // frustum_culling.cpp:82-141.  Mirrors planes_kernel in kernels.cu operation for operation.
  double x = pose[3], y = pose[4], z = pose[5], w = pose[6];
  double d = x * x + y * y + z * z + w * w;
  double s = 2.0 / d;
  double xs = x * s, ys = y * s, zs = z * s;
  double wx = w * xs, wy = w * ys, wz = w * zs;
  double xx = x * xs, xy = x * ys, xz = x * zs;
  double yy = y * ys, yz = y * zs, zz = z * zs;
  float view[3] = {(float)(1.0 - (yy + zz)), (float)(xy + wz), (float)(xz - wy)};
  float up[3] = {(float)(xy - wz), (float)(1.0 - (xx + zz)), (float)(yz + wx)};
  float right[3] = {(float)(xz + wy), (float)(yz - wx), (float)(1.0 - (xx + yy))};
  float T[3] = {(float)pose[0], (float)pose[1], (float)pose[2]};
  float fpc[3], npc[3], c8[8][3];
  for (int c = 0; c < 3; ++c) {
    float vf = view[c] * fc.far_dist;
    fpc[c] = T[c] + vf;
    float uh = (up[c] * fc.fp_h) / 2, rw = (right[c] * fc.fp_w) / 2;
    float fu = fpc[c] + uh, fd = fpc[c] - uh;
    c8[0][c] = fu - rw; c8[1][c] = fu + rw; c8[2][c] = fd - rw; c8[3][c] = fd + rw;
    float vn = view[c] * fc.near_dist;
    npc[c] = T[c] + vn;
    float uhn = (up[c] * fc.np_h) / 2, rwn = (right[c] * fc.np_w) / 2;
    float nu = npc[c] + uhn, nd = npc[c] - uhn;
    c8[4][c] = nu - rwn; c8[5][c] = nu + rwn; c8[6][c] = nd - rwn; c8[7][c] = nd + rwn;
  }
  if (corners)
    for (int k = 0; k < 8; ++k)
      for (int c = 0; c < 3; ++c) corners[3 * k + c] = c8[k][c];
  auto cross = [](const float* a, const float* b, float* o) {
    float t0 = a[1] * b[2], t1 = a[2] * b[1], t2 = a[2] * b[0], t3 = a[0] * b[2], t4 = a[0] * b[1], t5 = a[1] * b[0];
    o[0] = t0 - t1; o[1] = t2 - t3; o[2] = t4 - t5;
  };
  auto ndot = [](const float* a, const float* b) {
    float p0 = a[0] * b[0], p1 = a[1] * b[1], p2 = a[2] * b[2];
    float s12 = p1 + p2;
    return -(p0 + s12);
  };
  const float *fp_tl = c8[0], *fp_tr = c8[1], *fp_bl = c8[2], *fp_br = c8[3], *np_tr = c8[5], *np_bl = c8[6],
              *np_br = c8[7];
  float e1[3], e2[3], a[3], b[3], cc[3], dd[3];
  float *pl_l = planes, *pl_r = planes + 4, *pl_t = planes + 8, *pl_b = planes + 12, *pl_f = planes + 16,
        *pl_n = planes + 20;
  for (int c = 0; c < 3; ++c) { e1[c] = fp_bl[c] - fp_br[c]; e2[c] = fp_tr[c] - fp_br[c]; }
  cross(e1, e2, pl_f); pl_f[3] = ndot(fpc, pl_f);
  for (int c = 0; c < 3; ++c) { e1[c] = np_tr[c] - np_br[c]; e2[c] = np_bl[c] - np_br[c]; }
  cross(e1, e2, pl_n); pl_n[3] = ndot(npc, pl_n);
  for (int c = 0; c < 3; ++c) {
    a[c] = fp_bl[c] - T[c]; b[c] = fp_br[c] - T[c]; cc[c] = fp_tr[c] - T[c]; dd[c] = fp_tl[c] - T[c];
  }
  cross(b, cc, pl_r); cross(dd, a, pl_l); cross(cc, dd, pl_t); cross(a, b, pl_b);
  pl_r[3] = ndot(T, pl_r); pl_l[3] = ndot(T, pl_l); pl_t[3] = ndot(T, pl_t); pl_b[3] = ndot(T, pl_b);
  for (int c = 0; c < 3; ++c) origin[c] = T[c];
}

int preprocess_model(const float* xyz, int64_t n, const AsscppParams& p, ModelHost* out, std::string* err) {
  ModelHost& m = *out;
  m.n = n;
  m.inv_leaf = 1.0f / p.leaf;  // <PCL setLeafSize>
  m.inv_cov = 1.0f / p.cov_leaf;
  uint32_t lb, cb;
  std::memcpy(&lb, &p.leaf, 4);
  std::memcpy(&cb, &p.cov_leaf, 4);
  m.cov_same = (lb == cb);
  {
    // occluder rule (voxel_grid_occlusion_estimation.cpp:560-561): sqrt((double)sum) > (double)(leaf*2).
    // sqrt is monotone, so the rule is "sum > s_max" with s_max the largest float that fails it.
    float two_leaf = p.leaf * 2;
    double thr = (double)two_leaf;
    float s = (float)(thr * thr);
    while (std::sqrt((double)s) > thr) s = std::nextafterf(s, -INFINITY);
    while (std::sqrt((double)std::nextafterf(s, INFINITY)) <= thr) s = std::nextafterf(s, INFINITY);
    m.dist_thr_sq = s;
  }
  if (n == 0) {
    m.vox.push_back(Int4{0, 0, 0, 0});
    return ASSCPP_OK;
  }
  // ---- lattice coordinates ------------------------------------------------------------------
  const float lim = 1.0e9f;
  std::vector<int32_t> ijk(3 * n);
  int32_t lo[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, hi[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
  int32_t clo[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, chi[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
  for (int64_t i = 0; i < n; ++i)
    for (int c = 0; c < 3; ++c) {
      float v = xyz[3 * i + c];
      float q = v * m.inv_leaf, qc = v * m.inv_cov;
      if (!std::isfinite(v) || !(std::fabs(q) < lim) || !(std::fabs(qc) < lim)) {
        *err = "point coordinates must be finite and |x/leaf| < 1e9";
        return ASSCPP_ERANGE;
      }
      int32_t a = (int32_t)std::floor(q), b = (int32_t)std::floor(qc);
      ijk[3 * i + c] = a;
      lo[c] = std::min(lo[c], a); hi[c] = std::max(hi[c], a);
      clo[c] = std::min(clo[c], b); chi[c] = std::max(chi[c], b);
    }
  int64_t div[3], cdiv[3];
  for (int c = 0; c < 3; ++c) {
    m.min_b[c] = lo[c]; m.max_b[c] = hi[c];
    div[c] = (int64_t)hi[c] - lo[c] + 1;
    cdiv[c] = (int64_t)chi[c] - clo[c] + 1;
    m.cov_min_b[c] = clo[c];
    if (div[c] >= (1 << 21)) { *err = "model spans more than 2^21 voxels on an axis"; return ASSCPP_ERANGE; }
  }
  // <PCL VoxelGrid::applyFilter>: "Leaf size is too small for the input dataset" (int32 index overflow)
  if (div[0] * div[1] * div[2] > 0x7fffffffLL || cdiv[0] * cdiv[1] * cdiv[2] > 0x7fffffffLL) {
    *err = "leaf size is too small for the model (voxel grid exceeds 2^31 cells)";
    return ASSCPP_ERANGE;
  }
  for (int c = 0; c < 3; ++c) m.cov_div[c] = (int32_t)cdiv[c];
  // ---- sort by (morton(voxel), original index) ------------------------------------------------
  std::vector<KeyIdx> keys(n);
  for (int64_t i = 0; i < n; ++i) {
    keys[i].key = morton3((uint32_t)(ijk[3 * i] - lo[0]), (uint32_t)(ijk[3 * i + 1] - lo[1]),
                          (uint32_t)(ijk[3 * i + 2] - lo[2]));
    keys[i].idx = (uint32_t)i;
  }
  parallel_sort(keys.begin(), keys.end(),
                [](const KeyIdx& a, const KeyIdx& b) { return a.key < b.key || (a.key == b.key && a.idx < b.idx); });
  m.pts.resize(n);
  m.orig.resize(n);
  m.pos_of.resize(n);
  m.vox.clear();
  for (int64_t pos = 0; pos < n; ++pos) {
    uint32_t i = keys[pos].idx;
    if (pos == 0 || keys[pos].key != keys[pos - 1].key)
      m.vox.push_back(Int4{ijk[3 * (int64_t)i], ijk[3 * (int64_t)i + 1], ijk[3 * (int64_t)i + 2], (int32_t)pos});
    uint32_t u = (uint32_t)(m.vox.size() - 1);
    float wbits;
    std::memcpy(&wbits, &u, 4);
    m.pts[pos] = Float4{xyz[3 * (int64_t)i], xyz[3 * (int64_t)i + 1], xyz[3 * (int64_t)i + 2], wbits};
    m.orig[pos] = (int32_t)i;
    m.pos_of[i] = (int32_t)pos;
  }
  m.n_vox = (int64_t)m.vox.size();
  m.vox.push_back(Int4{0, 0, 0, (int32_t)n});
  // ---- coverage lattice: compact ids in ascending (k, j, i) order -----------------------------
  std::vector<int64_t> ckey(n);
  for (int64_t i = 0; i < n; ++i) {
    int64_t a = (int32_t)std::floor(xyz[3 * i] * m.inv_cov) - clo[0];
    int64_t b = (int32_t)std::floor(xyz[3 * i + 1] * m.inv_cov) - clo[1];
    int64_t c = (int32_t)std::floor(xyz[3 * i + 2] * m.inv_cov) - clo[2];
    ckey[i] = a + cdiv[0] * (b + cdiv[1] * c);
  }
  m.cov_keys = ckey;
  parallel_sort(m.cov_keys.begin(), m.cov_keys.end(), [](int64_t a, int64_t b) { return a < b; });
  m.cov_keys.erase(std::unique(m.cov_keys.begin(), m.cov_keys.end()), m.cov_keys.end());
  m.n_cov = (int64_t)m.cov_keys.size();
  m.pt_cov.resize(n);
  {
    unsigned hw = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    std::vector<std::thread> th;
    for (unsigned t = 0; t < hw; ++t)
      th.emplace_back([&, t]() {
        for (int64_t pos = n * t / hw; pos < n * (t + 1) / hw; ++pos) {
          int64_t k = ckey[m.orig[pos]];
          m.pt_cov[pos] = (int32_t)(std::lower_bound(m.cov_keys.begin(), m.cov_keys.end(), k) - m.cov_keys.begin());
        }
      });
    for (auto& t : th) t.join();
  }
  m.vox_cov.resize(m.n_vox);
  // coverage id shared by all points of the voxel, or -1 when they straddle coverage cells (possible whenever the
  // two lattices do not nest exactly in float32)
  for (int64_t u = 0; u < m.n_vox; ++u) {
    int32_t id = m.pt_cov[m.vox[u].w];
    for (int64_t q = m.vox[u].w + 1; q < m.vox[u + 1].w; ++q)
      if (m.pt_cov[q] != id) { id = -1; break; }
    m.vox_cov[u] = id;
  }
  // ---- chunks of whole voxels, tight AABBs ------------------------------------------------------
  auto bits_of = [](uint32_t v) { float f; std::memcpy(&f, &v, 4); return f; };
  int64_t u = 0;
  while (u < m.n_vox) {
    int64_t ub = u, pb = m.vox[u].w;
    ++u;
    while (u < m.n_vox && m.vox[u + 1].w - pb <= kChunkMaxPoints) ++u;
    int64_t pe = m.vox[u].w;
    Float4 l{INFINITY, INFINITY, INFINITY, bits_of((uint32_t)pb)}, h{-INFINITY, -INFINITY, -INFINITY, bits_of((uint32_t)pe)};
    for (int64_t q = pb; q < pe; ++q) {
      l.x = std::min(l.x, m.pts[q].x); l.y = std::min(l.y, m.pts[q].y); l.z = std::min(l.z, m.pts[q].z);
      h.x = std::max(h.x, m.pts[q].x); h.y = std::max(h.y, m.pts[q].y); h.z = std::max(h.z, m.pts[q].z);
    }
    m.chunk_lo.push_back(l);
    m.chunk_hi.push_back(h);
    m.chunk_vox.push_back((int32_t)ub);
  }
  m.chunk_vox.push_back((int32_t)m.n_vox);
  m.n_chunks = (int64_t)m.chunk_lo.size();
  m.n_super = (m.n_chunks + kSuperFanout - 1) / kSuperFanout;
  m.super_lo.resize(m.n_super);
  m.super_hi.resize(m.n_super);
  for (int64_t s = 0; s < m.n_super; ++s) {
    Float4 l{INFINITY, INFINITY, INFINITY, 0}, h{-INFINITY, -INFINITY, -INFINITY, 0};
    for (int64_t c = s * kSuperFanout; c < std::min<int64_t>(m.n_chunks, (s + 1) * kSuperFanout); ++c) {
      l.x = std::min(l.x, m.chunk_lo[c].x); l.y = std::min(l.y, m.chunk_lo[c].y); l.z = std::min(l.z, m.chunk_lo[c].z);
      h.x = std::max(h.x, m.chunk_hi[c].x); h.y = std::max(h.y, m.chunk_hi[c].y); h.z = std::max(h.z, m.chunk_hi[c].z);
    }
    m.super_lo[s] = l;
    m.super_hi[s] = h;
  }
  return ASSCPP_OK;
}

}  // namespace asscpp
