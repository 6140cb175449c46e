// kernels.cu -- sm_100a kernels of the coverage-evaluation path.
//
//   planes_kernel      pose -> camera matrix -> six frustum planes          (occlusion_culling.cpp:271-290,
//                                                                            frustum_culling.cpp:82-141)
//   eval_views_kernel  persistent, one CTA per viewpoint at a time:
//       P1 bucket pre-cull (conservative AABB-vs-plane reject of super-chunks / chunks)
//       P2 exact per-point frustum test + emission of the voxels occupied for this view
//                                                                           (frustum_culling.cpp:150-171)
//       P3 per-view occupancy bit-grid with PCL's dense linearisation      (voxel_grid_occlusion_estimation.cpp:43-65)
//       P4 one 3D-DDA ray per occupied voxel + second gate + gain popcounts (...:99-128,174-238,410-582;
//                                                                            occlusion_culling.cpp:317-370)
//       P5 optional OR of the visible set into a covered set               (heuristic.cpp:344-346)
//   ordered_views_kernel  exact replay of the reference's sticky `state` (quirk Q5) for the rare views in
//                         which occlusionEstimation failed (occlusion_culling.cpp:308,332-335)
//   set kernels        covered-set OR / popcount / permutation / coverage of an arbitrary cloud
//
// All float32 arithmetic that decides set membership uses the round-to-nearest intrinsics (never
// contracted into FMA) in the operation order fixed in SURVEY.md Appendix A.
#include <cuda_runtime.h>

#include <cmath>
#include <cstring>

#include "common.h"
#include "kernels.h"

namespace asscpp {

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }

// Model points are read once per view and never reused by the CTA; the per-CTA scratch (voxel lists, ray queues) is
// rewritten every view and should stay in L2.  The point reads are marked evict-first (ld.global.cs) so that
// they do not push dirty scratch lines out to DRAM.
// The per-CTA voxel lists (occupied voxels, free voxels) are written once and read back one or two phases later; in
// between the SM streams several hundred KB of model data through L2.  With -DASSCPP_KEEP_LISTS their stores carry an
// L2 evict-last policy.  Measured (profiles/r02aa_keep_lists_ab.txt, 32 MB more persisting carve-out): DRAM writes
// 1.25 -> 0.65 GB per 10 000-view launch, DRAM reads +0.67 GB (less L2 left for the model), 0.7 % slower -- not
// adopted; plain stores are the default.
__device__ __forceinline__ void st_list(uint32_t* a, uint32_t v) {
#ifndef ASSCPP_KEEP_LISTS
  *a = v;
#else
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(a), "r"(v), "l"(pol) : "memory");
#endif
}

__device__ __forceinline__ float4 ld_pt(const float4* p) {
#ifdef ASSCPP_NO_STREAM_POINTS
  return *p;
#else
  return __ldcs(p);  // profiles/r02d_stream_points_ab.txt: same speed, 16 % fewer DRAM bytes written, 9.5 % fewer read
#endif
}

// not inlined: the hash is optional (ASSCPP_EVAL_HASH) and its 64-bit multiplies would sit in every point loop, in
// the way of the instruction fetch of the paths that do not ask for it
__device__ __noinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

// frustum_culling.cpp:153-162: pt = (x, y, z, 1); inside iff pt.dot(plane) <= 0 for l r t b f n.
// <Eigen> 4-wide dot reduced as (a0 + a2) + (a1 + a3).
__device__ __forceinline__ bool point_in_frustum(const float* pl, float x, float y, float z) {
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float a = fmul(x, pl[4 * k]), b = fmul(y, pl[4 * k + 1]), c = fmul(z, pl[4 * k + 2]);
    float t = fadd(fadd(a, c), fadd(b, pl[4 * k + 3]));
    if (!(t <= 0.0f)) return false;
  }
  return true;
}

// Conservative bucket reject: true only if every point of the box fails the exact test of some
// plane.  The exact value differs from the real-valued plane function by at most a few ulps of the
// sum of magnitudes A; the margin 1e-6*A (16 ulps) + 1e-30 (underflow) covers it.
__device__ __forceinline__ bool box_outside(const float* pl, float lx, float ly, float lz, float hx, float hy,
                                            float hz) {
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float a = pl[4 * k], b = pl[4 * k + 1], c = pl[4 * k + 2], w = pl[4 * k + 3];
    float L = (a >= 0 ? a * lx : a * hx) + (b >= 0 ? b * ly : b * hy) + (c >= 0 ? c * lz : c * hz) + w;
    float A = fabsf(a) * fmaxf(fabsf(lx), fabsf(hx)) + fabsf(b) * fmaxf(fabsf(ly), fabsf(hy)) +
              fabsf(c) * fmaxf(fabsf(lz), fabsf(hz)) + fabsf(w);
    if (L > 1e-6f * A + 1e-30f) return true;
  }
  return false;
}

// Conservative bucket accept: true only if every point of the box passes the exact test of every plane (same
// margin argument as box_outside, on the maximum of the plane function over the box).
__device__ __forceinline__ bool box_inside(const float* pl, float lx, float ly, float lz, float hx, float hy,
                                           float hz) {
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float a = pl[4 * k], b = pl[4 * k + 1], c = pl[4 * k + 2], w = pl[4 * k + 3];
    float U = (a >= 0 ? a * hx : a * lx) + (b >= 0 ? b * hy : b * ly) + (c >= 0 ? c * hz : c * lz) + w;
    float A = fabsf(a) * fmaxf(fabsf(lx), fabsf(hx)) + fabsf(b) * fmaxf(fabsf(ly), fabsf(hy)) +
              fabsf(c) * fmaxf(fabsf(lz), fabsf(hz)) + fabsf(w);
    if (!(U < -(1e-6f * A + 1e-30f))) return false;
  }
  return true;
}

// ------------------------------------------------------------------------------------------------
// planes_kernel: one thread per pose.  Double part = <tf Matrix3x3::setRotation>; float part =
// frustum_culling.cpp:97-141 with the operation order of SURVEY.md A.2.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void compute_view(const double* __restrict__ ps, const FrustumConst& fc, ViewDev& out) {
  double x = ps[3], y = ps[4], z = ps[5], w = ps[6];
  double d = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)), __dmul_rn(w, w));
  double s = __ddiv_rn(2.0, d);
  double xs = __dmul_rn(x, s), ys = __dmul_rn(y, s), zs = __dmul_rn(z, s);
  double wx = __dmul_rn(w, xs), wy = __dmul_rn(w, ys), wz = __dmul_rn(w, zs);
  double xx = __dmul_rn(x, xs), xy = __dmul_rn(x, ys), xz = __dmul_rn(x, zs);
  double yy = __dmul_rn(y, ys), yz = __dmul_rn(y, zs), zz = __dmul_rn(z, zs);
  float view[3] = {(float)__dsub_rn(1.0, __dadd_rn(yy, zz)), (float)__dadd_rn(xy, wz), (float)__dsub_rn(xz, wy)};
  float up[3] = {(float)__dsub_rn(xy, wz), (float)__dsub_rn(1.0, __dadd_rn(xx, zz)), (float)__dadd_rn(yz, wx)};
  float right[3] = {(float)__dadd_rn(xz, wy), (float)__dsub_rn(yz, wx), (float)__dsub_rn(1.0, __dadd_rn(xx, yy))};
  float T[3] = {(float)ps[0], (float)ps[1], (float)ps[2]};
  float fpc[3], npc[3], c8[8][3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    fpc[c] = fadd(T[c], fmul(view[c], fc.far_dist));
    float uh = fdiv(fmul(up[c], fc.fp_h), 2.0f), rw = fdiv(fmul(right[c], fc.fp_w), 2.0f);
    float fu = fadd(fpc[c], uh), fd = fsub(fpc[c], uh);
    c8[0][c] = fsub(fu, rw); c8[1][c] = fadd(fu, rw); c8[2][c] = fsub(fd, rw); c8[3][c] = fadd(fd, rw);
    npc[c] = fadd(T[c], fmul(view[c], fc.near_dist));
    float uhn = fdiv(fmul(up[c], fc.np_h), 2.0f), rwn = fdiv(fmul(right[c], fc.np_w), 2.0f);
    float nu = fadd(npc[c], uhn), nd = fsub(npc[c], uhn);
    c8[4][c] = fsub(nu, rwn); c8[5][c] = fadd(nu, rwn); c8[6][c] = fsub(nd, rwn); c8[7][c] = fadd(nd, rwn);
  }
  auto cross = [](const float* a, const float* b, float* o) {
    o[0] = fsub(fmul(a[1], b[2]), fmul(a[2], b[1]));
    o[1] = fsub(fmul(a[2], b[0]), fmul(a[0], b[2]));
    o[2] = fsub(fmul(a[0], b[1]), fmul(a[1], b[0]));
  };
  auto ndot = [](const float* a, const float* b) {
    return -fadd(fmul(a[0], b[0]), fadd(fmul(a[1], b[1]), fmul(a[2], b[2])));
  };
  float *pl_l = out.planes, *pl_r = out.planes + 4, *pl_t = out.planes + 8, *pl_b = out.planes + 12,
        *pl_f = out.planes + 16, *pl_n = out.planes + 20;
  float e1[3], e2[3], a[3], b[3], cc[3], dd[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { e1[c] = fsub(c8[2][c], c8[3][c]); e2[c] = fsub(c8[1][c], c8[3][c]); }
  cross(e1, e2, pl_f); pl_f[3] = ndot(fpc, pl_f);
#pragma unroll
  for (int c = 0; c < 3; ++c) { e1[c] = fsub(c8[5][c], c8[7][c]); e2[c] = fsub(c8[6][c], c8[7][c]); }
  cross(e1, e2, pl_n); pl_n[3] = ndot(npc, pl_n);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    a[c] = fsub(c8[2][c], T[c]); b[c] = fsub(c8[3][c], T[c]); cc[c] = fsub(c8[1][c], T[c]); dd[c] = fsub(c8[0][c], T[c]);
  }
  cross(b, cc, pl_r); cross(dd, a, pl_l); cross(cc, dd, pl_t); cross(a, b, pl_b);
  pl_r[3] = ndot(T, pl_r); pl_l[3] = ndot(T, pl_l); pl_t[3] = ndot(T, pl_t); pl_b[3] = ndot(T, pl_b);
  out.origin[0] = T[0]; out.origin[1] = T[1]; out.origin[2] = T[2];
  out.pad = 0.0f;
}


__global__ void planes_kernel(const double* __restrict__ poses, int64_t nviews, FrustumConst fc,
                              ViewDev* __restrict__ views) {
  int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nviews) return;
  ViewDev out;
  compute_view(poses + 7 * v, fc, out);
  views[v] = out;
}

// ------------------------------------------------------------------------------------------------
// per-view state shared by the CTA
// ------------------------------------------------------------------------------------------------
struct GridView {  // geometry of the per-view grid (voxel_grid_occlusion_estimation.cpp:43-65)
  float leaf, inv, half, thr_sq;
  float base[3];  // b_min + leaf/2  (voxel_grid_occlusion_estimation.h:151-153)
  float b_min[3], b_max[3];
  int min_b[3], div[3];
  int mul1, mul2;     // PCL strides: div0, div0*div1 (linear index of quirk Q3)
  unsigned cells;     // div0*div1*div2
  int P0, P1, P2, P01;  // padded walk grid: P = div + 4 (one halo layer + one sentinel layer per side)
  unsigned mP0, mP01, mD0, mD01;  // floor(2^32 / d) for d = P0, P01, div0, div0*div1 (division by multiply-high)
};

// n / d for n < 2^31 with m = floor(2^32 / d): the multiply-high estimate is at most one too small.
__device__ __forceinline__ int fast_div(int n, int d, unsigned m) {
  int q = (int)__umulhi((unsigned)n, m);
  return q + ((n - q * d) >= d ? 1 : 0);
}

// Where the per-view bit-grid lives: shared memory (32-bit shared address) or the per-CTA global scratch.
struct GridMem {
  uint32_t* gptr;
  uint32_t saddr;
};
template <bool SMEM>
__device__ __forceinline__ uint32_t grid_load(const GridMem& gm, unsigned w) {
  if constexpr (SMEM) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(gm.saddr + (w << 2)));
    return v;
  } else {
    return gm.gptr[w];
  }
}

constexpr unsigned kDescShared = 1024;
// A voxel with at least this many points is not walked by one lane (a chain of dependent loads, one point after the
// other): it is set aside and the points are tested 32 at a time by a whole warp (coarse lattices hold hundreds of
// points per voxel; with few voxels per view that chain was most of a small batch's latency).
#ifndef ASSCPP_COOP_MIN_POINTS
#define ASSCPP_COOP_MIN_POINTS 8
#endif
constexpr uint32_t kCoopMinPoints = ASSCPP_COOP_MIN_POINTS;
// ... and with few voxels in all (a small view: fewer voxels than lane groups in the CTA) every voxel of two or more
// points is taken that way: idle warps cost nothing, a lane walking seven points one after the other costs microseconds
constexpr unsigned kCoopAllBelow = 128;
__device__ __forceinline__ uint32_t coop_min_points(unsigned nvoxels) { return nvoxels <= kCoopAllBelow ? 2u : kCoopMinPoints; }

struct ViewShared {
  float planes[24];
  float origin[3];
  float vox_hi[6];  // cull: voxel cube classification bound per plane
  GridView g;
  unsigned words;
  int bb_min[3], bb_max[3];
  int view;
  int use_smem;
  unsigned long long cand_state;  // cull: descriptors handed out (high word) | candidate voxels so far (low word)
  unsigned big_count, chunk_next, super_count, list_count, vis_count, ray_next;  // big_count: voxels handled by lane groups
  unsigned big_next;  // cull P1b: super-chunks handed out so far (beyond the first of every warp)
  unsigned survivors, visible, carry, unique, gain_p, gain_v, status;
  unsigned long long steps, vis_hash, surv_hash;
  unsigned long long acc[4];  // accuracy: fixed-point sums (depth hi / lo, z hi / lo)
  int acc_min, acc_max;       // accuracy: ordered-int keys of the depth extrema
  ViewTrig trig;              // accuracy: the view's transform constants, fetched while the cull runs
  int seg_last[1024];  // ordered path; cull: list of surviving super-chunks
  uint2 chunk_desc[kDescShared];  // cull: candidate chunk descriptors (the rest spill to the global chunk list)
};

// include/cscpp/voxel_grid_occlusion_estimation.h:143-157 with r = ijk - min_b
__device__ __forceinline__ float centre(const GridView& g, int c, int r) {
  return fadd(g.base[c], fmul((float)r, g.leaf));
}

// voxel_grid_occlusion_estimation.cpp:174-238
__device__ __forceinline__ float ray_box(const GridView& g, const float o[3], const float d[3]) {
  float tmin, tmax, tymin, tymax, tzmin, tzmax;
  if (d[0] >= 0) { tmin = fdiv(fsub(g.b_min[0], o[0]), d[0]); tmax = fdiv(fsub(g.b_max[0], o[0]), d[0]); }
  else           { tmin = fdiv(fsub(g.b_max[0], o[0]), d[0]); tmax = fdiv(fsub(g.b_min[0], o[0]), d[0]); }
  if (d[1] >= 0) { tymin = fdiv(fsub(g.b_min[1], o[1]), d[1]); tymax = fdiv(fsub(g.b_max[1], o[1]), d[1]); }
  else           { tymin = fdiv(fsub(g.b_max[1], o[1]), d[1]); tymax = fdiv(fsub(g.b_min[1], o[1]), d[1]); }
  if ((tmin > tymax) || (tymin > tmax)) return -1.0f;
  if (tymin > tmin) tmin = tymin;
  if (tymax < tmax) tmax = tymax;
  if (d[2] >= 0) { tzmin = fdiv(fsub(g.b_min[2], o[2]), d[2]); tzmax = fdiv(fsub(g.b_max[2], o[2]), d[2]); }
  else           { tzmin = fdiv(fsub(g.b_max[2], o[2]), d[2]); tzmax = fdiv(fsub(g.b_min[2], o[2]), d[2]); }
  if ((tmin > tzmax) || (tzmin > tmax)) return -1.0f;
  if (tzmin > tmin) tmin = tzmin;
  return tmin;
}

// (centre - origin).normalize() on a Vector4f whose w is 0 - 0; <Eigen> squaredNorm = (v0^2+v2^2)+(v1^2+0)
__device__ __forceinline__ void direction(const float ctr[3], const float o[3], float d[3]) {
  float v0 = fsub(ctr[0], o[0]), v1 = fsub(ctr[1], o[1]), v2 = fsub(ctr[2], o[2]);
  float n2 = fadd(fadd(fmul(v0, v0), fmul(v2, v2)), fadd(fmul(v1, v1), 0.0f));
  float nrm = __fsqrt_rn(n2);
  d[0] = fdiv(v0, nrm); d[1] = fdiv(v1, nrm); d[2] = fdiv(v2, nrm);
}

// occlusionEstimation + rayTraversal (voxel_grid_occlusion_estimation.cpp:99-128, 410-582) for the
// voxel at relative coordinates rt (= ijk - min_b) with cell centre tc.  Returns 0 free, 1 occluded,
// 2 failed (tmin == -1: the caller's `state` stays untouched, quirk Q5).  The walk stops at the first
// qualifying occluder: the reference keeps walking but its result can only stay 1 (:561-562).
//
// The walk runs on a PADDED copy of the grid (coordinate q = r + 2, P = div + 4 cells per axis) so the
// hot loop needs one integer per ray: the padded linear index p.
//   * interior cells (0 <= r < div) hold the occupancy bit,
//   * the halo layer (r == -1 or r == div) is where the reference still walks (:538-540) and where
//     PCL's unchecked linear index aliases into another row/slice (quirk Q3); its bits are set so the
//     rare visit drops into the slow path, which evaluates the aliased lookup exactly,
//   * the sentinel layer (r == -2 or r == div + 1) is set as well: arriving there is the loop exit.
// Target test = one compare (p is unique per cell), bounds test = none in the hot loop.
template <bool SMEM>
__device__ __forceinline__ int occlusion_estimation(const GridView& g, const GridMem& gm, const float origin[3],
                                                    const int rt[3], const float tc[3], unsigned& steps) {
  float d[3];
  direction(tc, origin, d);                 // :111-113
  float tmin = ray_box(g, origin, d);       // :116
  if (tmin == -1.0f) return 2;              // :118-122
  float t_max[3], t_delta[3];
  int r[3], step[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float start = fadd(origin[c], fmul(tmin, d[c]));                   // :422
    int ijk = (int)floorf(fadd(fmul(start, g.inv), 0.5f));             // :425 (round = floor(x + 0.5f))
    r[c] = ijk - g.min_b[c];
    float vm = centre(g, c, r[c]);                                     // :432
    if (d[c] >= 0) { vm = fadd(vm, g.half); step[c] = 1; }             // :436-465
    else           { vm = fsub(vm, g.half); step[c] = -1; }
    t_max[c] = fadd(tmin, fdiv(fsub(vm, start), d[c]));                // :513-515
    t_delta[c] = fdiv(g.leaf, fabsf(d[c]));                            // :517-519
  }
  // loop entry condition :538-540: ijk <= max_b + 1 && ijk + 1 >= min_b  <=>  -1 <= r <= div
  if (!((unsigned)(r[0] + 1) <= (unsigned)(g.div[0] + 1) && (unsigned)(r[1] + 1) <= (unsigned)(g.div[1] + 1) &&
        (unsigned)(r[2] + 1) <= (unsigned)(g.div[2] + 1)))
    return 0;
  const int P0 = g.P0, P01 = g.P01;
  const int k0 = step[0], k1 = step[1] * P0, k2 = step[2] * P01;
  const int pt = (rt[0] + 2) + (rt[1] + 2) * P0 + (rt[2] + 2) * P01;
  int p = (r[0] + 2) + (r[1] + 2) * P0 + (r[2] + 2) * P01;
  float tx = t_max[0], ty = t_max[1], tz = t_max[2];
  const float dtx = t_delta[0], dty = t_delta[1], dtz = t_delta[2];
  int result = 0;
  unsigned n = 0;
  for (;;) {
    ++n;
    if (p == pt) break;                                                 // :551-552
    const uint32_t word = grid_load<SMEM>(gm, (unsigned)p >> 5);
    if ((word >> (p & 31)) & 1u) {
      // rare: occupied interior cell, halo cell or sentinel
      const int q2 = fast_div(p, P01, g.mP01), rem = p - q2 * P01, q1 = fast_div(rem, P0, g.mP0), q0 = rem - q1 * P0;
      if (q0 == 0 || q0 == P0 - 1 || q1 == 0 || q1 == g.P1 - 1 || q2 == 0 || q2 == g.P2 - 1) { --n; break; }
      const int c0 = q0 - 2, c1 = q1 - 2, c2 = q2 - 2;
      bool occ = true;
      if (c0 < 0 || c0 == g.div[0] || c1 < 0 || c1 == g.div[1] || c2 < 0 || c2 == g.div[2]) {
        // <PCL getCentroidIndexAt>: idx = (ijk - min_b) . (1, div0, div0*div1), only 0 <= idx < size checked
        const int lin = c0 + c1 * g.mul1 + c2 * g.mul2;
        occ = false;
        if ((unsigned)lin < g.cells) {
          const int a2 = fast_div(lin, g.mul2, g.mD01), ar = lin - a2 * g.mul2, a1 = fast_div(ar, g.mul1, g.mD0),
                    a0 = ar - a1 * g.mul1;
          const int pa = (a0 + 2) + (a1 + 2) * P0 + (a2 + 2) * P01;
          occ = (grid_load<SMEM>(gm, (unsigned)pa >> 5) >> (pa & 31)) & 1u;
        }
      }
      if (occ) {                                                        // :555-556
        float dx = fsub(centre(g, 0, c0), tc[0]), dy = fsub(centre(g, 1, c1), tc[1]),
              dz = fsub(centre(g, 2, c2), tc[2]);
        float sum = fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz)); // :560
        if (sum > g.thr_sq) { result = 1; break; }                        // :561 sqrt(double(sum)) > leaf*2
      }
    }
    // :565-579, ties x -> y -> z by <=; NaN compares false and falls through to z
    const bool sx = (tx <= ty) & (tx <= tz);
    const bool sy = (!sx) & (ty <= tz) & (ty <= tx);
    const bool sz = !(sx | sy);
    tx = sx ? fadd(tx, dtx) : tx;
    ty = sy ? fadd(ty, dty) : ty;
    tz = sz ? fadd(tz, dtz) : tz;
    p += sx ? k0 : (sy ? k1 : k2);
  }
  steps += n;
  return result;
}

// second gate, occlusion_culling.cpp:338-342: ray from the MODEL cloud's sensor origin (Q6)
__device__ __forceinline__ bool second_gate(const GridView& g, const float pcd_origin[3], const float tc[3]) {
  float d[3];
  direction(tc, pcd_origin, d);
  return ray_box(g, pcd_origin, d) != -1.0f;
}

__device__ __forceinline__ bool bit_test(const uint32_t* b, uint32_t i) { return (b[i >> 5] >> (i & 31)) & 1u; }

__device__ __forceinline__ unsigned warp_sum(unsigned v) { return __reduce_add_sync(0xffffffffu, v); }

// ------------------------------------------------------------------------------------------------
// P1 + P2: cull.  Fills sh.survivors / bb_* / surv_hash and the per-CTA list of occupied voxels
// (duplicates possible when a voxel's points straddle two 32-point loads; P3 removes them).
// ------------------------------------------------------------------------------------------------
template <bool BY_CHUNK>
__device__ __forceinline__ void cull_view_t(const EvalArgs& A, ViewShared& sh, uint2* chunk_list, uint32_t* vox_list,
                                            uint32_t* cut_list, unsigned cut_cap) {
  const ModelDev& m = A.m;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const bool brute = (A.flags & ASSCPP_EVAL_BRUTE) != 0;
  // coarse lattices (a leaf bucket is a handful of voxels): the boundary chunks' points are tested chunk by chunk with the
  // brute loop below instead of voxel by voxel (P2) -- the per-voxel cube test can reject nothing a chunk's tight box did not
  const bool by_chunk = BY_CHUNK && !brute;  // (a template parameter: the fine-lattice path compiles exactly as without it)
  const bool want_hash = (A.flags & ASSCPP_EVAL_HASH) != 0;
  if (threadIdx.x < 6) {
    // Voxel cube vs plane k, centre / radius form: every point of a voxel lies within h = leaf/2 + slack of the
    // cube centre on each axis (slack: float32 rounding of x * inv_leaf and of the centre itself, < 1e-6 of the
    // largest model coordinate), so the plane function over the cube stays within r = h (|a|+|b|+|c|) of its value v
    // at the centre; the exact float32 test value differs from the real one by less than M = 1e-6 A (16 ulps of
    // the sum of magnitudes A).  v > r + M: all points fail plane k;  v < -(r + M): all pass it.
    const float* pl = sh.planes + 4 * threadIdx.x;
    const float n1 = fabsf(pl[0]) + fabsf(pl[1]) + fabsf(pl[2]);
    const float h = 0.5f * m.leaf + 1e-6f * m.max_abs_coord;
    sh.vox_hi[threadIdx.x] = h * n1 + (1e-6f * (n1 * (m.max_abs_coord + m.leaf) + fabsf(pl[3])) + 1e-30f);
  }
  int bmin[3] = {INT_MAX, INT_MAX, INT_MAX}, bmax[3] = {INT_MIN, INT_MIN, INT_MIN};
  unsigned surv = 0, surv_in = 0;
  unsigned long long shash = 0;
  if (brute) __syncthreads();
  if (!brute) {
    // P1a: every thread tests super-chunks; survivors go to a shared list (sh.seg_last doubles as that list)
    constexpr unsigned kSuperListCap = 1024;
    const bool prof = A.phase_cycles != nullptr && threadIdx.x == 0;
    const long long c0 = prof ? clock64() : 0;
    // a model with no more super-chunks than warps: P1b tests each super-chunk itself, no list, one barrier less
    const bool direct_supers = m.n_super <= (int64_t)nwarps;
    for (int64_t s0 = (int64_t)warp * 32; s0 < m.n_super && !direct_supers; s0 += (int64_t)nwarps * 32) {
      const int64_t s = s0 + lane;
      bool keep = false;
      if (s < m.n_super) {
        float4 lo = m.super_lo[s], hi = m.super_hi[s];
        keep = !box_outside(sh.planes, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z);
      }
      const unsigned sb = __ballot_sync(0xffffffffu, keep);
      if (sb) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(&sh.super_count, (unsigned)__popc(sb));
        base = __shfl_sync(0xffffffffu, base, 0);
        const unsigned at = base + __popc(sb & ((1u << lane) - 1));
        if (keep && at < kSuperListCap) sh.seg_last[at] = (int)s;
      }
    }
    if (!direct_supers) __syncthreads();
    const long long c1 = prof ? clock64() : 0;
    // P1b: the 32 children of each surviving super-chunk, one super per warp step, spread evenly over the warps
    const unsigned nsup = sh.super_count;
    const bool listed = nsup <= kSuperListCap && !direct_supers;  // otherwise (huge frustum): walk all super-chunks again
    const int64_t nstep = listed ? (int64_t)nsup : m.n_super;
    // the first super-chunk of a warp is its own, the others are claimed (inside chunks make the cost per super uneven)
    bool first_super = true;
    for (int64_t k = warp;;) {
      if (!first_super) {
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(&sh.big_next, 1u);
        k = (int64_t)nwarps + __shfl_sync(0xffffffffu, t, 0);
      }
      first_super = false;
      if (k >= nstep) break;
      int64_t sup = k;
      if (listed) {
        sup = sh.seg_last[k];
      } else {
        float4 lo = m.super_lo[k], hi = m.super_hi[k];
        if (box_outside(sh.planes, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z)) continue;
      }
      const int64_t c = sup * kSuperFanout + lane;
      bool ck = false;
      float4 lo, hi;
      uint32_t vb0 = 0, vb1 = 0;
      if (c < m.n_chunks) {
        lo = m.chunk_lo[c]; hi = m.chunk_hi[c];
        vb0 = (uint32_t)m.chunk_vox[c]; vb1 = (uint32_t)m.chunk_vox[c + 1];  // with the boxes, not after the test: one round trip less
        ck = !box_outside(sh.planes, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z);
      }
      // Chunks wholly inside the frustum: every point passes (box_inside is conservative), so nothing of them needs the
      // per-voxel pass below.  Survivors = the chunk's points; their voxel box = the box of the chunk's tight point
      // AABB (floor(x * inv_leaf) is monotone: the corners' voxels are the extreme voxels); their voxels go straight
      // into the list, flagged "every point survives".
      const bool in = ck && box_inside(sh.planes, lo.x, lo.y, lo.z, hi.x, hi.y, hi.z);
      const unsigned ib = __ballot_sync(0xffffffffu, in);
      if (ib) {
        const uint32_t nvi = in ? vb1 - vb0 : 0u;
        uint32_t incl = nvi;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(&sh.list_count, tot);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (in) {
          const uint32_t pb = __float_as_uint(lo.w), pe = __float_as_uint(hi.w);
          surv_in += pe - pb;
          if (want_hash)
            for (uint32_t q = pb; q < pe; ++q) shash += mix64((uint64_t)(uint32_t)m.orig[q]);
          bmin[0] = min(bmin[0], (int)floorf(fmul(lo.x, m.inv_leaf))); bmax[0] = max(bmax[0], (int)floorf(fmul(hi.x, m.inv_leaf)));
          bmin[1] = min(bmin[1], (int)floorf(fmul(lo.y, m.inv_leaf))); bmax[1] = max(bmax[1], (int)floorf(fmul(hi.y, m.inv_leaf)));
          bmin[2] = min(bmin[2], (int)floorf(fmul(lo.z, m.inv_leaf))); bmax[2] = max(bmax[2], (int)floorf(fmul(hi.z, m.inv_leaf)));
        }
        for (unsigned rest = ib; rest; rest &= rest - 1) {
          const int src = __ffs(rest) - 1;
          const uint32_t v0 = __shfl_sync(0xffffffffu, vb0, src), nvv = __shfl_sync(0xffffffffu, nvi, src);
          const unsigned at0 = base + __shfl_sync(0xffffffffu, incl - nvi, src);
          for (uint32_t k = lane; k < nvv; k += 32)
            if (at0 + k < A.list_cap) st_list(vox_list + at0 + k, (v0 + k) | kVoxInside);
        }
      }
      const bool bd = ck && !in;  // cut by a plane (or not provably inside): its voxels are classified one by one
      const unsigned cb = __ballot_sync(0xffffffffu, bd);
      if (cb) {
        // boundary chunk -> descriptor {first voxel, offset of its voxels in the flat candidate order};
        // one 64-bit atomic hands out descriptor slots (high word) and voxel offsets (low word) together, so
        // offsets ascend with the slot index
        uint32_t vb = 0, nv = 0;
        if (bd) { vb = vb0; nv = vb1 - vb0; }
        uint32_t incl = nv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += t;
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(&sh.cand_state, ((unsigned long long)__popc(cb) << 32) | tot);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (bd) {
          const unsigned slot = (unsigned)(base >> 32) + __popc(cb & ((1u << lane) - 1));
          const uint2 d = by_chunk ? make_uint2((uint32_t)c, 0u) : make_uint2(vb, (uint32_t)base + incl - nv);
          if (slot < kDescShared) sh.chunk_desc[slot] = d;
          else chunk_list[slot] = d;
        }
      }
    }
    __syncthreads();
    if (prof) {
      atomicAdd(&A.phase_cycles[6], (unsigned long long)(c1 - c0));
      atomicAdd(&A.phase_cycles[7], (unsigned long long)(clock64() - c1));
    }
  }
  if (!brute && !by_chunk) {
    // P2: the voxels of the candidate chunks in one flat order, cut into one contiguous segment per warp, one
    // lane per voxel, kCullUnroll blocks of 32 in flight.  Chunk inside the frustum: every point survives.
    // Otherwise the voxel's cube is classified against the planes (centre / radius form with the per-view
    // bounds sh.vox_hi); voxels cut by a plane are set aside and their points tested one by one
    // (frustum_culling.cpp:153-166) afterwards, one lane per such voxel.
    constexpr int kCullUnroll = 2;
    const unsigned total = (unsigned)sh.cand_state, ndesc = (unsigned)(sh.cand_state >> 32);
    const uint32_t coop_min = coop_min_points(total);
    auto desc = [&](unsigned i) { return i < kDescShared ? sh.chunk_desc[i] : chunk_list[i]; };
    auto emit_voxels = [&](bool emit, uint32_t entry) {
      const unsigned ob = __ballot_sync(0xffffffffu, emit);
      if (!ob) return;
      unsigned slot = 0;
      if (lane == 0) slot = atomicAdd(&sh.list_count, (unsigned)__popc(ob));
      slot = __shfl_sync(0xffffffffu, slot, 0);
      if (emit) {
        const unsigned at = slot + __popc(ob & ((1u << lane) - 1));
        if (at < A.list_cap) st_list(vox_list + at, entry);
      }
    };
    uint32_t* cut = cut_list;  // this warp's list of voxels cut by a plane
    unsigned ncut = 0;         // warp-uniform
    auto flush_cut = [&]() {
      __syncwarp();
      for (unsigned i0 = 0; i0 < ncut; i0 += 32) {
        const unsigned i = i0 + lane;
        unsigned cnt = 0;
        uint32_t u = 0;
        if (i < ncut) {
          u = __ldcg(cut + i);
          const int4 vx = m.vox[u];
          const uint32_t q0 = (uint32_t)vx.w, q1 = (uint32_t)m.vox[u + 1].w;
          // the points of a voxel four at a time: the loads, not the plane tests, are what this loop waits for
          for (uint32_t qb = q0; qb < q1; qb += 4) {
            float4 p[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) p[k] = ld_pt(&m.pts[min(qb + k, q1 - 1)]);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (qb + k >= q1 || !point_in_frustum(sh.planes, p[k].x, p[k].y, p[k].z)) continue;
              ++cnt;
              if (want_hash) shash += mix64((uint64_t)(uint32_t)m.orig[qb + k]);
            }
          }
          if (cnt) {
            surv_in += cnt;
            bmin[0] = min(bmin[0], vx.x); bmin[1] = min(bmin[1], vx.y); bmin[2] = min(bmin[2], vx.z);
            bmax[0] = max(bmax[0], vx.x); bmax[1] = max(bmax[1], vx.y); bmax[2] = max(bmax[2], vx.z);
          }
        }
        emit_voxels(cnt != 0, u);
      }
      ncut = 0;
      __syncwarp();
    };
    const bool prof2 = A.phase_cycles != nullptr && threadIdx.x == 0;
    const long long d0 = prof2 ? clock64() : 0;
    // warps claim runs of kCullClaim candidates (cut voxels make the cost per candidate uneven)
#ifndef ASSCPP_CULL_CLAIM
#define ASSCPP_CULL_CLAIM 256
#endif
    constexpr unsigned kCullClaim = ASSCPP_CULL_CLAIM;
    for (;;) {
    unsigned j_begin = 0;
    if (lane == 0) j_begin = atomicAdd(&sh.chunk_next, kCullClaim);
    j_begin = __shfl_sync(0xffffffffu, j_begin, 0);
    if (j_begin >= total) break;
    const unsigned j_end = min(total, j_begin + kCullClaim);
    // c0 / d0: the last descriptor starting at or before the current block of 32 candidates (warp-uniform)
    unsigned c0 = 0;
    {
      unsigned lo = 0, hi = ndesc;  // last descriptor starting at or before j_begin
      while (hi - lo > 1) {
        const unsigned mid = (lo + hi) >> 1;
        if (desc(mid).y <= j_begin) lo = mid; else hi = mid;
      }
      c0 = lo;
    }
    uint2 d0 = desc(c0);
    for (unsigned jb = j_begin; jb < j_end; jb += 32 * kCullUnroll) {
      uint32_t uu[kCullUnroll], q1[kCullUnroll];
      int4 vx[kCullUnroll];
      bool on[kCullUnroll];
#pragma unroll
      for (int b = 0; b < kCullUnroll; ++b) {
        const unsigned j0 = jb + b * 32, j = j0 + lane;
        on[b] = j < j_end; uu[b] = 0; q1[b] = 0; vx[b] = make_int4(0, 0, 0, 0);
        // The chunk of every candidate of the block, found by the warp together: a chunk holds at least one voxel,
        // so the chunks that start inside the block are among the 32 descriptors after c0, one per lane; the bit
        // mask of their starting positions turns "my descriptor" into a population count and two shuffles.
        const unsigned ci = c0 + 1 + lane;
        const uint2 dl = ci < ndesc ? desc(ci) : make_uint2(0u, 0xffffffffu);
        const unsigned rel = dl.y - j0;
        const unsigned starts = __reduce_or_sync(0xffffffffu, rel < 32u ? 1u << rel : 0u);
        const unsigned idx = (unsigned)__popc(starts & (0xffffffffu >> (31 - lane)));  // starts at or before j
        const unsigned sx = __shfl_sync(0xffffffffu, dl.x, (idx - 1u) & 31u);
        const unsigned sy = __shfl_sync(0xffffffffu, dl.y, (idx - 1u) & 31u);
        const uint2 d = idx ? make_uint2(sx, sy) : d0;
        const unsigned nst = (unsigned)__popc(starts);
        const unsigned lx = __shfl_sync(0xffffffffu, dl.x, (nst - 1u) & 31u);
        const unsigned ly = __shfl_sync(0xffffffffu, dl.y, (nst - 1u) & 31u);
        if (nst) { d0 = make_uint2(lx, ly); c0 += nst; }
        if (on[b]) {
          uu[b] = d.x + (j - d.y);
          vx[b] = m.vox[uu[b]];
          q1[b] = (uint32_t)m.vox[uu[b] + 1].w;
        }
      }
#pragma unroll
      for (int b = 0; b < kCullUnroll; ++b) {
        if (jb + b * 32 >= j_end) break;
        int cls = 0;  // 0 outside, 1 cut by a plane, 2 inside
        if (on[b]) {
          {
            const float cx = ((float)vx[b].x + 0.5f) * m.leaf, cy = ((float)vx[b].y + 0.5f) * m.leaf,
                        cz = ((float)vx[b].z + 0.5f) * m.leaf;
            // conservative classifier (bounds in sh.vox_hi), not reference arithmetic: FMA allowed; the verdict is
            // accumulated in two predicates (compare instructions run on the ALU pipe, the busiest one)
            bool out = false, cut = false;
#pragma unroll
            for (int k = 0; k < 6; ++k) {
              const float v = __fmaf_rn(sh.planes[4 * k], cx, __fmaf_rn(sh.planes[4 * k + 1], cy,
                              __fmaf_rn(sh.planes[4 * k + 2], cz, sh.planes[4 * k + 3])));
              out |= v > sh.vox_hi[k];
              cut |= !(v < -sh.vox_hi[k]);
            }
            cls = out ? 0 : (cut ? 1 : 2);
          }
          if (cls == 2) {
            const uint32_t q0 = (uint32_t)vx[b].w;
            surv_in += q1[b] - q0;
            if (want_hash)
              for (uint32_t q = q0; q < q1[b]; ++q) shash += mix64((uint64_t)(uint32_t)m.orig[q]);
            bmin[0] = min(bmin[0], vx[b].x); bmin[1] = min(bmin[1], vx[b].y); bmin[2] = min(bmin[2], vx[b].z);
            bmax[0] = max(bmax[0], vx[b].x); bmax[1] = max(bmax[1], vx[b].y); bmax[2] = max(bmax[2], vx[b].z);
          }
        }
        emit_voxels(cls == 2, uu[b] | kVoxInside);
        // cut by a plane: voxels with many points go to the CTA's list (one warp, or part of one, per voxel below),
        // the others to this warp's list (one lane per voxel)
        const bool bigcut = cls == 1 && q1[b] - (uint32_t)vx[b].w >= coop_min;
        const unsigned bb = __ballot_sync(0xffffffffu, bigcut);
        if (bb) {
          unsigned base = 0;
          if (lane == 0) base = atomicAdd(&sh.big_count, (unsigned)__popc(bb));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (bigcut) __stcg(vox_list + (A.list_cap - 1 - (int64_t)(base + __popc(bb & ((1u << lane) - 1)))), uu[b]);
        }
        const unsigned sb = __ballot_sync(0xffffffffu, cls == 1 && !bigcut);
        if (sb) {
          if (cls == 1 && !bigcut) {
            __stcg(cut + ncut + __popc(sb & ((1u << lane) - 1)), uu[b]);
            // its points are read when this warp's claims are done: start them on their way from DRAM now
            asm volatile("prefetch.global.L2 [%0];" ::"l"(m.pts + (uint32_t)vx[b].w));
          }
          ncut += (unsigned)__popc(sb);
          if (ncut + 32 > cut_cap) flush_cut();
        }
      }
    }
    }  // claims
    const long long d1 = prof2 ? clock64() : 0;
    if (prof2) atomicAdd(&A.phase_cycles[10], (unsigned long long)ncut);
    flush_cut();
    const long long d2 = prof2 ? clock64() : 0;
    // the cut voxels with many points (the top of the list): W lanes per voxel -- a whole warp each when there are
    // few, so that a handful of voxels does not become a chain of dependent loads, a quarter warp each when there are many
    __syncthreads();
    if (prof2) atomicAdd(&A.phase_cycles[14], (unsigned long long)(clock64() - d2));  // wait for the other warps' claims
    {
      const unsigned nbig = sh.big_count;
      const int W = nbig <= (unsigned)nwarps ? 32 : (nbig <= 2u * nwarps ? 16 : 8);
      const int sub = lane / W, sl = lane - sub * W, per = 32 / W;
      const unsigned gmask = W == 32 ? 0xffffffffu : (((1u << W) - 1u) << (sub * W));
      for (unsigned i = warp * per + sub; i < nbig; i += nwarps * per) {
        const uint32_t u = __ldcg(vox_list + (A.list_cap - 1 - (int64_t)i));
        const int4 vx = m.vox[u];
        const uint32_t q0 = (uint32_t)vx.w, q1 = (uint32_t)m.vox[u + 1].w;
        unsigned c = 0;
        for (uint32_t qb = q0; qb < q1; qb += 4 * W) {
          float4 p[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t q = qb + k * W + sl;
            p[k] = q < q1 ? ld_pt(&m.pts[q]) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t q = qb + k * W + sl;
            if (q >= q1 || !point_in_frustum(sh.planes, p[k].x, p[k].y, p[k].z)) continue;
            ++c;
            if (want_hash) shash += mix64((uint64_t)(uint32_t)m.orig[q]);
          }
        }
        c = __reduce_add_sync(gmask, c);
        if (c && sl == 0) {
          surv_in += c;
          bmin[0] = min(bmin[0], vx.x); bmin[1] = min(bmin[1], vx.y); bmin[2] = min(bmin[2], vx.z);
          bmax[0] = max(bmax[0], vx.x); bmax[1] = max(bmax[1], vx.y); bmax[2] = max(bmax[2], vx.z);
          const unsigned at = atomicAdd(&sh.list_count, 1u);
          if (at < A.list_cap) st_list(vox_list + at, u);
        }
      }
      __syncwarp();
    }
    if (prof2) {
      atomicAdd(&A.phase_cycles[8], (unsigned long long)(d1 - d0));
      atomicAdd(&A.phase_cycles[9], (unsigned long long)(clock64() - d1));
      atomicAdd(&A.phase_cycles[11], (unsigned long long)total);
    }
  }
  // brute mode: exact test of every point, 32 points per warp step
  const int64_t nwork = brute ? m.n_chunks : (by_chunk ? (int64_t)(sh.cand_state >> 32) : 0);
  for (;;) {
    // chunks are handed out one at a time: inside chunks are cheap, boundary chunks are not
    unsigned w = 0;
    if (lane == 0) w = atomicAdd(&sh.chunk_next, 1u);
    w = __shfl_sync(0xffffffffu, w, 0);
    if ((int64_t)w >= nwork) break;
    const int64_t ch = by_chunk ? (int64_t)(w < kDescShared ? sh.chunk_desc[w].x : chunk_list[w].x) : (int64_t)w;
    uint32_t pb, pe;
    pb = __float_as_uint(m.chunk_lo[ch].w); pe = __float_as_uint(m.chunk_hi[ch].w);
    for (uint32_t base0 = pb; base0 < pe; base0 += 128) {
      // four independent 512-byte loads in flight per warp before any is consumed (latency, not bandwidth,
      // bounds this phase: a chunk holds at most kChunkMaxPoints = 128 points unless it is one big voxel)
      float4 pt[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint32_t pos = base0 + 32 * k + lane;
        pt[k] = pos < pe ? ld_pt(&m.pts[pos]) : make_float4(0.f, 0.f, 0.f, __uint_as_float(0xffffffffu));
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t base = base0 + 32 * k;
        if (base >= pe) break;
        const uint32_t pos = base + lane;
        const float4 p = pt[k];
        const uint32_t u = __float_as_uint(p.w);
        const bool pass = pos < pe && point_in_frustum(sh.planes, p.x, p.y, p.z);
        unsigned pm = __ballot_sync(0xffffffffu, pass);
        if (!pm) continue;
        surv += (lane == 0) ? (unsigned)__popc(pm) : 0u;
        if (want_hash && pass) shash += mix64((uint64_t)(uint32_t)m.orig[pos]);
        // first passing lane of every run of equal voxel ids emits the voxel
        uint32_t u_prev = __shfl_up_sync(0xffffffffu, u, 1);
        bool head = (lane == 0) || (u != u_prev);
        unsigned hb = __ballot_sync(0xffffffffu, head);
        unsigned below = hb & ((2u << lane) - 1);          // heads at or below this lane
        int my_head = 31 - __clz(below);
        unsigned run_lower = pm & ((1u << lane) - 1) & ~((1u << my_head) - 1);  // passing lanes of my run below me
        bool emit = pass && (run_lower == 0);
        unsigned eb = __ballot_sync(0xffffffffu, emit);
        unsigned slot = 0;
        if (lane == 0) slot = atomicAdd(&sh.list_count, (unsigned)__popc(eb));
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (emit) {
          unsigned at = slot + __popc(eb & ((1u << lane) - 1));
          if (at < A.list_cap) st_list(vox_list + at, u);
          // voxel coordinates as <PCL getGridCoordinates>: floor(x * inverse_leaf) (same value as m.vox[u])
          int vi = (int)floorf(fmul(p.x, m.inv_leaf)), vj = (int)floorf(fmul(p.y, m.inv_leaf)),
              vk = (int)floorf(fmul(p.z, m.inv_leaf));
          bmin[0] = min(bmin[0], vi); bmin[1] = min(bmin[1], vj); bmin[2] = min(bmin[2], vk);
          bmax[0] = max(bmax[0], vi); bmax[1] = max(bmax[1], vj); bmax[2] = max(bmax[2], vk);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    int lo = __reduce_min_sync(0xffffffffu, bmin[c]), hi = __reduce_max_sync(0xffffffffu, bmax[c]);
    if (lane == 0 && lo <= hi) { atomicMin(&sh.bb_min[c], lo); atomicMax(&sh.bb_max[c], hi); }
  }
  surv_in = warp_sum(surv_in);
  if (lane == 0 && (surv | surv_in)) atomicAdd(&sh.survivors, surv + surv_in);
  if (want_hash) {
    for (int o = 16; o; o >>= 1) shash += __shfl_xor_sync(0xffffffffu, shash, o);
    if (lane == 0 && shash) atomicAdd(&sh.surv_hash, shash);
  }
  __syncthreads();
}

// two instances, chosen per model: the fine-lattice one compiles exactly as it did before the coarse form existed
__device__ __forceinline__ void cull_view(const EvalArgs& A, ViewShared& sh, uint2* chunk_list, uint32_t* vox_list,
                                          uint32_t* cut_list, unsigned cut_cap) {
  if (A.chunk_cull) cull_view_t<true>(A, sh, chunk_list, vox_list, cut_list, cut_cap);
  else cull_view_t<false>(A, sh, chunk_list, vox_list, cut_list, cut_cap);
}

// ------------------------------------------------------------------------------------------------
// P3: grid geometry from the survivors' box, zero + fill the padded bit-grid, drop duplicate list
// entries.  Returns where the grid lives (shared or global).  Must be called by all threads.
// ------------------------------------------------------------------------------------------------
template <class F>
__device__ __forceinline__ void or_bit_range(F&& or_word, unsigned lo, unsigned hi) {  // bits [lo, hi)
  if (hi <= lo) return;
  unsigned w0 = lo >> 5, w1 = (hi - 1) >> 5;
  for (unsigned w = w0; w <= w1; ++w) {
    uint32_t mask = 0xffffffffu;
    if (w == w0) mask &= 0xffffffffu << (lo & 31);
    if (w == w1) mask &= 0xffffffffu >> (31 - ((hi - 1) & 31));
    or_word(w, mask);
  }
}

__device__ __forceinline__ bool build_grid(const EvalArgs& A, ViewShared& sh, uint32_t* smem_grid, uint32_t* glob_grid,
                           uint32_t* vox_list, GridMem* gm_out) {
  const ModelDev& m = A.m;
  if (threadIdx.x == 0) {
    sh.big_count = 0;  // the cull's list of many-point voxels is done with; P4b starts its own
    GridView& g = sh.g;
    g.leaf = m.leaf; g.inv = m.inv_leaf; g.half = m.half_leaf; g.thr_sq = m.dist_thr_sq;
    unsigned long long cells = 1, pcells = 1;
    for (int c = 0; c < 3; ++c) {
      // <PCL applyFilter>: min_b = floor(min_p * inv), max_b likewise; floor(x*inv) is monotone in x, so the
      // box of the survivors' voxel coordinates is the same box.
      g.min_b[c] = sh.bb_min[c];
      g.div[c] = sh.bb_max[c] - sh.bb_min[c] + 1;
      g.b_min[c] = fmul((float)sh.bb_min[c], m.leaf);        // voxel_grid_occlusion_estimation.cpp:53-55
      g.b_max[c] = fmul((float)(sh.bb_max[c] + 1), m.leaf);  // :56-58
      g.base[c] = fadd(g.b_min[c], m.half_leaf);
      cells *= (unsigned long long)g.div[c];
      pcells *= (unsigned long long)(g.div[c] + 4);
    }
    g.mul1 = g.div[0];
    g.mul2 = g.div[0] * g.div[1];
    g.cells = (unsigned)cells;
    g.P0 = g.div[0] + 4; g.P1 = g.div[1] + 4; g.P2 = g.div[2] + 4; g.P01 = g.P0 * g.P1;
    // floor(2^32 / d) for d >= 2 in 32-bit arithmetic (a 64-bit division is several hundred cycles on this
    // one-thread path): floor((2^32 - 1) / d), plus one when the remainder is d - 1
    auto inv32 = [](unsigned d) { const unsigned q = 0xffffffffu / d; return q + ((0xffffffffu - q * d) == d - 1u ? 1u : 0u); };
    g.mP0 = inv32((unsigned)g.P0);
    g.mP01 = inv32((unsigned)g.P01);
    g.mD0 = g.mul1 > 1 ? inv32((unsigned)g.mul1) : 0xffffffffu;
    g.mD01 = g.mul2 > 1 ? inv32((unsigned)g.mul2) : 0xffffffffu;
    unsigned long long words = (pcells + 31) / 32;
    sh.words = (unsigned)words;
    bool fits_smem = words <= A.smem_grid_words && !(A.flags & ASSCPP_EVAL_FORCE_GLOBAL);
    sh.use_smem = fits_smem ? 1 : 0;
    sh.status |= fits_smem ? ASSCPP_VS_GRID_SHARED : ASSCPP_VS_GRID_GLOBAL;
    if ((!fits_smem && words > (unsigned long long)A.grid_cap_words) || sh.list_count > A.list_cap ||
        pcells > 0x7fffffffULL) {
      sh.status |= kVsError;  // scratch bound violated: report, do not touch memory out of range
      sh.words = 0;
      sh.list_count = 0;
    }
  }
  __syncthreads();
  const bool use_smem = sh.use_smem != 0;
  uint32_t* grid = use_smem ? smem_grid : glob_grid;
  gm_out->gptr = grid;
  gm_out->saddr = (uint32_t)__cvta_generic_to_shared(smem_grid);
  for (unsigned w = threadIdx.x; w < sh.words; w += blockDim.x) grid[w] = 0;
  __syncthreads();
  if (sh.words == 0) return use_smem;
  const GridView& g = sh.g;
  const int P0 = g.P0, P1 = g.P1, P2 = g.P2, P01 = g.P01;
  // interior: occupancy of the survivors' voxels (four independent voxel-table loads in flight per thread: the
  // gather from L2 / HBM is what this loop waits for)
  unsigned uniq = 0;
  const unsigned nlist = sh.list_count;
  for (unsigned e0 = threadIdx.x; e0 < nlist; e0 += 4 * blockDim.x) {
    int4 vx[4];
    bool on[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned e = e0 + k * blockDim.x;
      on[k] = e < nlist;
      vx[k] = on[k] ? m.vox[vox_list[e] & kVoxMask] : make_int4(0, 0, 0, 0);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (!on[k]) continue;
      int p = (vx[k].x - g.min_b[0] + 2) + (vx[k].y - g.min_b[1] + 2) * P0 + (vx[k].z - g.min_b[2] + 2) * P01;
      uint32_t bit = 1u << (p & 31);
      uint32_t old = atomicOr(&grid[p >> 5], bit);
      if (old & bit) vox_list[e0 + k * blockDim.x] = 0xffffffffu;  // same voxel emitted by another 32-point load
      else ++uniq;
    }
  }
  // halo + sentinel layers: every cell with a padded coordinate in {0, 1, P-2, P-1}
  auto orw = [&](unsigned w, uint32_t mask) { atomicOr(&grid[w], mask); };
  {  // the two bottom and two top z-slabs, word-parallel
    const unsigned slab = 2u * (unsigned)P01, top = (unsigned)(P2 - 2) * (unsigned)P01, end = (unsigned)P2 * (unsigned)P01;
    for (unsigned w = threadIdx.x; w <= ((slab - 1) >> 5); w += blockDim.x)
      or_bit_range(orw, max(w << 5, 0u), min((w + 1) << 5, slab));
    for (unsigned w = (top >> 5) + threadIdx.x; w <= ((end - 1) >> 5); w += blockDim.x)
      or_bit_range(orw, max(w << 5, top), min((w + 1) << 5, end));
  }
  const int rows = P1 * (P2 - 4);
  for (int t = threadIdx.x; t < rows; t += blockDim.x) {
    const int q2 = 2 + t / P1, q1 = t - (q2 - 2) * P1;
    const unsigned base = (unsigned)q2 * (unsigned)P01 + (unsigned)q1 * (unsigned)P0;
    if (q1 < 2 || q1 >= P1 - 2) {
      or_bit_range(orw, base, base + (unsigned)P0);
    } else {
      or_bit_range(orw, base, base + 2u);
      or_bit_range(orw, base + (unsigned)P0 - 2u, base + (unsigned)P0);
    }
  }
  uniq = warp_sum(uniq);
  if ((threadIdx.x & 31) == 0 && uniq) atomicAdd(&sh.unique, uniq);
  __syncthreads();
  return use_smem;
}

// ------------------------------------------------------------------------------------------------
// visit the survivors of a voxel that turned out visible: counts, gains, hash, optional OR.
// mode 0: count/gain/hash; mode 1: OR into A.accum.
// ------------------------------------------------------------------------------------------------
// number of set bits of `bits` in positions [lo, hi)
__device__ __forceinline__ unsigned count_bit_range(const uint32_t* bits, uint32_t lo, uint32_t hi) {
  unsigned n = 0;
  or_bit_range([&](unsigned w, uint32_t mask) { n += __popc(bits[w] & mask); }, lo, hi);
  return n;
}

__device__ __forceinline__ void visit_voxel_points(const EvalArgs& A, const ViewShared& sh, uint32_t uf, int mode,
                                                   uint32_t* cov_scratch, unsigned& vis, unsigned& gain_p,
                                                   unsigned& gain_v, unsigned long long& vhash) {
  const ModelDev& m = A.m;
  const uint32_t u = uf & kVoxMask;
  const uint32_t pb = (uint32_t)m.vox[u].w, pe = (uint32_t)m.vox[u + 1].w;
  const bool want_hash = (A.flags & ASSCPP_EVAL_HASH) != 0;
  const int32_t vcov = m.vox_cov[u];  // coverage voxel shared by all points of the voxel, or -1
  bool any = false;
  if (uf & kVoxInside) {
    // every point of the voxel passed the frustum test (chunk inside the frustum): no coordinates needed
    any = true;
    if (mode == 0) {
      vis += pe - pb;
      gain_p += (pe - pb) - (A.covered.pts ? count_bit_range(A.covered.pts, pb, pe) : 0u);
      if (want_hash)
        for (uint32_t pos = pb; pos < pe; ++pos) vhash += mix64((uint64_t)(uint32_t)m.orig[pos]);
    } else {
      or_bit_range([&](unsigned w, uint32_t mask) { atomicOr(&A.accum.pts[w], mask); }, pb, pe);
    }
    if (vcov < 0) {
      for (uint32_t pos = pb; pos < pe; ++pos) {
        uint32_t id = (uint32_t)m.pt_cov[pos], bit = 1u << (id & 31);
        if (mode == 0) {
          uint32_t old = atomicOr(&cov_scratch[id >> 5], bit);
          if (!(old & bit) && (!A.covered.vox || !bit_test(A.covered.vox, id))) ++gain_v;
        } else {
          atomicOr(&A.accum.vox[id >> 5], bit);
        }
      }
    }
  } else {
    // four loads in flight: the loads, not the tests, are what this loop waits for
    for (uint32_t p0 = pb; p0 < pe; p0 += 4) {
    float4 pp[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) pp[k] = ld_pt(&m.pts[min(p0 + k, pe - 1)]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t pos = p0 + k;
      if (pos >= pe) break;
      const float4 p = pp[k];
      if (!point_in_frustum(sh.planes, p.x, p.y, p.z)) continue;
      any = true;
      if (mode == 0) {
        ++vis;
        if (!A.covered.pts || !bit_test(A.covered.pts, pos)) ++gain_p;
        if (want_hash) vhash += mix64((uint64_t)(uint32_t)m.orig[pos]);
      } else {
        atomicOr(&A.accum.pts[pos >> 5], 1u << (pos & 31));
      }
      if (vcov < 0) {
        uint32_t id = (uint32_t)m.pt_cov[pos], bit = 1u << (id & 31);
        if (mode == 0) {
          uint32_t old = atomicOr(&cov_scratch[id >> 5], bit);
          if (!(old & bit) && (!A.covered.vox || !bit_test(A.covered.vox, id))) ++gain_v;
        } else {
          atomicOr(&A.accum.vox[id >> 5], bit);
        }
      }
    }
    }
  }
  if (any && vcov >= 0) {
    const uint32_t id = (uint32_t)vcov, bit = 1u << (id & 31);
    if (mode == 0) {
      // several occlusion voxels share a coverage voxel: cov_scratch marks the ones already counted for this view
      uint32_t old = A.cov_direct ? 0u : atomicOr(&cov_scratch[id >> 5], bit);
      if (!(old & bit) && (!A.covered.vox || !bit_test(A.covered.vox, id))) ++gain_v;
    } else {
      atomicOr(&A.accum.vox[id >> 5], bit);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// P4a: one DDA ray per voxel occupied for this view (quirk Q1: every survivor of a voxel shares the
// ray to its cell centre).  Fast path (vstate == nullptr): visible voxels (state != 1 and second gate)
// are appended to vis_list, failed rays (quirk Q5) are counted in sh.carry.  Ordered path: the state
// of every voxel is stored in vstate for the replay in original index order.
// ------------------------------------------------------------------------------------------------
template <bool SMEM>
__device__ __forceinline__ void trace_rays(const EvalArgs& A, ViewShared& sh, const GridMem& gm,
                                           const uint32_t* vox_list, uint32_t* vis_list, uint8_t* vstate = nullptr) {
  const ModelDev& m = A.m;
  const GridView g = sh.g;
  const float origin[3] = {sh.origin[0], sh.origin[1], sh.origin[2]};
  const float pcd_o[3] = {m.pcd_origin[0], m.pcd_origin[1], m.pcd_origin[2]};
  const int lane = threadIdx.x & 31;
  unsigned steps = 0, fails = 0;
  const unsigned count = sh.list_count;
  for (;;) {
    // warps take 32-ray tiles from a shared counter: ray lengths vary, static striding leaves a long tail
    unsigned e0 = 0;
    if (lane == 0) e0 = atomicAdd(&sh.ray_next, 32u);
    e0 = __shfl_sync(0xffffffffu, e0, 0);
    if (e0 >= count) break;
    const unsigned e = e0 + lane;
    uint32_t u = e < count ? vox_list[e] : 0xffffffffu;
    bool visible = false;
    if (u != 0xffffffffu) {
      u &= kVoxMask;
      int4 vx = m.vox[u];
      int rt[3] = {vx.x - g.min_b[0], vx.y - g.min_b[1], vx.z - g.min_b[2]};
      float tc[3] = {centre(g, 0, rt[0]), centre(g, 1, rt[1]), centre(g, 2, rt[2])};
      int state = occlusion_estimation<SMEM>(g, gm, origin, rt, tc, steps);
      if (vstate) {
        vstate[u] = (uint8_t)(state | (second_gate(g, pcd_o, tc) ? 4 : 0));
      } else {
        if (state == 2) ++fails;
        visible = (state == 0) && second_gate(g, pcd_o, tc);  // occlusion_culling.cpp:335-342
      }
    }
    if (!vstate) {
      unsigned vb = __ballot_sync(0xffffffffu, visible);
      if (vb) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(&sh.vis_count, (unsigned)__popc(vb));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (visible) vis_list[base + __popc(vb & ((1u << lane) - 1))] = u;
      }
    }
  }
  fails = warp_sum(fails);
  unsigned long long st64 = steps;
  for (int o = 16; o; o >>= 1) st64 += __shfl_xor_sync(0xffffffffu, st64, o);
  if (lane == 0) {
    if (fails) atomicAdd(&sh.carry, fails);
    if (st64) atomicAdd(&sh.steps, st64);
  }
}

// ------------------------------------------------------------------------------------------------
// P4a, fast path: the same rays as trace_rays, reorganised so that lanes stay busy.
//
//   setup_ray    one thread per occupied voxel, fully converged (a warp prepares kRayBatch list entries at a time): direction, slab test, start cell, t_max /
//                t_delta (:99-128, :410-519) and the first loop iterations (:538-579): the start cell, whose
//                coordinates are still at hand, and up to three more while the ray is in the halo layer.
//                Rays that are already decided leave here (failed / free / occluded); the others are written
//                as 48-byte records to the warp's private queue (7.5 KB per warp, L2-resident).
//   walk         lanes pull records from the queue; a lane that finishes its ray takes the next record instead
//                of idling until the slowest ray of a 32-ray tile is done.  When the queue runs dry the rays still
//                in flight are checkpointed as records (the record format is resumable) and the next batch is
//                prepared, so the setup code never runs with walk state live in registers.  The step loop holds nothing but
//                the bit test and the t_max update: a set bit is first classified on the linear index
//                difference to the target (the 3x3x3 neighbourhood of the target can never occlude, quirk Q4:
//                d^2 <= 3 < 4) and everything else (target reached, occluder, halo alias, sentinel) is parked
//                and resolved for all parked lanes together.  DDA step counts come from the Manhattan
//                distance between start and end cell (every iteration moves one cell along one axis).
//
// Record: a = {p, u, qs0 | qs1 << 16, qs2 | qt2 << 16}   p = padded linear index the walk starts on,
//         b = {t_max x y z, qt0 | qt1 << 16}              qs / qt = padded start / target coordinates,
//         c = {t_delta x y z with the step sign in the (otherwise unused) sign bit, 0}
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dda_step(float& tx, float& ty, float& tz, int& p, float dtx, float dty, float dtz,
                                         int k0, int k1, int k2) {
  // :565-579, ties x -> y -> z by <=; NaN compares false and falls through to z
  const bool sx = (tx <= ty) & (tx <= tz);
  const bool sy = (!sx) & (ty <= tz) & (ty <= tx);
  const bool sz = !(sx | sy);
  tx = sx ? fadd(tx, dtx) : tx;
  ty = sy ? fadd(ty, dty) : ty;
  tz = sz ? fadd(tz, dtz) : tz;
  p += sx ? k0 : (sy ? k1 : k2);
}

// occupancy of the cell at relative coordinates c (halo cells: <PCL getCentroidIndexAt> aliasing, quirk Q3)
template <bool SMEM>
__device__ __forceinline__ bool cell_occupied(const GridView& g, const GridMem& gm, int c0, int c1, int c2) {
  int a0 = c0, a1 = c1, a2 = c2;
  // c in [-1, div]: halo <=> c == -1 or c == div <=> (unsigned)c >= div
  if ((unsigned)c0 >= (unsigned)g.div[0] || (unsigned)c1 >= (unsigned)g.div[1] || (unsigned)c2 >= (unsigned)g.div[2]) {
    const int lin = c0 + c1 * g.mul1 + c2 * g.mul2;
    if ((unsigned)lin >= g.cells) return false;
    a2 = fast_div(lin, g.mul2, g.mD01);
    const int ar = lin - a2 * g.mul2;
    a1 = fast_div(ar, g.mul1, g.mD0);
    a0 = ar - a1 * g.mul1;
  }
  const int pa = (a0 + 2) + (a1 + 2) * g.P0 + (a2 + 2) * g.P01;
  return (grid_load<SMEM>(gm, (unsigned)pa >> 5) >> (pa & 31)) & 1u;
}

// :558-562 for the occupied cell c against the target cell t (relative coordinates)
__device__ __forceinline__ bool is_occluder(const GridView& g, int c0, int c1, int c2, int t0, int t1, int t2) {
  const int o0 = abs(c0 - t0), o1 = abs(c1 - t1), o2 = abs(c2 - t2);
  if (max(o0, max(o1, o2)) >= 3) return true;  // d^2 >= 9 leaf^2: no float32 rounding can bring it under 4 leaf^2
  float dx = fsub(centre(g, 0, c0), centre(g, 0, t0)), dy = fsub(centre(g, 1, c1), centre(g, 1, t1)),
        dz = fsub(centre(g, 2, c2), centre(g, 2, t2));
  float sum = fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));  // :560
  return sum > g.thr_sq;                                             // :561
}

// Record planes of one warp's private queue
struct RayQueue {
  uint4* a;
  float4* b;
  float4* c;
};

// Ray setup for one list entry (u = voxel id or 0xffffffff).  Returns 0 nothing to emit (duplicate entry, failed
// or occluded by its start cell), 1 free (candidate for the second gate), 2 walk (record in ra / rb / rc).
template <bool SMEM>
__device__ __forceinline__ int setup_ray(const ModelDev& m, const GridView& g, const GridMem& gm, const float origin[3],
                                         uint32_t u, uint4& ra, float4& rb, float4& rc, unsigned& steps,
                                         unsigned& fails, int setup_iters, int halo_iters) {
  if (u == 0xffffffffu) return 0;
  const int4 vx = m.vox[u & kVoxMask];
  const int rt[3] = {vx.x - g.min_b[0], vx.y - g.min_b[1], vx.z - g.min_b[2]};
  const float tc[3] = {centre(g, 0, rt[0]), centre(g, 1, rt[1]), centre(g, 2, rt[2])};
  float d[3];
  direction(tc, origin, d);                  // :111-113
  const float tmin = ray_box(g, origin, d);  // :116
  if (tmin == -1.0f) { ++fails; return 0; }  // :118-122, quirk Q5
  float t_max[3], t_delta[3];
  int r[3], step[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float start = fadd(origin[c], fmul(tmin, d[c]));                   // :422
    int ijk = (int)floorf(fadd(fmul(start, g.inv), 0.5f));             // :425
    r[c] = ijk - g.min_b[c];
    float vm = centre(g, c, r[c]);                                     // :432
    if (d[c] >= 0) { vm = fadd(vm, g.half); step[c] = 1; }             // :436-465
    else           { vm = fsub(vm, g.half); step[c] = -1; }
    t_max[c] = fadd(tmin, fdiv(fsub(vm, start), d[c]));                // :513-515
    t_delta[c] = fdiv(g.leaf, fabsf(d[c]));                            // :517-519
  }
  if (!((unsigned)(r[0] + 1) <= (unsigned)(g.div[0] + 1) && (unsigned)(r[1] + 1) <= (unsigned)(g.div[1] + 1) &&
        (unsigned)(r[2] + 1) <= (unsigned)(g.div[2] + 1)))
    return 1;                                // :538-540 false on entry: result 0, no iteration
  // the first iterations (:538-579) here, where the cell coordinates are at hand and all lanes are converged: at
  // least one, then on while the walk is still in the halo layer (a cell outside the survivors' box, the expensive
  // case for the walk) or has done fewer than setup_iters.  Short rays never reach the queue.  Rays that end here
  // count their iterations; rays that go on are counted in the walk from their start cell.
  const int P0 = g.P0, P01 = g.P01;
  const int rs[3] = {r[0], r[1], r[2]};
  float tx = t_max[0], ty = t_max[1], tz = t_max[2];
#pragma unroll 1
  for (int it = 0;; ++it) {
    if (r[0] == rt[0] && r[1] == rt[1] && r[2] == rt[2]) { steps += it + 1; return 1; }  // :551-552
    if (cell_occupied<SMEM>(g, gm, r[0], r[1], r[2]) && is_occluder(g, r[0], r[1], r[2], rt[0], rt[1], rt[2])) {
      steps += it + 1;                       // occluded (early exit: the result stays 1)
      return 0;
    }
    // :565-579, ties x -> y -> z by <=; NaN compares false and falls through to z
    const bool sx = (tx <= ty) & (tx <= tz);
    const bool sy = (!sx) & (ty <= tz) & (ty <= tx);
    if (sx) { tx = fadd(tx, t_delta[0]); r[0] += step[0]; }
    else if (sy) { ty = fadd(ty, t_delta[1]); r[1] += step[1]; }
    else { tz = fadd(tz, t_delta[2]); r[2] += step[2]; }
    if (!((unsigned)(r[0] + 1) <= (unsigned)(g.div[0] + 1) && (unsigned)(r[1] + 1) <= (unsigned)(g.div[1] + 1) &&
          (unsigned)(r[2] + 1) <= (unsigned)(g.div[2] + 1))) {
      steps += it + 1;                       // :538-540 false: left the expanded box, result 0
      return 1;
    }
    const bool interior = (unsigned)r[0] < (unsigned)g.div[0] && (unsigned)r[1] < (unsigned)g.div[1] &&
                          (unsigned)r[2] < (unsigned)g.div[2];
    if ((interior && it + 1 >= setup_iters) || it + 1 >= setup_iters + halo_iters) break;
  }
  if (tx != tx) ty = tx;  // tx NaN: the reference steps z from here on; see dda_burst for why ty says so too
  const int p = (r[0] + 2) + (r[1] + 2) * P0 + (r[2] + 2) * P01;
  ra = make_uint4((unsigned)p, u, (unsigned)(rs[0] + 2) | ((unsigned)(rs[1] + 2) << 16),
                  (unsigned)(rs[2] + 2) | ((unsigned)(rt[2] + 2) << 16));
  rb = make_float4(tx, ty, tz, __uint_as_float((unsigned)(rt[0] + 2) | ((unsigned)(rt[1] + 2) << 16)));
  rc = make_float4(__uint_as_float(__float_as_uint(t_delta[0]) & 0x7fffffffu | (step[0] < 0 ? 0x80000000u : 0u)),
                   __uint_as_float(__float_as_uint(t_delta[1]) & 0x7fffffffu | (step[1] < 0 ? 0x80000000u : 0u)),
                   __uint_as_float(__float_as_uint(t_delta[2]) & 0x7fffffffu | (step[2] < 0 ? 0x80000000u : 0u)),
                   0.0f);
  return 2;
}

// The burst loop in PTX: three DDA iterations per lane, then a warp vote; leaves when fewer than `thresh`
// lanes are still walking.  A lane stops walking at the first set bit (pbit then addresses that cell).
// Branch-free per iteration: bit test, the x -> y -> z `<=` cascade of :565-579 as predicates, predicated FADD
// (add.rn: never fused) and index updates.  pbit = padded linear index (+ 8 * shared byte address of the grid in
// the SMEM variant, so pbit >> 3 is the address of the cell's byte and pbit & 7 its bit).
//
// The loop is bound by the ALU pipe (integer / logic / compare: half the rate of the FMA pipe), so it is kept to
// eight ALU instructions per iteration:
//   * the cell's BYTE is loaded (no word-address mask) and replicated into the four bytes of a register by an
//     integer multiply (FMA pipe), so that the funnel shift by pbit (mod 32) finds bit pbit & 7;
//   * `walking` is not carried from one iteration to the next: a lane that is not walking rests on a set bit (a
//     parked ray on the bit that stopped it, a lane without a ray on that bit or on padded cell 0, which is in the
//     sentinel layer), so the bit test alone says whether the lane walks -- every caller keeps that invariant;
//   * sy = !sx & (ty <= tz) & (ty <= tx) is evaluated without its last term.  With sx false and ty <= tz, either
//     tx > ty (the term holds) or tx <= ty and tx > tz, which contradicts ty <= tz -- unless tx is NaN, where the
//     reference takes z for ever; rays are handed to the walk with ty = NaN in that case (setup_ray), which
//     forces the same choice, and tx cannot become NaN later (it only grows by a t_delta that is NaN itself or
//     leaves -inf + inf to the very first step, which setup_ray takes).
#define ASSCPP_DDA_ITER(LOAD)                          \
  " shr.u32 a, %0, 3;\n" LOAD                          \
  " mul.lo.u32 word, word, 0x01010101;\n"              \
  " shf.r.wrap.b32 b, word, word, %0;\n"               \
  " and.b32 b, b, 1;\n"                                \
  " setp.eq.u32 pw, b, 0;\n"                           \
  " setp.le.and.f32 q1, %1, %2, pw;\n"                 \
  " setp.le.and.f32 px, %1, %3, q1;\n"                 \
  " xor.pred pn, pw, px;\n"                            \
  " setp.le.and.f32 py, %2, %3, pn;\n"                 \
  " xor.pred pz, pn, py;\n"                            \
  " @px add.rn.f32 %1, %1, %5;\n"                      \
  " @py add.rn.f32 %2, %2, %6;\n"                      \
  " @pz add.rn.f32 %3, %3, %7;\n"                      \
  " @px add.s32 %0, %0, %8;\n"                         \
  " @py add.s32 %0, %0, %9;\n"                         \
  " @pz add.s32 %0, %0, %10;\n"

template <bool SMEM>
__device__ __forceinline__ void dda_burst(unsigned& pbit, float& tx, float& ty, float& tz, unsigned& walking,
                                          float dtx, float dty, float dtz, int k0, int k1, int k2, unsigned thresh,
                                          const uint32_t* gptr) {
  if constexpr (SMEM) {
    asm volatile(
        "{\n"
        " .reg .pred pw, px, py, pz, pn, q1, pc;\n"
        " .reg .u32 a, word, b, bal, cnt;\n"
        "LB_%=:\n"
        ASSCPP_DDA_ITER(" ld.shared.u8 word, [a];\n")
        ASSCPP_DDA_ITER(" ld.shared.u8 word, [a];\n")
        ASSCPP_DDA_ITER(" ld.shared.u8 word, [a];\n")
        " vote.sync.ballot.b32 bal, pw, 0xffffffff;\n"
        " popc.b32 cnt, bal;\n"
        " setp.ge.u32 pc, cnt, %11;\n"
        " @pc bra.uni LB_%=;\n"
        " selp.u32 %4, 1, 0, pw;\n"
        "}\n"
        : "+r"(pbit), "+f"(tx), "+f"(ty), "+f"(tz), "+r"(walking)
        : "f"(dtx), "f"(dty), "f"(dtz), "r"(k0), "r"(k1), "r"(k2), "r"(thresh)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        " .reg .pred pw, px, py, pz, pn, q1, pc;\n"
        " .reg .u32 a, word, b, bal, cnt;\n"
        " .reg .u64 ga;\n"
        "LB_%=:\n"
        ASSCPP_DDA_ITER(" cvt.u64.u32 ga, a;\n add.u64 ga, ga, %12;\n ld.global.u8 word, [ga];\n")
        ASSCPP_DDA_ITER(" cvt.u64.u32 ga, a;\n add.u64 ga, ga, %12;\n ld.global.u8 word, [ga];\n")
        ASSCPP_DDA_ITER(" cvt.u64.u32 ga, a;\n add.u64 ga, ga, %12;\n ld.global.u8 word, [ga];\n")
        " vote.sync.ballot.b32 bal, pw, 0xffffffff;\n"
        " popc.b32 cnt, bal;\n"
        " setp.ge.u32 pc, cnt, %11;\n"
        " @pc bra.uni LB_%=;\n"
        " selp.u32 %4, 1, 0, pw;\n"
        "}\n"
        : "+r"(pbit), "+f"(tx), "+f"(ty), "+f"(tz), "+r"(walking)
        : "f"(dtx), "f"(dty), "f"(dtz), "r"(k0), "r"(k1), "r"(k2), "r"(thresh), "l"(gptr)
        : "memory");
  }
}

constexpr int kRayBatch = 128;              // list entries a warp claims and prepares at a time
constexpr int kRayQueueCap = kRayBatch + 32;  // + at most 32 checkpointed rays

// One warp: alternate between preparing a batch of rays into its private queue and walking them.
template <bool SMEM>
__device__ __forceinline__ void trace_rays_fast(const EvalArgs& A, ViewShared& sh, const GridMem& gm,
                                                const uint32_t* vox_list, RayQueue q, uint32_t* cand_list) {
  const ModelDev& m = A.m;
  const GridView& g = sh.g;
  const float origin[3] = {sh.origin[0], sh.origin[1], sh.origin[2]};
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1;
  const unsigned total = sh.list_count;
  const int P0 = g.P0, P01 = g.P01;
  const int h0 = P0 >> 1, h01 = P01 >> 1, near_max = P01 + P0 + 1;
  const unsigned bit_base = SMEM ? (gm.saddr << 3) : 0u;  // see dda_burst
  const unsigned T = (unsigned)A.walk_thresh;
  unsigned pbit = bit_base;
  int pt = 0, k0 = 0, k1 = 0, k2 = 0;
  float tx = 0, ty = 0, tz = 0, dtx = 0, dty = 0, dtz = 0;
  uint32_t u = 0, w_s01 = 0, w_s2t2 = 0, w_t01 = 0;
  unsigned walking = 0, have = 0;  // have: the lane owns an unfinished ray (walking or parked)
  unsigned qhead = 0, qtail = 0;   // private queue, warp-uniform
  bool more = true;                // list entries not yet claimed by any warp (warp-uniform)
  unsigned steps_lo = 0, fails = 0;
  unsigned long long steps = 0;
  for (;;) {
    // ---- service: resolve parked lanes, emit free rays, refill idle lanes (converged) ----------------
    unsigned freed = 0;
    if (have && !walking) {
      // parked on a set bit: target, occupied cell, halo or sentinel
      const int p = (int)(pbit - bit_base);
      const int delta = p - pt;
      // cells of the target's 3x3x3 neighbourhood (delta = a + b P0 + c P01, |a|,|b|,|c| <= 1) never occlude
      // (quirk Q4: d^2 <= 3 < 4), whatever their occupancy: step over them
      auto in_near27 = [&](int dl) {
        const int c = (dl > h01) - (dl < -h01);
        const int rem = dl - c * P01;
        const int b = (rem > h0) - (rem < -h0);
        const int a = rem - b * P0;
        return (abs(dl) <= near_max) & (abs(rem) <= P0 + 1) & (abs(a) <= 1);
      };
      const bool near27 = in_near27(delta);
      bool go_on = near27 & (delta != 0);
      if (!go_on) {
        const int qs0 = w_s01 & 0xffff, qs1 = w_s01 >> 16, qs2 = w_s2t2 & 0xffff;
        const int qt0 = w_t01 & 0xffff, qt1 = w_t01 >> 16, qt2 = w_s2t2 >> 16;
        if (delta == 0) {                                                  // :551-552
          steps_lo += (unsigned)(abs(qt0 - qs0) + abs(qt1 - qs1) + abs(qt2 - qs2) + 1);
          freed = 1;
        } else {
          const int q2 = fast_div(p, P01, g.mP01), rm = p - q2 * P01, q1 = fast_div(rm, P0, g.mP0), q0 = rm - q1 * P0;
          const unsigned walked = (unsigned)(abs(q0 - qs0) + abs(q1 - qs1) + abs(q2 - qs2));
          // q in [0, P - 1]: sentinel layer <=> q == 0 or q == P - 1 <=> (unsigned)(q - 1) >= P - 2
          if ((unsigned)(q0 - 1) >= (unsigned)(P0 - 2) || (unsigned)(q1 - 1) >= (unsigned)(g.P1 - 2) ||
              (unsigned)(q2 - 1) >= (unsigned)(g.P2 - 2)) {
            steps_lo += walked;                                            // left the expanded box (:538-540): result 0
            freed = 1;
          } else if ((((unsigned)(q0 - 2) < (unsigned)(P0 - 4) && (unsigned)(q1 - 2) < (unsigned)(g.P1 - 4) && (unsigned)(q2 - 2) < (unsigned)(g.P2 - 4)) ||  // interior: the set bit IS the occupancy
                      cell_occupied<SMEM>(g, gm, q0 - 2, q1 - 2, q2 - 2)) &&                              // halo: aliased lookup (Q3)
                     is_occluder(g, q0 - 2, q1 - 2, q2 - 2, qt0 - 2, qt1 - 2, qt2 - 2)) {
            steps_lo += walked + 1;                                        // :561-562 result 1
          } else {
            go_on = true;
          }
        }
      }
      if (go_on) {
        int pp = p;
        dda_step(tx, ty, tz, pp, dtx, dty, dtz, k0, k1, k2);
        if (near27) {
          // Inside the target's neighbourhood nothing occludes, occupied or not: walk on here -- the target is at most
          // three cells away -- instead of parking on every occupied neighbour (surface voxels have several).  The
          // same float steps in the same order; a ray that leaves the neighbourhood goes back to the step loop.
#pragma unroll 1
          for (int it = 0; it < 4; ++it) {
            const int d2 = pp - pt;
            if (d2 == 0) {                                                  // :551-552
              steps_lo += (unsigned)(abs((int)(w_t01 & 0xffff) - (int)(w_s01 & 0xffff)) + abs((int)(w_t01 >> 16) - (int)(w_s01 >> 16)) +
                                     abs((int)(w_s2t2 >> 16) - (int)(w_s2t2 & 0xffff)) + 1);
              freed = 1;
              go_on = false;
              break;
            }
            if (!in_near27(d2)) break;
            dda_step(tx, ty, tz, pp, dtx, dty, dtz, k0, k1, k2);
          }
        }
        if (go_on) {
          pbit = (unsigned)pp + bit_base;
          walking = 1;
        } else {
          have = 0;
        }
      } else {
        have = 0;
      }
    }
    const unsigned fb = __ballot_sync(0xffffffffu, freed != 0);
    if (fb) {
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(&sh.vis_count, (unsigned)__popc(fb));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (freed) st_list(cand_list + base + __popc(fb & lt), u);
    }
    unsigned idle = __ballot_sync(0xffffffffu, have == 0);
    if (idle && qhead < qtail) {
      if (!have) {
        const unsigned at = qhead + __popc(idle & lt);
        if (at < qtail) {
          const uint4 ra = __ldcg(q.a + at);
          const float4 rb = __ldcg(q.b + at);
          const float4 rc = __ldcg(q.c + at);
          pbit = ra.x + bit_base; u = ra.y; w_s01 = ra.z; w_s2t2 = ra.w;
          tx = rb.x; ty = rb.y; tz = rb.z; w_t01 = __float_as_uint(rb.w);
          dtx = fabsf(rc.x); dty = fabsf(rc.y); dtz = fabsf(rc.z);
          k0 = (__float_as_uint(rc.x) >> 31) ? -1 : 1;
          k1 = (__float_as_uint(rc.y) >> 31) ? -P0 : P0;
          k2 = (__float_as_uint(rc.z) >> 31) ? -P01 : P01;
          pt = (int)(w_t01 & 0xffff) + (int)(w_t01 >> 16) * P0 + (int)(w_s2t2 >> 16) * P01;
          walking = 1; have = 1;
        }
      }
      qhead = min(qtail, qhead + (unsigned)__popc(idle));
      idle = __ballot_sync(0xffffffffu, have == 0);
    }
    const unsigned nhave = 32u - (unsigned)__popc(idle);
    if (qhead == qtail && more && nhave < T) {
      // queue empty and too few rays in flight: checkpoint them as records, then prepare the next batch
      const unsigned hb = ~idle;
      if (have) {
        const unsigned at = (unsigned)__popc(hb & lt);
        __stcg(q.a + at, make_uint4(pbit - bit_base, u, w_s01, w_s2t2));
        __stcg(q.b + at, make_float4(tx, ty, tz, __uint_as_float(w_t01)));
        __stcg(q.c + at, make_float4(__uint_as_float(__float_as_uint(dtx) | (k0 < 0 ? 0x80000000u : 0u)),
                                     __uint_as_float(__float_as_uint(dty) | (k1 < 0 ? 0x80000000u : 0u)),
                                     __uint_as_float(__float_as_uint(dtz) | (k2 < 0 ? 0x80000000u : 0u)), 0.0f));
      }
      qhead = 0;
      qtail = nhave;
      have = 0; walking = 0;  // nothing of the walk state is live across the setup code below
      pbit = bit_base; pt = 0; k0 = k1 = k2 = 0; tx = ty = tz = dtx = dty = dtz = 0; u = w_s01 = w_s2t2 = w_t01 = 0;
      // guided claim: full batches while plenty of entries remain, smaller ones near the end of the list so
      // that the warps finish together
      unsigned e0 = 0, bsz = 0;
      if (lane == 0) {
        const unsigned seen = *(volatile unsigned*)&sh.ray_next;
        const unsigned rem = seen < total ? total - seen : 0u;
        bsz = max(32u, min((unsigned)kRayBatch, (rem / (2u * (blockDim.x >> 5))) & ~31u));
        e0 = atomicAdd(&sh.ray_next, bsz);
      }
      e0 = __shfl_sync(0xffffffffu, e0, 0);
      bsz = __shfl_sync(0xffffffffu, bsz, 0);
      if (e0 >= total) {
        more = false;
      } else {
        const unsigned e1 = min(total, e0 + bsz);
        for (unsigned eb = e0; eb < e1; eb += 32) {
          const unsigned e = eb + lane;
          const uint32_t uu = e < e1 ? vox_list[e] : 0xffffffffu;
          uint4 ra = make_uint4(0, 0, 0, 0);
          float4 rb = make_float4(0, 0, 0, 0), rc = make_float4(0, 0, 0, 0);
          const int kind = setup_ray<SMEM>(m, g, gm, origin, uu, ra, rb, rc, steps_lo, fails, A.setup_iters, A.halo_iters);
          const unsigned cb = __ballot_sync(0xffffffffu, kind == 1);
          if (cb) {
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(&sh.vis_count, (unsigned)__popc(cb));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (kind == 1) st_list(cand_list + base + __popc(cb & lt), uu);
          }
          const unsigned wb = __ballot_sync(0xffffffffu, kind == 2);
          if (kind == 2) {
            const unsigned at = qtail + __popc(wb & lt);
            __stcg(q.a + at, ra);
            __stcg(q.b + at, rb);
            __stcg(q.c + at, rc);
          }
          qtail += (unsigned)__popc(wb);
        }
      }
      __syncwarp();
      steps += steps_lo; steps_lo = 0;
      continue;
    }
    if (nhave == 0) break;  // nothing in flight, queue empty, no entries left
    // ---- burst: step until too few lanes are still walking -----------------------------------------
    if (__any_sync(0xffffffffu, walking != 0))
      dda_burst<SMEM>(pbit, tx, ty, tz, walking, dtx, dty, dtz, k0, k1, k2, (qhead == qtail && !more) ? 1u : T, gm.gptr);
  }
  steps += steps_lo;
  fails = warp_sum(fails);
  for (int o = 16; o; o >>= 1) steps += __shfl_xor_sync(0xffffffffu, steps, o);
  if (lane == 0) {
    if (fails) atomicAdd(&sh.carry, fails);
    if (steps) atomicAdd(&sh.steps, steps);
  }
}

// ------------------------------------------------------------------------------------------------
// accuracy terms (ASSCPP_EVAL_ACCURACY): occlusion_culling.cpp:445-475, 505-513, 538-618
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int float_key(float f) {  // order-preserving map float -> int (for atomicMin / atomicMax)
  int k = __float_as_int(f);
  return k >= 0 ? k : k ^ 0x7fffffff;
}

struct AccLocal {
  unsigned long long d_hi = 0, d_lo = 0, z_hi = 0, z_lo = 0;
  float mn = INFINITY, mx = -INFINITY;
};

// value * scale (scale = 2^S, exact) split into hi * 2^32 + lo and added to integer accumulators: integer adds
// commute, so the sum does not depend on the order in which threads visit the points
__device__ __forceinline__ void acc_fixed(double e, double scale, unsigned long long& hi, unsigned long long& lo) {
  const double v = __dmul_rn(e, scale);
  const unsigned long long h = __double2ull_rz(__dmul_rn(v, 2.3283064365386963e-10));  // v * 2^-32
  hi += h;
  lo += __double2ull_rz(__dsub_rn(v, __dmul_rn((double)h, 4294967296.0)));
}

// one visible point: depth = x of pointCloudViewportTransform (translate, rotate about z, rotate about y; every
// stage in double on float32 inputs and narrowed to float32, transformPointMatVec :521-531), then the two error
// terms 0.0000285 * val * val of calcAvgAccuracy (:452, :469)
__device__ __forceinline__ void acc_point(const ViewTrig& tr, double sc_d, double sc_z, float px, float py, float pz,
                                          AccLocal& a) {
  const float a0 = (float)__dadd_rn((double)px, tr.t[0]);
  const float a1 = (float)__dadd_rn((double)py, tr.t[1]);
  const float a2 = (float)__dadd_rn((double)pz, tr.t[2]);
  const float b0 = (float)__dadd_rn(__dmul_rn(tr.cz, (double)a0), __dmul_rn(-tr.sz, (double)a1));
  const float b2 = a2;
  const float depth = (float)__dadd_rn(__dmul_rn(tr.cy, (double)b0), __dmul_rn(tr.sy, (double)b2));
  const double val = (double)depth;
  acc_fixed(__dmul_rn(__dmul_rn(0.0000285, val), val), sc_d, a.d_hi, a.d_lo);
  const double vz = (double)pz;
  acc_fixed(__dmul_rn(__dmul_rn(0.0000285, vz), vz), sc_z, a.z_hi, a.z_lo);
  a.mn = fminf(a.mn, depth);
  a.mx = fmaxf(a.mx, depth);
}

__device__ __forceinline__ void acc_commit(ViewShared& sh, AccLocal& a) {
  for (int o = 16; o; o >>= 1) {
    a.d_hi += __shfl_xor_sync(0xffffffffu, a.d_hi, o);
    a.d_lo += __shfl_xor_sync(0xffffffffu, a.d_lo, o);
    a.z_hi += __shfl_xor_sync(0xffffffffu, a.z_hi, o);
    a.z_lo += __shfl_xor_sync(0xffffffffu, a.z_lo, o);
    a.mn = fminf(a.mn, __shfl_xor_sync(0xffffffffu, a.mn, o));
    a.mx = fmaxf(a.mx, __shfl_xor_sync(0xffffffffu, a.mx, o));
  }
  if ((threadIdx.x & 31) == 0 && (a.d_hi | a.d_lo | a.z_hi | a.z_lo | (a.mn <= a.mx))) {
    atomicAdd(&sh.acc[0], a.d_hi); atomicAdd(&sh.acc[1], a.d_lo);
    atomicAdd(&sh.acc[2], a.z_hi); atomicAdd(&sh.acc[3], a.z_lo);
    atomicMin(&sh.acc_min, float_key(a.mn));
    atomicMax(&sh.acc_max, float_key(a.mx));
  }
}

// ------------------------------------------------------------------------------------------------
// One warp, one voxel (every lane passes the same entry): visit_voxel_points with the points taken 32 at a time, four
// loads in flight per lane.  MODE 0: counts / gains / hash; 1: OR into A.accum; 2: accuracy terms.
// ------------------------------------------------------------------------------------------------
struct LaneGroup {  // W consecutive lanes of a warp working on one voxel
  int W, sub, sl;     // lanes per group, group index in the warp, lane index in the group
  unsigned mask;
};
__device__ __forceinline__ LaneGroup lane_group(unsigned nvoxels, int nwarps) {
  LaneGroup g;
  const int lane = threadIdx.x & 31;
  g.W = nvoxels <= (unsigned)nwarps ? 32 : (nvoxels <= 2u * nwarps ? 16 : 8);
  g.sub = lane / g.W;
  g.sl = lane - g.sub * g.W;
  g.mask = g.W == 32 ? 0xffffffffu : (((1u << g.W) - 1u) << (g.sub * g.W));
  return g;
}

template <int MODE>
__device__ __forceinline__ void visit_voxel_warp(const EvalArgs& A, const ViewShared& sh, const LaneGroup& lg, uint32_t uf,
                                                 uint32_t* cov_scratch, unsigned& vis, unsigned& gain_p, unsigned& gain_v,
                                                 unsigned long long& vhash, const ViewTrig* tr, AccLocal* acc) {
  const ModelDev& m = A.m;
  const int lane = lg.sl, W = lg.W;
  const uint32_t u = uf & kVoxMask;
  const uint32_t pb = (uint32_t)m.vox[u].w, pe = (uint32_t)m.vox[u + 1].w;
  const bool inside = (uf & kVoxInside) != 0;  // every point of the voxel passed the frustum test
  const bool want_hash = (A.flags & ASSCPP_EVAL_HASH) != 0;
  const int32_t vcov = MODE == 2 ? 0 : m.vox_cov[u];
  bool any = false;
  constexpr int K = 4;
  for (uint32_t p0 = pb; p0 < pe; p0 += W * K) {
    float4 p[K];
    bool in[K];
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const uint32_t pos = p0 + W * k + lane;
      in[k] = pos < pe;
      p[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (in[k] && (!inside || MODE == 2)) p[k] = ld_pt(&m.pts[pos]);
    }
#pragma unroll
    for (int k = 0; k < K; ++k) {
      if (p0 + W * k >= pe) break;  // uniform over the group
      const uint32_t pos = p0 + W * k + lane;
      const bool on = in[k] && (inside || point_in_frustum(sh.planes, p[k].x, p[k].y, p[k].z));
      any |= on;
      if (MODE == 1) {
        // the W positions straddle at most two words of the point bitset: two atomics per step of the group
        const unsigned mb = __ballot_sync(lg.mask, on) >> (lg.sub * W);
        const uint32_t base = p0 + W * k, sft = base & 31;
        if (lane == 0 && (mb << sft)) atomicOr(&A.accum.pts[base >> 5], mb << sft);
        if (lane == 1 && sft && (mb >> (32 - sft))) atomicOr(&A.accum.pts[(base >> 5) + 1], mb >> (32 - sft));
      }
      if (!on) continue;
      if (MODE == 0) {
        ++vis;
        if (!A.covered.pts || !bit_test(A.covered.pts, pos)) ++gain_p;
        if (want_hash) vhash += mix64((uint64_t)(uint32_t)m.orig[pos]);
      } else if (MODE == 2) {
        acc_point(*tr, A.acc_scale_d, A.acc_scale_z, p[k].x, p[k].y, p[k].z, *acc);
      }
      if (MODE != 2 && vcov < 0) {
        const uint32_t id = (uint32_t)m.pt_cov[pos], bit = 1u << (id & 31);
        if (MODE == 0) {
          const uint32_t old = atomicOr(&cov_scratch[id >> 5], bit);
          if (!(old & bit) && (!A.covered.vox || !bit_test(A.covered.vox, id))) ++gain_v;
        } else {
          atomicOr(&A.accum.vox[id >> 5], bit);
        }
      }
    }
  }
  if (MODE != 2 && vcov >= 0 && __any_sync(lg.mask, any) && lane == 0) {
    const uint32_t id = (uint32_t)vcov, bit = 1u << (id & 31);
    if (MODE == 0) {
      const uint32_t old = A.cov_direct ? 0u : atomicOr(&cov_scratch[id >> 5], bit);
      if (!(old & bit) && (!A.covered.vox || !bit_test(A.covered.vox, id))) ++gain_v;
    } else {
      atomicOr(&A.accum.vox[id >> 5], bit);
    }
  }
}

// pointCloudViewportTransform constants from device-resident poses (asscpp_eval_batch_device).  The host path
// computes them with libm instead (abi.cu): CUDA's double asin / atan2 / sin / cos may differ from libm in the
// last ulp.
__global__ void trig_kernel(const double* __restrict__ poses, int64_t nviews, ViewTrig* __restrict__ out) {
  int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nviews) return;
  const double* ps = poses + 7 * v;
  double x = ps[3], y = ps[4], z = ps[5], w = ps[6];
  double d = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z)), __dmul_rn(w, w));
  double s = __ddiv_rn(2.0, d);
  double xs = __dmul_rn(x, s), ys = __dmul_rn(y, s), zs = __dmul_rn(z, s);
  double wy = __dmul_rn(w, ys), wz = __dmul_rn(w, zs);
  double xy = __dmul_rn(x, ys), xz = __dmul_rn(x, zs);
  double yy = __dmul_rn(y, ys), zz = __dmul_rn(z, zs);
  double r00 = __dsub_rn(1.0, __dadd_rn(yy, zz)), r10 = __dadd_rn(xy, wz), r20 = __dsub_rn(xz, wy);
  double yaw, pitch;
  if (fabs(r20) >= 1) {
    yaw = 0;
    pitch = r20 < 0 ? 1.5707963267948966 : -1.5707963267948966;
  } else {
    pitch = -asin(r20);
    yaw = atan2(__ddiv_rn(r10, cos(pitch)), __ddiv_rn(r00, cos(pitch)));
  }
  ViewTrig t;
  t.t[0] = -ps[0]; t.t[1] = -ps[1]; t.t[2] = -ps[2];
  t.cz = cos(-yaw); t.sz = sin(-yaw); t.cy = cos(-pitch); t.sy = sin(-pitch);
  t.pad = 0;
  out[v] = t;
}

__device__ __forceinline__ void reset_view(const EvalArgs& A, ViewShared& sh, const double* pose, const ViewDev* param_view,
                                           int v) {
  if (param_view) {
    // tiny batches: the planes themselves are kernel parameters (computed by the host, same arithmetic)
    if (threadIdx.x < 24) sh.planes[threadIdx.x] = param_view->planes[threadIdx.x];
    if (threadIdx.x < 3) sh.origin[threadIdx.x] = param_view->origin[threadIdx.x];
  } else if (pose) {
    // small batches: the planes of this view, computed here instead of by a planes_kernel launch
    if (threadIdx.x == 32) {
      ViewDev vd;
      compute_view(pose, A.fc, vd);
#pragma unroll
      for (int k = 0; k < 24; ++k) sh.planes[k] = vd.planes[k];
      sh.origin[0] = vd.origin[0]; sh.origin[1] = vd.origin[1]; sh.origin[2] = vd.origin[2];
    }
  } else {
    const ViewDev& vd = A.views[v];
    if (threadIdx.x < 24) sh.planes[threadIdx.x] = vd.planes[threadIdx.x];
    if (threadIdx.x < 3) sh.origin[threadIdx.x] = vd.origin[threadIdx.x];
  }
  if (threadIdx.x < 3) {
    sh.bb_min[threadIdx.x] = INT_MAX;
    sh.bb_max[threadIdx.x] = INT_MIN;
  }
  if (A.acc && threadIdx.x >= 64 && threadIdx.x < 72)
    reinterpret_cast<double*>(&sh.trig)[threadIdx.x - 64] = reinterpret_cast<const double*>(&A.trig[v])[threadIdx.x - 64];
  if (threadIdx.x == 0) {
    sh.view = v;
    sh.cand_state = 0; sh.big_count = 0; sh.chunk_next = 0; sh.super_count = 0; sh.list_count = 0; sh.vis_count = 0; sh.ray_next = 0; sh.big_next = 0;
    sh.survivors = 0; sh.visible = 0; sh.carry = 0; sh.unique = 0; sh.gain_p = 0; sh.gain_v = 0; sh.status = 0;
    sh.steps = 0; sh.vis_hash = 0; sh.surv_hash = 0;
    sh.acc[0] = sh.acc[1] = sh.acc[2] = sh.acc[3] = 0;
    sh.acc_min = 0x7f800000; sh.acc_max = (int)0xff800000 ^ 0x7fffffff;  // float_key(+inf), float_key(-inf)
    sh.words = 0; sh.use_smem = 1;
    sh.g.cells = 0;
  }
}

__device__ __forceinline__ void write_result(const EvalArgs& A, const ViewShared& sh, int v) {
  AsscppViewResult r;
  r.survivors = sh.survivors; r.visible = sh.visible; r.carry = sh.carry; r.unique_rays = sh.unique;
  r.gain_pts = sh.gain_p; r.gain_vox = sh.gain_v; r.status = sh.status; r.reserved = 0;
  r.cells = sh.survivors ? (uint64_t)sh.g.cells : 0;
  r.steps = sh.steps; r.vis_hash = sh.vis_hash; r.surv_hash = sh.surv_hash;
  A.results[v] = r;
  A.gains[v] = make_uint2(sh.gain_p, sh.gain_v);
  if (A.small) {
    A.host_results[v] = r;
    if (!A.view_cov) A.host_gains[v] = make_uint2(sh.gain_p, sh.gain_v);
  }
  if (sh.status & kVsError) atomicAdd(A.error_count, 1);
  if (A.acc) {
    AccRaw a;
    a.d_hi = sh.acc[0]; a.d_lo = sh.acc[1]; a.z_hi = sh.acc[2]; a.z_lo = sh.acc[3];
    a.min_key = sh.acc_min; a.max_key = sh.acc_max;
    a.visible = sh.visible; a.pad = 0;
    A.acc[v] = a;
    if (A.small) A.host_acc[v] = a;
  }
}

// ------------------------------------------------------------------------------------------------
// eval_views_kernel
// ------------------------------------------------------------------------------------------------
extern __shared__ __align__(16) unsigned char smem_raw[];

// Split-group mode inside the small-batch kernel: called by all threads of the CTA when view v is finished.  The last
// view of a group to finish (a ticket that wraps to zero by itself) computes the group's union gains from the per-view
// coverage bitsets: gain_vox of view s = voxels of its visible cloud neither in the covered set nor seen by an earlier
// view of the group (heuristic.cpp:439-445, 473-483), exactly what group_gain_kernel does for larger batches.
__device__ __forceinline__ void finish_group_view(const EvalArgs& A, int v, int* s_last, unsigned* s_gain) {
  const int G = A.split_group;
  const int grp = v / G, v0 = grp * G;
  const int gsz = (int)min((int64_t)G, A.nviews - v0);
  __syncthreads();  // the view's result and bitset are written
  if (threadIdx.x == 0) {
    __threadfence();
    *s_last = atomicInc(A.group_ticket + grp, (unsigned)gsz - 1u) == (unsigned)gsz - 1u;
  }
  if (threadIdx.x < kSmallMaxGroup) s_gain[threadIdx.x] = 0;
  __syncthreads();
  if (!*s_last) return;
  __threadfence();
  // view s of the group: bits of its row that are neither covered nor in an earlier row (rows re-read per view: the
  // groups are small and rolled loops keep this rarely run code short)
#pragma unroll 1
  for (int s = 0; s < gsz; ++s) {
    unsigned n = 0;
    for (int64_t w = threadIdx.x; w < A.cov_words; w += blockDim.x) {
      uint32_t seen = A.covered.vox ? A.covered.vox[w] : 0u;
#pragma unroll 1
      for (int t = 0; t < s; ++t) seen |= __ldcg(A.view_cov + (int64_t)(v0 + t) * A.cov_words + w);
      n += __popc(__ldcg(A.view_cov + (int64_t)(v0 + s) * A.cov_words + w) & ~seen);
    }
    n = warp_sum(n);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(&s_gain[s], n);
  }
  __syncthreads();
  if ((int)threadIdx.x < gsz) {
    const int vs = v0 + (int)threadIdx.x;
    const unsigned gp = __ldcg(&A.gains[vs].x), gv = s_gain[threadIdx.x];
    A.results[vs].gain_vox = gv;
    A.host_results[vs].gain_vox = gv;
    A.gains[vs] = make_uint2(gp, gv);
    A.host_gains[vs] = make_uint2(gp, gv);
  }
}

// param_poses: the poses as kernel parameters (eval_views_small_kernel) or nullptr (planes_kernel has filled A.views)
__device__ __forceinline__ void eval_views_body(const EvalArgs& A, const double* param_poses, const ViewDev* param_views) {
  ViewShared& sh = *reinterpret_cast<ViewShared*>(smem_raw);
  uint32_t* smem_grid = reinterpret_cast<uint32_t*>(smem_raw + ((sizeof(ViewShared) + 15) & ~size_t(15)));
  const ModelDev& m = A.m;
  uint2* chunk_list = A.chunk_list + (int64_t)blockIdx.x * A.chunk_cap;
  uint32_t* vox_list = A.vox_list + (int64_t)blockIdx.x * A.list_cap;
  uint32_t* glob_grid = A.grid_scratch + (int64_t)blockIdx.x * A.grid_cap_words;
  uint32_t* cov_scratch = A.cov_scratch + (int64_t)blockIdx.x * A.cov_words;
  uint32_t* vis_list = A.vis_list + (int64_t)blockIdx.x * A.vis_cap;
  // views are handed out in groups of A.group consecutive views (the sensors of one candidate); within a group
  // the coverage-voxel marks are kept, so gain_vox of a later view counts only voxels no earlier view of the
  // group has seen: the group's gain_vox sum is the gain of the union (heuristic.cpp:439-445, 473-483)
  __shared__ int s_view, s_end, s_first, s_dirty, s_group_last;
  __shared__ unsigned s_group_gain[kSmallMaxGroup];
  if (threadIdx.x == 0) { s_view = -1; s_end = 0; s_dirty = 0; }
  const long long k_start = clock64();
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) {
      int nv = s_view + 1;
      s_first = 0;
      if (nv >= s_end) {
        const long long g = atomicAdd(A.work_counter, 1);
        const long long b = g * A.group;
        nv = (int)(b < A.nviews ? b : A.nviews);
        s_end = (int)(b + A.group < A.nviews ? b + A.group : A.nviews);
        s_first = 1;
      }
      s_view = nv;
    }
    __syncthreads();
    const int v = s_view;
    if (v >= A.nviews) break;
    if (s_first && s_dirty && !A.view_cov) {
      for (int64_t w = threadIdx.x; w < A.cov_words; w += blockDim.x) cov_scratch[w] = 0;
      __syncthreads();
      if (threadIdx.x == 0) s_dirty = 0;
    }
    if (A.view_cov) cov_scratch = A.view_cov + (int64_t)v * A.cov_words;  // split-group mode: marks are per view, zeroed by the host
    long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    const bool prof = A.phase_cycles != nullptr && threadIdx.x == 0;
    if (prof) t0 = clock64();
    if (A.view_cov && A.small)  // the view's coverage bitset starts empty (larger split-group batches: zeroed by the host)
      for (int64_t w = threadIdx.x; w < A.cov_words; w += blockDim.x) cov_scratch[w] = 0;
    reset_view(A, sh, param_poses ? param_poses + 7 * (int64_t)v : nullptr, param_views ? param_views + v : nullptr, v);
    __syncthreads();
    if (prof) atomicAdd(&A.phase_cycles[12], (unsigned long long)(clock64() - t0));
    cull_view(A, sh, chunk_list, vox_list,
              reinterpret_cast<uint32_t*>(A.ray_buf + ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 3 * kRayQueueCap),
              3u * kRayQueueCap * 4u);  // the warp's ray queue is idle during the cull
    if (prof) t1 = clock64();
    if (sh.survivors == 0) {  // empty frustum cloud: extractVisibleSurface returns an empty cloud
      if (threadIdx.x == 0) write_result(A, sh, v);
      if (prof) atomicAdd(&A.phase_cycles[0], (unsigned long long)(t1 - t0));
      if (A.small && A.view_cov) finish_group_view(A, v, &s_group_last, s_group_gain);
      continue;
    }
    GridMem gm;
    const bool in_smem = build_grid(A, sh, smem_grid, glob_grid, vox_list, &gm);
    if (prof) t2 = clock64();
    // P4a: one ray per occupied voxel; every warp prepares batches of rays into its private queue and walks
    // them; voxels whose ray comes back free are collected in vis_list
    {
      uint4* qb = A.ray_buf + ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 3 * kRayQueueCap;
      RayQueue q;
      q.a = qb;
      q.b = reinterpret_cast<float4*>(qb + kRayQueueCap);
      q.c = reinterpret_cast<float4*>(qb + 2 * kRayQueueCap);
      if (in_smem) trace_rays_fast<true>(A, sh, gm, vox_list, q, vis_list);
      else trace_rays_fast<false>(A, sh, gm, vox_list, q, vis_list);
    }
    __syncthreads();
    if (prof) t3 = clock64();
    // P4b: second gate (occlusion_culling.cpp:338-342) per free voxel, then its survivors -> counts, gains, hash.
    // One lane per voxel; voxels with many points are set aside (the front of vox_list is free now) and taken one
    // warp per voxel afterwards.
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    {
      unsigned vis = 0, gain_p = 0, gain_v = 0;
      unsigned long long vhash = 0;
      const unsigned nvis = sh.vis_count;
      const uint32_t coop_min = coop_min_points(nvis);
      const GridView& g = sh.g;
      const float pcd_o[3] = {m.pcd_origin[0], m.pcd_origin[1], m.pcd_origin[2]};
      for (unsigned e0 = threadIdx.x & ~31u; e0 < nvis; e0 += blockDim.x) {
        const unsigned e = e0 + lane;
        bool big = false;
        uint32_t u = 0;
        if (e < nvis) {
          u = vis_list[e];
          const int4 vx = m.vox[u & kVoxMask];
          const float tc[3] = {centre(g, 0, vx.x - g.min_b[0]), centre(g, 1, vx.y - g.min_b[1]),
                               centre(g, 2, vx.z - g.min_b[2])};
          if (!second_gate(g, pcd_o, tc)) {
            vis_list[e] = 0xffffffffu;
          } else {
            big = (uint32_t)m.vox[(u & kVoxMask) + 1].w - (uint32_t)vx.w >= coop_min;
            if (!big) visit_voxel_points(A, sh, u, 0, cov_scratch, vis, gain_p, gain_v, vhash);
          }
        }
        const unsigned bb = __ballot_sync(0xffffffffu, big);
        if (bb) {
          unsigned base = 0;
          if (lane == 0) base = atomicAdd(&sh.big_count, (unsigned)__popc(bb));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (big) __stcg(vox_list + base + __popc(bb & ((1u << lane) - 1)), u);
        }
      }
      if (prof) atomicAdd(&A.phase_cycles[13], (unsigned long long)(clock64() - t3));
      __syncthreads();
      const unsigned nbig = sh.big_count;
      {
        const LaneGroup lg = lane_group(nbig, nwarps);
        for (unsigned i = warp * (32 / lg.W) + lg.sub; i < nbig; i += nwarps * (32 / lg.W))
          visit_voxel_warp<0>(A, sh, lg, __ldcg(vox_list + i), cov_scratch, vis, gain_p, gain_v, vhash, nullptr, nullptr);
        __syncwarp();
      }
      vis = warp_sum(vis); gain_p = warp_sum(gain_p); gain_v = warp_sum(gain_v);
      for (int o = 16; o; o >>= 1) vhash += __shfl_xor_sync(0xffffffffu, vhash, o);
      if (lane == 0) {
        if (vis) atomicAdd(&sh.visible, vis);
        if (gain_p) atomicAdd(&sh.gain_p, gain_p);
        if (gain_v) atomicAdd(&sh.gain_v, gain_v);
        if (vhash) atomicAdd(&sh.vis_hash, vhash);
      }
    }
    __syncthreads();
    if (A.acc && sh.visible && !sh.carry) {
      // P4c: accuracy terms of the visible points (occlusion_culling.cpp:445-475, 505-513)
      const ViewTrig tr = sh.trig;
      AccLocal acc;
      const unsigned nvis = sh.vis_count;
      for (unsigned e = threadIdx.x; e < nvis; e += blockDim.x) {
        const uint32_t uf = vis_list[e];
        if (uf == 0xffffffffu) continue;
        const uint32_t u = uf & kVoxMask;
        const uint32_t pb = (uint32_t)m.vox[u].w, pe = (uint32_t)m.vox[u + 1].w;
        if (pe - pb >= coop_min_points(nvis)) continue;  // in the list of P4b
        for (uint32_t pos = pb; pos < pe; ++pos) {
          const float4 p = ld_pt(&m.pts[pos]);
          if (!(uf & kVoxInside) && !point_in_frustum(sh.planes, p.x, p.y, p.z)) continue;
          acc_point(tr, A.acc_scale_d, A.acc_scale_z, p.x, p.y, p.z, acc);
        }
      }
      {
        unsigned d0 = 0, d1 = 0, d2 = 0;
        unsigned long long d3 = 0;
        const unsigned nbig = sh.big_count;
        const LaneGroup lg = lane_group(nbig, nwarps);
        for (unsigned i = warp * (32 / lg.W) + lg.sub; i < nbig; i += nwarps * (32 / lg.W))
          visit_voxel_warp<2>(A, sh, lg, __ldcg(vox_list + i), cov_scratch, d0, d1, d2, d3, &tr, &acc);
        __syncwarp();
      }
      acc_commit(sh, acc);
      __syncthreads();
    }
    if (!A.cov_direct && !A.view_cov && sh.visible && threadIdx.x == 0) s_dirty = 1;  // marks are cleared before the next group
    if (sh.carry) {
      // quirk Q5: at least one ray failed; the sticky `state` makes the result order dependent.
      // Hand the view to ordered_views_kernel, which replays the reference loop exactly.
      if (threadIdx.x == 0) {
        int at = atomicAdd(A.ordered_count, 1);
        A.ordered_list[at] = v;
        sh.status |= ASSCPP_VS_ORDERED;
        write_result(A, sh, v);
      }
      if (A.small && A.view_cov) finish_group_view(A, v, &s_group_last, s_group_gain);
      continue;
    }
    // P5: OR the visible set into the accumulator (heuristic.cpp:344-346)
    if (A.accum.pts && (int64_t)sh.visible > A.accum_min_visible) {
      unsigned d0 = 0, d1 = 0, d2 = 0;
      unsigned long long d3 = 0;
      const unsigned nvis = sh.vis_count;
      for (unsigned e = threadIdx.x; e < nvis; e += blockDim.x) {
        const uint32_t uf = vis_list[e];
        if (uf == 0xffffffffu) continue;
        if ((uint32_t)m.vox[(uf & kVoxMask) + 1].w - (uint32_t)m.vox[uf & kVoxMask].w >= coop_min_points(nvis)) continue;
        visit_voxel_points(A, sh, uf, 1, cov_scratch, d0, d1, d2, d3);
      }
      const unsigned nbig = sh.big_count;
      const LaneGroup lg = lane_group(nbig, nwarps);
      for (unsigned i = warp * (32 / lg.W) + lg.sub; i < nbig; i += nwarps * (32 / lg.W))
        visit_voxel_warp<1>(A, sh, lg, __ldcg(vox_list + i), cov_scratch, d0, d1, d2, d3, nullptr, nullptr);
      __syncwarp();
    }
    if (threadIdx.x == 0) write_result(A, sh, v);
    const long long tg = prof ? clock64() : 0;
    if (A.small && A.view_cov) finish_group_view(A, v, &s_group_last, s_group_gain);
    if (prof) {
      long long t4 = clock64();
      atomicAdd(&A.phase_cycles[15], (unsigned long long)(t4 - tg));
      t4 = tg;
      atomicAdd(&A.phase_cycles[0], (unsigned long long)(t1 - t0));  // cull (P1 + P2)
      atomicAdd(&A.phase_cycles[1], (unsigned long long)(t2 - t1));  // grid build (P3)
      atomicAdd(&A.phase_cycles[2], (unsigned long long)(t3 - t2));  // DDA (P4a)
      atomicAdd(&A.phase_cycles[3], (unsigned long long)(t4 - t3));  // point visit + OR (P4b, P5)
      atomicAdd(&A.phase_cycles[4], 1ull);
    }
  }
  if (s_dirty && !A.view_cov)  // ordered_views_kernel expects clean marks
    for (int64_t w = threadIdx.x; w < A.cov_words; w += blockDim.x) (A.cov_scratch + (int64_t)blockIdx.x * A.cov_words)[w] = 0;
  if (A.phase_cycles != nullptr && threadIdx.x == 0) atomicAdd(&A.phase_cycles[5], (unsigned long long)clock64() - (unsigned long long)k_start);
  if (A.small) {
    // the last CTA to leave publishes the counters the host looks at and resets the device counters, so the next
    // small launch needs no memset: {work counter, ordered views, error views, CTAs done}
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();  // this CTA's results are in host memory before it counts as done
      if (atomicAdd(A.work_counter + 3, 1) == (int)gridDim.x - 1) {
        A.host_counters[0] = *(volatile int32_t*)A.ordered_count;
        A.host_counters[1] = *(volatile int32_t*)A.error_count;
        A.work_counter[0] = 0; A.work_counter[1] = 0; A.work_counter[2] = 0; A.work_counter[3] = 0;
        __threadfence_system();
        *(volatile int32_t*)(A.host_counters + 2) = A.epoch;  // the host may spin on this instead of synchronising
      }
    }
  }
}

__global__ void __launch_bounds__(1024, 1) eval_views_kernel(const EvalArgs A) { eval_views_body(A, nullptr, nullptr); }

// Small batches: the poses travel with the launch (7 doubles per view) instead of through a host-to-device copy and a
// planes_kernel launch.
template <int NVIEWS>
struct PoseBlock {
  double p[7 * NVIEWS];
};
template <int NVIEWS>
__global__ void __launch_bounds__(1024, 1) eval_views_small_kernel(const __grid_constant__ EvalArgs A,
                                                                   const __grid_constant__ PoseBlock<NVIEWS> P) {
  eval_views_body(A, P.p, nullptr);
}
constexpr int kTinyMaxViews = 8;  // at most this many views: the host computes the planes (under a microsecond)
struct PlaneBlock {
  ViewDev v[kTinyMaxViews];
};
__global__ void __launch_bounds__(1024, 1) eval_views_tiny_kernel(const __grid_constant__ EvalArgs A,
                                                                  const __grid_constant__ PlaneBlock P) {
  eval_views_body(A, nullptr, P.v);
}


// ------------------------------------------------------------------------------------------------
// ordered_views_kernel: exact replay of occlusion_culling.cpp:308-370 in ORIGINAL index order for
// the views in which occlusionEstimation failed.  Slow by design (two passes over all N points,
// one CTA per view); such views need tmin == -1.0f exactly and are rare.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1) ordered_views_kernel(const EvalArgs A) {
  ViewShared& sh = *reinterpret_cast<ViewShared*>(smem_raw);
  uint32_t* smem_grid = reinterpret_cast<uint32_t*>(smem_raw + ((sizeof(ViewShared) + 15) & ~size_t(15)));
  const ModelDev& m = A.m;
  uint2* chunk_list = A.chunk_list + (int64_t)blockIdx.x * A.chunk_cap;
  uint32_t* vox_list = A.vox_list + (int64_t)blockIdx.x * A.list_cap;
  uint32_t* glob_grid = A.grid_scratch + (int64_t)blockIdx.x * A.grid_cap_words;
  uint32_t* cov_scratch = A.cov_scratch + (int64_t)blockIdx.x * A.cov_words;
  uint8_t* vstate = A.vstate + (int64_t)blockIdx.x * m.n_vox;
  const int nflag = *A.ordered_count;
  const bool want_hash = (A.flags & ASSCPP_EVAL_HASH) != 0;
  for (int q = blockIdx.x; q < nflag; q += gridDim.x) {
    const int v = A.ordered_list[q];
    __syncthreads();
    if (A.view_cov) {  // split-group mode: this view's own bitset, which the fast path has already written into
      cov_scratch = A.view_cov + (int64_t)v * A.cov_words;
      for (int64_t w = threadIdx.x; w < A.cov_words; w += blockDim.x) cov_scratch[w] = 0;
    }
    reset_view(A, sh, nullptr, nullptr, v);
    __syncthreads();
    cull_view(A, sh, chunk_list, vox_list,
              reinterpret_cast<uint32_t*>(A.ray_buf + ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 3 * kRayQueueCap),
              3u * kRayQueueCap * 4u);  // the warp's ray queue is idle during the cull
    GridMem gm;
    const bool in_smem = build_grid(A, sh, smem_grid, glob_grid, vox_list, &gm);
    // per-voxel state: bits 0-1 = 0 free / 1 occluded / 2 failed, bit 2 = second gate passes
    if (in_smem) trace_rays<true>(A, sh, gm, vox_list, nullptr, vstate);
    else trace_rays<false>(A, sh, gm, vox_list, nullptr, vstate);
    __syncthreads();
    // ordered scan over original indices, one contiguous segment per thread
    const int64_t seg = (m.n + blockDim.x - 1) / blockDim.x;
    const int64_t j0 = (int64_t)threadIdx.x * seg, j1 = min(m.n, j0 + seg);
    int last = -1;
    for (int64_t j = j0; j < j1; ++j) {
      uint32_t pos = (uint32_t)m.pos_of[j];
      float4 p = m.pts[pos];
      uint8_t s = vstate[__float_as_uint(p.w)];
      if (s == 0xff) continue;
      if (!point_in_frustum(sh.planes, p.x, p.y, p.z)) continue;
      if ((s & 3) != 2) last = s & 3;
    }
    sh.seg_last[threadIdx.x] = last;
    __syncthreads();
    int state = 0;  // canonical initial value of the uninitialised `state` (occlusion_culling.cpp:308)
    for (int t = (int)threadIdx.x - 1; t >= 0; --t)
      if (sh.seg_last[t] >= 0) { state = sh.seg_last[t]; break; }
    for (int pass = 0; pass < 2; ++pass) {
      // pass 0 counts; pass 1 ORs into the accumulator once the visible count is known
      if (pass == 1 && !(A.accum.pts && (int64_t)sh.visible > A.accum_min_visible)) break;
      int st = state;
      unsigned vis = 0, gain_p = 0, gain_v = 0, carry = 0;
      unsigned long long vhash = 0;
      AccLocal acc;
      ViewTrig tr;
      if (A.acc && pass == 0) tr = A.trig[v];
      for (int64_t j = j0; j < j1; ++j) {
        uint32_t pos = (uint32_t)m.pos_of[j];
        float4 p = m.pts[pos];
        uint8_t s = vstate[__float_as_uint(p.w)];
        if (s == 0xff) continue;
        if (!point_in_frustum(sh.planes, p.x, p.y, p.z)) continue;
        if ((s & 3) != 2) st = s & 3; else ++carry;       // :332 ret = -1 leaves state untouched
        if (st != 1 && (s & 4)) {                         // :335, :342
          uint32_t id = (uint32_t)m.pt_cov[pos], bit = 1u << (id & 31);
          if (pass == 0) {
            ++vis;
            if (!A.covered.pts || !bit_test(A.covered.pts, pos)) ++gain_p;
            if (want_hash) vhash += mix64((uint64_t)j);
            if (A.acc) acc_point(tr, A.acc_scale_d, A.acc_scale_z, p.x, p.y, p.z, acc);
            uint32_t old = atomicOr(&cov_scratch[id >> 5], bit);
            if (!(old & bit) && (!A.covered.vox || !bit_test(A.covered.vox, id))) ++gain_v;
          } else {
            atomicOr(&A.accum.pts[pos >> 5], 1u << (pos & 31));
            atomicOr(&A.accum.vox[id >> 5], bit);
          }
        }
      }
      if (pass == 0) {
        if (A.acc) acc_commit(sh, acc);
        if (vis) atomicAdd(&sh.visible, vis);
        if (gain_p) atomicAdd(&sh.gain_p, gain_p);
        if (gain_v) atomicAdd(&sh.gain_v, gain_v);
        if (carry) atomicAdd(&sh.carry, carry);
        if (vhash) atomicAdd(&sh.vis_hash, vhash);
      }
      __syncthreads();
    }
    if (!A.view_cov)
      for (int64_t w = threadIdx.x; w < A.cov_words; w += blockDim.x) cov_scratch[w] = 0;
    for (unsigned e = threadIdx.x; e < sh.list_count; e += blockDim.x) {
      uint32_t u = vox_list[e];
      if (u != 0xffffffffu) vstate[u & kVoxMask] = 0xff;
    }
    if (threadIdx.x == 0) {
      sh.status |= ASSCPP_VS_ORDERED;
      write_result(A, sh, v);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// group_gain_kernel (split-group mode): one warp per group.  gain_vox of view s of a group = coverage voxels of its
// visible cloud that are neither in the covered set nor seen by an earlier view of the group
// (heuristic.cpp:439-445, 473-483: the group's sum is the gain of the union of the sensors' clouds).
// ------------------------------------------------------------------------------------------------
__global__ void group_gain_kernel(uint32_t* __restrict__ view_cov, int64_t cov_words, const uint32_t* __restrict__ covered,
                                  int64_t nviews, int group, AsscppViewResult* __restrict__ results,
                                  uint2* __restrict__ gains) {
  const int lane = threadIdx.x & 31;
  const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t v0 = g * group;
  if (v0 >= nviews) return;
  const int64_t v1 = min(nviews, v0 + group);
  if (group <= kSmallMaxGroup) {
    // small groups: the earlier rows are re-read per view and every row stays what its view saw (asscpp_get_view_cells)
    for (int64_t v = v0; v < v1; ++v) {
      unsigned n = 0;
      for (int64_t w = lane; w < cov_words; w += 32) {
        uint32_t seen = covered ? covered[w] : 0u;
        for (int64_t t = v0; t < v; ++t) seen |= view_cov[t * cov_words + w];
        n += __popc(view_cov[v * cov_words + w] & ~seen);
      }
      n = warp_sum(n);
      if (lane == 0) { results[v].gain_vox = n; gains[v].y = n; }
    }
    return;
  }
  uint32_t* acc = view_cov + v0 * cov_words;  // the first view's bitset becomes the running union
  for (int64_t v = v0; v < v1; ++v) {
    const uint32_t* mine = view_cov + v * cov_words;
    unsigned n = 0;
    for (int64_t w = lane; w < cov_words; w += 32) {
      const uint32_t m = mine[w], c = covered ? covered[w] : 0u, a = v == v0 ? 0u : acc[w];
      n += __popc(m & ~c & ~a);
      if (v != v0) acc[w] = a | m;
    }
    n = warp_sum(n);
    if (lane == 0) { results[v].gain_vox = n; gains[v].y = n; }
  }
}

// ------------------------------------------------------------------------------------------------
// covered-set kernels
// ------------------------------------------------------------------------------------------------
__global__ void or_words_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, int64_t words) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] |= src[i];
}

// dst |= src[0] | src[1] | ... (nsrc bitsets laid out back to back): the local half of the OR-allreduce
__global__ void or_reduce_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, int64_t words,
                                 int nsrc) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (int64_t)gridDim.x * blockDim.x) {
    uint32_t v = dst[i];
    for (int r = 0; r < nsrc; ++r) v |= src[(int64_t)r * words + i];
    dst[i] = v;
  }
}

__global__ void popcount_kernel(const uint32_t* __restrict__ bits, int64_t words, unsigned long long* out) {
  unsigned long long c = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (int64_t)gridDim.x * blockDim.x)
    c += __popc(bits[i]);
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// point bitset in sorted-position space <-> original index space
__global__ void permute_bits_kernel(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst,
                                    const int32_t* __restrict__ map, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    if ((src[i >> 5] >> (i & 31)) & 1u) {
      uint32_t j = (uint32_t)map[i];
      atomicOr(&dst[j >> 5], 1u << (j & 31));
    }
}

// set bits of the listed ORIGINAL indices (and their coverage voxels)
__global__ void or_points_kernel(CoveredDev set, const int32_t* __restrict__ idx, int64_t n, ModelDev m) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    int32_t j = idx[i];
    if (j < 0 || j >= m.n) continue;
    uint32_t pos = (uint32_t)m.pos_of[j];
    atomicOr(&set.pts[pos >> 5], 1u << (pos & 31));
    uint32_t id = (uint32_t)m.pt_cov[pos];
    atomicOr(&set.vox[id >> 5], 1u << (id & 31));
  }
}

// calcCoveragePercent of an arbitrary cloud (occlusion_culling.cpp:387-428): voxelise at the coverage
// leaf and mark the model coverage voxels hit.  keys = sorted linear keys of the model's coverage voxels.
__global__ void cloud_cover_kernel(const float* __restrict__ xyz, int64_t n, float inv_cov, int3 cmin, int3 cdiv,
                                   const long long* __restrict__ keys, int64_t nkeys, uint32_t* __restrict__ marks) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    float qx = fmul(x, inv_cov), qy = fmul(y, inv_cov), qz = fmul(z, inv_cov);
    if (!(fabsf(qx) < 1e9f) || !(fabsf(qy) < 1e9f) || !(fabsf(qz) < 1e9f)) continue;
    long long a = (long long)((int)floorf(qx)) - cmin.x, b = (long long)((int)floorf(qy)) - cmin.y,
              c = (long long)((int)floorf(qz)) - cmin.z;
    if (a < 0 || b < 0 || c < 0 || a >= cdiv.x || b >= cdiv.y || c >= cdiv.z) continue;
    long long key = a + (long long)cdiv.x * (b + (long long)cdiv.y * c);
    int64_t lo = 0, hi = nkeys;
    while (lo < hi) {
      int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < key) lo = mid + 1; else hi = mid;
    }
    if (lo < nkeys && keys[lo] == key) atomicOr(&marks[lo >> 5], 1u << (lo & 31));
  }
}

// flip the first nbits bits (the model's coverage voxels); the padding bits of the last word stay zero
__global__ void invert_bits_kernel(uint32_t* __restrict__ bits, int64_t nbits) {
  const int64_t words = (nbits + 31) / 32;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    uint32_t mask = 0xffffffffu;
    if (w == words - 1 && (nbits & 31)) mask = (1u << (nbits & 31)) - 1u;
    bits[w] = (~bits[w]) & mask;
  }
}

__global__ void fill_bytes_kernel(uint8_t* p, int64_t n, uint8_t v) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// ------------------------------------------------------------------------------------------------
// host-side launchers
// ------------------------------------------------------------------------------------------------
size_t eval_smem_bytes(uint32_t grid_words) { return ((sizeof(ViewShared) + 15) & ~size_t(15)) + (size_t)grid_words * 4; }
int ray_queue_cap() { return kRayQueueCap; }
int tiny_max_views() { return kTinyMaxViews; }
size_t eval_smem_header_bytes() { return ((sizeof(ViewShared) + 15) & ~size_t(15)); }

cudaError_t configure_kernels(size_t smem_bytes) {
  cudaError_t e = cudaFuncSetAttribute(eval_views_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(eval_views_tiny_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(eval_views_small_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(eval_views_small_kernel<kSmallMaxViews>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(ordered_views_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
}

void launch_planes(const double* d_poses, int64_t nviews, const FrustumConst& fc, ViewDev* views, cudaStream_t s) {
  if (nviews <= 0) return;
  int threads = 128;
  int blocks = (int)((nviews + threads - 1) / threads);
  planes_kernel<<<blocks, threads, 0, s>>>(d_poses, nviews, fc, views);
}
void launch_trig(const double* d_poses, int64_t nviews, ViewTrig* out, cudaStream_t s) {
  if (nviews <= 0) return;
  trig_kernel<<<(int)((nviews + 127) / 128), 128, 0, s>>>(d_poses, nviews, out);
}
void launch_eval(const EvalArgs& a, int ctas, int threads, size_t smem_bytes, cudaStream_t s) {
  eval_views_kernel<<<ctas, threads, smem_bytes, s>>>(a);
}
void launch_eval_small(const EvalArgs& a, const double* h_poses, const ViewDev* h_views, int ctas, int threads,
                       size_t smem_bytes, cudaStream_t s) {
  // the launch copies the parameter block: the poses need not outlive this call
  if (h_views && a.nviews <= kTinyMaxViews) {
    PlaneBlock pb;
    memcpy(pb.v, h_views, (size_t)a.nviews * sizeof(ViewDev));
    eval_views_tiny_kernel<<<ctas, threads, smem_bytes, s>>>(a, pb);
  } else if (a.nviews <= 64) {
    PoseBlock<64> pb;
    memcpy(pb.p, h_poses, (size_t)a.nviews * 7 * sizeof(double));
    eval_views_small_kernel<64><<<ctas, threads, smem_bytes, s>>>(a, pb);
  } else {
    PoseBlock<kSmallMaxViews> pb;
    memcpy(pb.p, h_poses, (size_t)a.nviews * 7 * sizeof(double));
    eval_views_small_kernel<kSmallMaxViews><<<ctas, threads, smem_bytes, s>>>(a, pb);
  }
}
void launch_group_gain(const EvalArgs& a, int group, cudaStream_t s) {
  const int64_t ngroups = (a.nviews + group - 1) / group;
  if (ngroups <= 0) return;
  const int threads = 128;
  group_gain_kernel<<<(int)((ngroups * 32 + threads - 1) / threads), threads, 0, s>>>(a.view_cov, a.cov_words, a.covered.vox,
                                                                                    a.nviews, group, a.results, a.gains);
}
void launch_ordered(const EvalArgs& a, int ctas, size_t smem_bytes, cudaStream_t s) {
  ordered_views_kernel<<<ctas, 1024, smem_bytes, s>>>(a);
}
static inline int blocks_for(int64_t n, int threads, int cap) {
  int64_t b = (n + threads - 1) / threads;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
void launch_or_words(uint32_t* dst, const uint32_t* src, int64_t words, cudaStream_t s) {
  if (words > 0) or_words_kernel<<<blocks_for(words, 256, 1184), 256, 0, s>>>(dst, src, words);
}
void launch_or_reduce(uint32_t* dst, const uint32_t* src, int64_t words, int nsrc, cudaStream_t s) {
  if (words > 0) or_reduce_kernel<<<blocks_for(words, 256, 1184), 256, 0, s>>>(dst, src, words, nsrc);
}
void launch_popcount(const uint32_t* bits, int64_t words, unsigned long long* out, cudaStream_t s) {
  if (words > 0) popcount_kernel<<<blocks_for(words, 256, 1184), 256, 0, s>>>(bits, words, out);
}
void launch_permute_bits(const uint32_t* src, uint32_t* dst, const int32_t* map, int64_t n, cudaStream_t s) {
  if (n > 0) permute_bits_kernel<<<blocks_for(n, 256, 2368), 256, 0, s>>>(src, dst, map, n);
}
void launch_or_points(CoveredDev set, const int32_t* idx, int64_t n, const ModelDev& m, cudaStream_t s) {
  if (n > 0) or_points_kernel<<<blocks_for(n, 256, 1184), 256, 0, s>>>(set, idx, n, m);
}
void launch_cloud_cover(const float* xyz, int64_t n, float inv_cov, const int32_t cmin[3], const int32_t cdiv[3],
                        const int64_t* keys, int64_t nkeys, uint32_t* marks, cudaStream_t s) {
  if (n > 0)
    cloud_cover_kernel<<<blocks_for(n, 256, 1184), 256, 0, s>>>(xyz, n, inv_cov, make_int3(cmin[0], cmin[1], cmin[2]),
                                                                make_int3(cdiv[0], cdiv[1], cdiv[2]),
                                                                (const long long*)keys, nkeys, marks);
}
void launch_invert_bits(uint32_t* bits, int64_t nbits, cudaStream_t s) {
  if (nbits > 0) invert_bits_kernel<<<blocks_for((nbits + 31) / 32, 256, 1184), 256, 0, s>>>(bits, nbits);
}
void launch_fill_bytes(uint8_t* p, int64_t n, uint8_t v, cudaStream_t s) {
  if (n > 0) fill_bytes_kernel<<<blocks_for(n, 256, 1184), 256, 0, s>>>(p, n, v);
}

}  // namespace asscpp
