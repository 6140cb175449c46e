// mesh.cu -- the geometric sample filters of the adaptive loop against the structure's triangle mesh
// (SURVEY.md 8f rank 3).  The reference answers them with a CGAL AABB tree over Simple_cartesian<double>
// triangles (cscpp/include/cscpp/coverage_path_planning_heuristic.h:59-69):
//   isFilteringConditionSatisfied  coverage_path_planning_heuristic.cpp:317-335
//       inside test : number_of_intersected_primitives(Ray(a, b = (100, 10, 56))) odd  -> reject
//       stand-off   : minDist^2 <= squared_distance(a) <= maxDist^2
//   isConnectionConditionSatisfied :300-315   number_of_intersected_primitives(Segment(a, b)) == 0
// Here every query visits every triangle (32 k triangles x a few thousand candidates is milliseconds on a B200):
// one CTA per query, triangles strided over its threads, double precision, no FMA contraction, no tree.
// Intersection = division-free Moeller-Trumbore with inclusive edges; distance = closest point on the triangle
// by Voronoi regions.  Both differ from CGAL's predicates only for queries that graze an edge or a vertex.
#include <cuda_runtime.h>

#include <cmath>
#include <string>
#include <vector>

#include "../../include/asscpp_gpu.h"

namespace {

struct Tri { double p[9]; };  // p0 p1 p2

__device__ __forceinline__ double dot3d(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

// ray (tmax < 0) or segment (t in [0, 1]) o + t d against one triangle
__device__ __forceinline__ bool hit(const double* o, const double* d, const Tri& t, bool segment) {
  const double e1[3] = {t.p[3] - t.p[0], t.p[4] - t.p[1], t.p[5] - t.p[2]};
  const double e2[3] = {t.p[6] - t.p[0], t.p[7] - t.p[1], t.p[8] - t.p[2]};
  const double h[3] = {d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0]};
  double det = dot3d(e1, h);
  if (det == 0.0) return false;  // parallel to the triangle's plane
  const double s[3] = {o[0] - t.p[0], o[1] - t.p[1], o[2] - t.p[2]};
  double u = dot3d(s, h);
  const double q[3] = {s[1] * e1[2] - s[2] * e1[1], s[2] * e1[0] - s[0] * e1[2], s[0] * e1[1] - s[1] * e1[0]};
  double v = dot3d(d, q), w = dot3d(e2, q);
  if (det < 0.0) { det = -det; u = -u; v = -v; w = -w; }
  if (u < 0.0 || v < 0.0 || u + v > det || w < 0.0) return false;
  return !segment || w <= det;
}

// squared distance from p to the triangle (closest point by Voronoi region of the triangle)
__device__ __forceinline__ double sq_dist(const double* p, const Tri& t) {
  const double* a = t.p;
  const double* b = t.p + 3;
  const double* c = t.p + 6;
  const double ab[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, ac[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
  const double ap[3] = {p[0] - a[0], p[1] - a[1], p[2] - a[2]};
  const double d1 = dot3d(ab, ap), d2 = dot3d(ac, ap);
  double cl[3];
  bool done = false;
  if (d1 <= 0.0 && d2 <= 0.0) { cl[0] = a[0]; cl[1] = a[1]; cl[2] = a[2]; done = true; }
  double d3 = 0, d4 = 0, d5 = 0, d6 = 0;
  if (!done) {
    const double bp[3] = {p[0] - b[0], p[1] - b[1], p[2] - b[2]};
    d3 = dot3d(ab, bp); d4 = dot3d(ac, bp);
    if (d3 >= 0.0 && d4 <= d3) { cl[0] = b[0]; cl[1] = b[1]; cl[2] = b[2]; done = true; }
  }
  if (!done) {
    const double vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
      const double v = d1 / (d1 - d3);
      cl[0] = a[0] + v * ab[0]; cl[1] = a[1] + v * ab[1]; cl[2] = a[2] + v * ab[2];
      done = true;
    }
    if (!done) {
      const double cp[3] = {p[0] - c[0], p[1] - c[1], p[2] - c[2]};
      d5 = dot3d(ab, cp); d6 = dot3d(ac, cp);
      if (d6 >= 0.0 && d5 <= d6) { cl[0] = c[0]; cl[1] = c[1]; cl[2] = c[2]; done = true; }
    }
    if (!done) {
      const double vb = d5 * d2 - d1 * d6;
      if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
        const double w = d2 / (d2 - d6);
        cl[0] = a[0] + w * ac[0]; cl[1] = a[1] + w * ac[1]; cl[2] = a[2] + w * ac[2];
        done = true;
      }
      if (!done) {
        const double va = d3 * d6 - d5 * d4;
        if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
          const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
          cl[0] = b[0] + w * (c[0] - b[0]); cl[1] = b[1] + w * (c[1] - b[1]); cl[2] = b[2] + w * (c[2] - b[2]);
        } else {
          const double denom = 1.0 / (va + vb + vc);
          const double v = vb * denom, w = vc * denom;
          cl[0] = a[0] + ab[0] * v + ac[0] * w; cl[1] = a[1] + ab[1] * v + ac[1] * w; cl[2] = a[2] + ab[2] * v + ac[2] * w;
        }
      }
    }
  }
  const double e[3] = {p[0] - cl[0], p[1] - cl[1], p[2] - cl[2]};
  return dot3d(e, e);
}

// one CTA per query: hits of the ray / segment a -> b and the smallest squared distance from a
__global__ void __launch_bounds__(256) mesh_query_kernel(const Tri* __restrict__ tris, int64_t nt,
                                                          const double* __restrict__ a_pts, const double* __restrict__ b_pts,
                                                          int b_stride, int segment, int want_dist,
                                                          int32_t* __restrict__ hits, double* __restrict__ sqd) {
  const int64_t qi = blockIdx.x;
  const double o[3] = {a_pts[3 * qi], a_pts[3 * qi + 1], a_pts[3 * qi + 2]};
  const double* bp = b_pts + (int64_t)b_stride * qi;
  const double d[3] = {bp[0] - o[0], bp[1] - o[1], bp[2] - o[2]};
  int cnt = 0;
  double best = INFINITY;
  for (int64_t k = threadIdx.x; k < nt; k += blockDim.x) {
    const Tri t = tris[k];
    cnt += hit(o, d, t, segment != 0) ? 1 : 0;
    if (want_dist) best = fmin(best, sq_dist(o, t));
  }
  for (int s = 16; s; s >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
    best = fmin(best, __shfl_xor_sync(0xffffffffu, best, s));
  }
  __shared__ int s_cnt[8];
  __shared__ double s_best[8];
  if ((threadIdx.x & 31) == 0) { s_cnt[threadIdx.x >> 5] = cnt; s_best[threadIdx.x >> 5] = best; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { cnt += s_cnt[w]; best = fmin(best, s_best[w]); }
    hits[qi] = cnt;
    if (sqd) sqd[qi] = best;
  }
}

}  // namespace

struct AsscppMesh {
  int device = 0;
  int64_t nt = 0;
  Tri* d_tris = nullptr;
  cudaStream_t stream = nullptr;
  int64_t cap = 0;
  double *d_a = nullptr, *d_b = nullptr, *d_sqd = nullptr;
  int32_t* d_hits = nullptr;
  int64_t launches = 0;
  std::string err;
};

namespace {
std::string g_mesh_error;

int mesh_reserve(AsscppMesh* m, int64_t n) {
  if (n <= m->cap) return ASSCPP_OK;
  cudaFree(m->d_a); cudaFree(m->d_b); cudaFree(m->d_sqd); cudaFree(m->d_hits);
  m->d_a = m->d_b = m->d_sqd = nullptr; m->d_hits = nullptr; m->cap = 0;
  if (cudaMalloc(&m->d_a, (size_t)n * 24) != cudaSuccess || cudaMalloc(&m->d_b, (size_t)n * 24) != cudaSuccess ||
      cudaMalloc(&m->d_sqd, (size_t)n * 8) != cudaSuccess || cudaMalloc(&m->d_hits, (size_t)n * 4) != cudaSuccess) {
    m->err = "cudaMalloc failed for mesh query buffers";
    return ASSCPP_ENOMEM;
  }
  m->cap = n;
  return ASSCPP_OK;
}
}  // namespace

extern "C" {

const char* asscpp_mesh_last_error(const AsscppMesh* m) { return m ? m->err.c_str() : g_mesh_error.c_str(); }

int asscpp_mesh_create(int device, const float* vertices, int64_t nv, const int32_t* triangles, int64_t nt,
                       AsscppMesh** out) {
  if (!out) return ASSCPP_EINVAL;
  *out = nullptr;
  if ((!vertices && nv > 0) || (!triangles && nt > 0) || nv < 0 || nt < 0) { g_mesh_error = "invalid arguments"; return ASSCPP_EINVAL; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    g_mesh_error = "no usable CUDA device (this library has no CPU path)";
    return ASSCPP_ENODEVICE;
  }
  std::vector<Tri> h;
  h.reserve((size_t)nt);
  for (int64_t k = 0; k < nt; ++k) {
    Tri t;
    for (int c = 0; c < 3; ++c) {
      const int32_t vi = triangles[3 * k + c];
      if (vi < 0 || vi >= nv) { g_mesh_error = "triangle refers to a vertex out of range"; return ASSCPP_EINVAL; }
      for (int a = 0; a < 3; ++a) t.p[3 * c + a] = (double)vertices[3 * (int64_t)vi + a];  // float -> Point (double), :1405-1407
    }
    // loadOBJFile drops collinear triangles (CGAL::collinear, coverage_path_planning_heuristic.cpp:1409, 1430):
    // (p1 - p0) x (p2 - p0) == 0 in double
    const double u[3] = {t.p[3] - t.p[0], t.p[4] - t.p[1], t.p[5] - t.p[2]}, v[3] = {t.p[6] - t.p[0], t.p[7] - t.p[1], t.p[8] - t.p[2]};
    const double cx = u[1] * v[2] - u[2] * v[1], cy = u[2] * v[0] - u[0] * v[2], cz = u[0] * v[1] - u[1] * v[0];
    if (cx == 0.0 && cy == 0.0 && cz == 0.0) continue;
    h.push_back(t);
  }
  AsscppMesh* m = new AsscppMesh();
  m->device = device;
  m->nt = (int64_t)h.size();
  cudaSetDevice(device);
  if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaMalloc(&m->d_tris, std::max<size_t>(h.size(), 1) * sizeof(Tri)) != cudaSuccess) {
    g_mesh_error = "CUDA allocation failed";
    delete m;
    return ASSCPP_ECUDA;
  }
  if (!h.empty()) cudaMemcpy(m->d_tris, h.data(), h.size() * sizeof(Tri), cudaMemcpyHostToDevice);
  *out = m;
  return ASSCPP_OK;
}

int asscpp_mesh_destroy(AsscppMesh* m) {
  if (!m) return ASSCPP_OK;
  cudaSetDevice(m->device);
  cudaFree(m->d_tris); cudaFree(m->d_a); cudaFree(m->d_b); cudaFree(m->d_sqd); cudaFree(m->d_hits);
  if (m->stream) cudaStreamDestroy(m->stream);
  delete m;
  return ASSCPP_OK;
}

int64_t asscpp_mesh_num_triangles(const AsscppMesh* m) { return m ? m->nt : -1; }
int64_t asscpp_mesh_launch_count(const AsscppMesh* m) { return m ? m->launches : -1; }

int asscpp_mesh_filter(AsscppMesh* m, const double* positions, int64_t n, const double ray_through[3], double min_dist,
                       double max_dist, int32_t* n_hits, double* sq_dist_out, uint8_t* pass) {
  if (!m || (!positions && n > 0) || !ray_through || n < 0) return ASSCPP_EINVAL;
  if (n == 0) return ASSCPP_OK;
  cudaSetDevice(m->device);
  int rc = mesh_reserve(m, n);
  if (rc) return rc;
  cudaMemcpyAsync(m->d_a, positions, (size_t)n * 24, cudaMemcpyHostToDevice, m->stream);
  cudaMemcpyAsync(m->d_b, ray_through, 24, cudaMemcpyHostToDevice, m->stream);
  mesh_query_kernel<<<(unsigned)n, 256, 0, m->stream>>>(m->d_tris, m->nt, m->d_a, m->d_b, 0, 0, 1, m->d_hits, m->d_sqd);
  m->launches += 1;
  std::vector<int32_t> hh((size_t)n);
  std::vector<double> hd((size_t)n);
  cudaMemcpyAsync(hh.data(), m->d_hits, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream);
  cudaMemcpyAsync(hd.data(), m->d_sqd, (size_t)n * 8, cudaMemcpyDeviceToHost, m->stream);
  cudaError_t e = cudaStreamSynchronize(m->stream);
  if (e != cudaSuccess) { m->err = cudaGetErrorString(e); return ASSCPP_ECUDA; }
  for (int64_t i = 0; i < n; ++i) {
    if (n_hits) n_hits[i] = hh[(size_t)i];
    if (sq_dist_out) sq_dist_out[i] = hd[(size_t)i];
    // heuristic.cpp:328-333: outside the model (even number of crossings) and within the stand-off band
    if (pass)
      pass[i] = (hh[(size_t)i] % 2 != 1) && hd[(size_t)i] >= (min_dist * min_dist) && hd[(size_t)i] <= (max_dist * max_dist);
  }
  return ASSCPP_OK;
}

int asscpp_mesh_segments_free(AsscppMesh* m, const double* a, const double* b, int64_t n, uint8_t* free_out) {
  if (!m || ((!a || !b || !free_out) && n > 0) || n < 0) return ASSCPP_EINVAL;
  if (n == 0) return ASSCPP_OK;
  cudaSetDevice(m->device);
  int rc = mesh_reserve(m, n);
  if (rc) return rc;
  cudaMemcpyAsync(m->d_a, a, (size_t)n * 24, cudaMemcpyHostToDevice, m->stream);
  cudaMemcpyAsync(m->d_b, b, (size_t)n * 24, cudaMemcpyHostToDevice, m->stream);
  mesh_query_kernel<<<(unsigned)n, 256, 0, m->stream>>>(m->d_tris, m->nt, m->d_a, m->d_b, 3, 1, 0, m->d_hits, nullptr);
  m->launches += 1;
  std::vector<int32_t> hh((size_t)n);
  cudaMemcpyAsync(hh.data(), m->d_hits, (size_t)n * 4, cudaMemcpyDeviceToHost, m->stream);
  cudaError_t e = cudaStreamSynchronize(m->stream);
  if (e != cudaSuccess) { m->err = cudaGetErrorString(e); return ASSCPP_ECUDA; }
  for (int64_t i = 0; i < n; ++i) free_out[i] = hh[(size_t)i] == 0;  // heuristic.cpp:311-314
  return ASSCPP_OK;
}

}  // extern "C"
