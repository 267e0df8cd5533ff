// tesselate.cuh — Catmull-Clark subdivision, vertex normals and displacement on the device
// (reference: tesselate_catmullclark / tesselate_surface,
// yocto_pathtrace.cpp:1119-1273, with make_edge_map / get_edges / get_boundary
// yocto_shape.cpp:1755-1814 and quads_normals / triangles_normals :1478-1510).
//
// Results equal the reference's bit for bit. What that takes on a parallel machine:
//   * edge numbering: the reference numbers edges in the order insert_edge first meets them
//     while walking the faces. Here every face side becomes a (key = min << 32 | max, position)
//     pair, a stable radix sort groups equal keys with their first position in front, and a
//     second sort of the unique edges by first position gives the same numbers;
//   * every per-vertex accumulation (`avert[vid] += c` over the new quads, `normals[v] += n`
//     over the faces) is a chain of float additions in face order. The (face, corner) entries
//     are stable-sorted by vertex and ONE thread per vertex adds its entries in that order;
//   * boundary terms: a vertex of a manifold boundary receives two crease contributions, and
//     a + b = b + a exactly, so their order (the iteration order of an unordered_map in the
//     reference) does not matter; they are added with atomics. Non-manifold boundaries
//     (three or more) would depend on that unknowable order, there as here.
// dim = 3 for positions (lock_boundary = false), 2 for texcoords (lock_boundary = true).
#pragma once

#include <cuda_runtime.h>

#include <cub/cub.cuh>

#include "scene_eval.cuh"

namespace vpt {

constexpr int tess_block = 256;

// face sides in the reference's insertion order: (x,y) (y,z) (z,w) (w,x); a triangle is a quad
// with z == w and has no (z,w) side
__device__ __forceinline__ bool tess_side(const int4& q, int k, int& a, int& b) {
  if (k == 0) a = q.x, b = q.y;
  else if (k == 1) a = q.y, b = q.z;
  else if (k == 2) a = q.z, b = q.w;
  else a = q.w, b = q.x;
  return !(k == 2 && q.z == q.w);
}

__global__ void __launch_bounds__(tess_block)
    k_tess_sides(const int4* __restrict__ quads, int nf, unsigned long long* __restrict__ keys,
        int* __restrict__ pos, int* __restrict__ face_count) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= 4 * nf) return;
  auto q = quads[s >> 2];
  int  a, b;
  bool valid = tess_side(q, s & 3, a, b);
  int  lo = a < b ? a : b, hi = a < b ? b : a;
  keys[s] = valid ? (((unsigned long long)(unsigned)lo << 32) | (unsigned)hi) : ~0ull;
  pos[s]  = s;
  if ((s & 3) == 0) face_count[s >> 2] = q.z != q.w ? 4 : 3;
}

// heads of the runs of equal keys in the sorted side list
__global__ void __launch_bounds__(tess_block)
    k_tess_heads(const unsigned long long* __restrict__ keys, int n, int* __restrict__ head) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  head[s] = keys[s] != ~0ull && (s == 0 || keys[s] != keys[s - 1]) ? 1 : 0;
}

// per unique edge u (run): first position, number of faces, its vertices
__global__ void __launch_bounds__(tess_block)
    k_tess_uniques(const unsigned long long* __restrict__ keys, const int* __restrict__ pos,
        const int* __restrict__ head, const int* __restrict__ run, int n,
        int* __restrict__ first_pos, int* __restrict__ uid, int* __restrict__ run_start) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n || !head[s]) return;
  int u        = run[s];  // exclusive scan of head
  first_pos[u] = pos[s];  // stable sort: the earliest side of the run comes first
  uid[u]       = u;
  run_start[u] = s;
}

// edge r (numbered by first appearance) = unique order[r]
__global__ void __launch_bounds__(tess_block)
    k_tess_edges(const unsigned long long* __restrict__ keys, const int* __restrict__ order,
        const int* __restrict__ run_start, int ne, int nsides_valid, int2* __restrict__ edges,
        int* __restrict__ nfaces, int* __restrict__ index_of_unique) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= ne) return;
  int  u     = order[r];
  int  s     = run_start[u];
  auto key   = keys[s];
  edges[r]   = make_int2((int)(key >> 32), (int)(key & 0xffffffffu));
  int  count = 1;
  while (s + count < nsides_valid && keys[s + count] == key) count++;
  nfaces[r]          = count;
  index_of_unique[u] = r;
}

// edge index of every face side
__global__ void __launch_bounds__(tess_block)
    k_tess_side_edges(const unsigned long long* __restrict__ keys, const int* __restrict__ pos,
        const int* __restrict__ head, const int* __restrict__ run,
        const int* __restrict__ index_of_unique, int n, int* __restrict__ side_edge) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n || keys[s] == ~0ull) return;
  int u             = run[s] + head[s] - 1;  // run id of a non-head = scan value - 1
  side_edge[pos[s]] = index_of_unique[u];
}

template <int D>
struct tess_vec {
  float v[D];
};
template <int D>
__device__ __forceinline__ tess_vec<D> tess_ld(const float* p, int i) {
  tess_vec<D> r;
  for (int k = 0; k < D; k++) r.v[k] = p[(size_t)D * i + k];
  return r;
}
template <int D>
__device__ __forceinline__ tess_vec<D> operator+(tess_vec<D> a, tess_vec<D> b) {
  for (int k = 0; k < D; k++) a.v[k] = a.v[k] + b.v[k];
  return a;
}
template <int D>
__device__ __forceinline__ tess_vec<D> operator/(tess_vec<D> a, float s) {
  for (int k = 0; k < D; k++) a.v[k] = a.v[k] / s;
  return a;
}

// new vertices: [0, nv) old, [nv, nv + ne) edge midpoints, [nv + ne, ..) face centres
template <int D>
__global__ void __launch_bounds__(tess_block)
    k_tess_vertices(const float* __restrict__ vert, int nv, const int2* __restrict__ edges, int ne,
        const int4* __restrict__ quads, int nf, float* __restrict__ tverts) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nv + ne + nf) return;
  tess_vec<D> r;
  if (i < nv) {
    r = tess_ld<D>(vert, i);
  } else if (i < nv + ne) {
    auto e = edges[i - nv];
    r      = (tess_ld<D>(vert, e.x) + tess_ld<D>(vert, e.y)) / 2.0f;
  } else {
    auto q = quads[i - nv - ne];
    if (q.z != q.w)
      r = (tess_ld<D>(vert, q.x) + tess_ld<D>(vert, q.y) + tess_ld<D>(vert, q.z) + tess_ld<D>(vert, q.w)) / 4.0f;
    else
      r = (tess_ld<D>(vert, q.x) + tess_ld<D>(vert, q.y) + tess_ld<D>(vert, q.z)) / 3.0f;
  }
  for (int k = 0; k < D; k++) tverts[(size_t)D * i + k] = r.v[k];
}

// new faces (yocto_pathtrace.cpp:1147-1167)
__global__ void __launch_bounds__(tess_block)
    k_tess_faces(const int4* __restrict__ quads, int nf, const int* __restrict__ side_edge,
        const int* __restrict__ face_offset, int nv, int ne, int4* __restrict__ tquads) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nf) return;
  auto q = quads[i];
  int  o = face_offset[i], c = nv + ne + i;
  int  e0 = nv + side_edge[4 * i], e1 = nv + side_edge[4 * i + 1], e3 = nv + side_edge[4 * i + 3];
  if (q.z != q.w) {
    int e2        = nv + side_edge[4 * i + 2];
    tquads[o]     = make_int4(q.x, e0, c, e3);
    tquads[o + 1] = make_int4(q.y, e1, c, e0);
    tquads[o + 2] = make_int4(q.z, e2, c, e1);
    tquads[o + 3] = make_int4(q.w, e3, c, e2);
  } else {
    tquads[o]     = make_int4(q.x, e0, c, e3);
    tquads[o + 1] = make_int4(q.y, e1, c, e0);
    tquads[o + 2] = make_int4(q.z, e3, c, e1);
  }
}

// boundary edges: vertex valence classes and the crease contributions
template <int D>
__global__ void __launch_bounds__(tess_block)
    k_tess_boundary(const int2* __restrict__ edges, const int* __restrict__ nfaces, int ne, int nv,
        int lock_boundary, const float* __restrict__ tverts, signed char* __restrict__ val,
        float* __restrict__ avert, int* __restrict__ acount, int pass) {
  int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= ne || nfaces[r] >= 2) return;
  int a = edges[r].x, m = nv + r, b = edges[r].y;
  if (pass == 0) {  // tverts_val (yocto_pathtrace.cpp:1186-1190)
    val[a] = val[m] = val[b] = lock_boundary ? 0 : 1;
    return;
  }
  // tboundary = {a, m}, {m, b}
  if (lock_boundary) {
    // crease vertices a, m, m, b: avert[p] += tverts[p] (:1195-1199)
    int ps[4] = {a, m, m, b};
    for (int p : ps) {
      for (int k = 0; k < D; k++) atomicAdd(&avert[(size_t)D * p + k], tverts[(size_t)D * p + k]);
      atomicAdd(&acount[p], 1);
    }
  } else {
    // crease edges: both ends get the edge's midpoint (:1200-1207)
    int2 es[2] = {make_int2(a, m), make_int2(m, b)};
    for (auto e : es) {
      auto c = (tess_ld<D>(tverts, e.x) + tess_ld<D>(tverts, e.y)) / 2.0f;
      for (int p : {e.x, e.y}) {
        for (int k = 0; k < D; k++) atomicAdd(&avert[(size_t)D * p + k], c.v[k]);
        atomicAdd(&acount[p], 1);
      }
    }
  }
}

// (vertex, entry) pairs of the new quads for the ordered per-vertex sums
__global__ void __launch_bounds__(tess_block)
    k_tess_corners(const int4* __restrict__ tquads, int nq, unsigned* __restrict__ vkey,
        int* __restrict__ entry) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= 4 * nq) return;
  auto q   = tquads[s >> 2];
  int  c   = s & 3;
  vkey[s]  = (unsigned)(c == 0 ? q.x : (c == 1 ? q.y : (c == 2 ? q.z : q.w)));
  entry[s] = s;
}

// averaging + correction (yocto_pathtrace.cpp:1209-1224): one thread per vertex walks its
// entries in (quad, corner) order
template <int D>
__global__ void __launch_bounds__(tess_block)
    k_tess_average(const int4* __restrict__ tquads, const unsigned* __restrict__ vkey,
        const int* __restrict__ entry, int nentries, const float* __restrict__ tverts, int ntv,
        const signed char* __restrict__ val, float* __restrict__ avert, int* __restrict__ acount) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ntv) return;
  tess_vec<D> sum;
  for (int k = 0; k < D; k++) sum.v[k] = avert[(size_t)D * i + k];
  int count = acount[i];
  if (val[i] == 2) {
    // first entry of vertex i in the sorted list (binary search)
    int lo = 0, hi = nentries;
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (vkey[mid] < (unsigned)i) lo = mid + 1; else hi = mid;
    }
    for (int s = lo; s < nentries && vkey[s] == (unsigned)i; s++) {
      auto q = tquads[entry[s] >> 2];
      auto c = (tess_ld<D>(tverts, q.x) + tess_ld<D>(tverts, q.y) + tess_ld<D>(tverts, q.z) +
                   tess_ld<D>(tverts, q.w)) / 4.0f;
      sum = sum + c;
      count += 1;
    }
  }
  sum = sum / (float)count;
  if (val[i] == 2) {
    float w = 4 / (float)count;
    for (int k = 0; k < D; k++) {
      float t  = tverts[(size_t)D * i + k];
      sum.v[k] = t + (sum.v[k] - t) * w;
    }
  }
  for (int k = 0; k < D; k++) avert[(size_t)D * i + k] = sum.v[k];
}

// ---- vertex normals (quads_normals / triangles_normals, yocto_shape.cpp:1478-1510) ----------
__global__ void __launch_bounds__(tess_block)
    k_tess_elem_corners(const int* __restrict__ elems, int ne, int verts_per,
        unsigned* __restrict__ vkey, int* __restrict__ entry) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= verts_per * ne) return;
  int e = s / verts_per, c = s % verts_per;
  int v = elems[s];
  // a triangle stored as a quad (z == w) contributes to w only once (:1506)
  bool skip = verts_per == 4 && c == 3 && elems[verts_per * e + 2] == v;
  vkey[s]   = skip ? 0xffffffffu : (unsigned)v;
  entry[s]  = s;
}

__device__ __forceinline__ f3 tess_p(const float* p, int i) {
  return {p[3 * (size_t)i], p[3 * (size_t)i + 1], p[3 * (size_t)i + 2]};
}

__global__ void __launch_bounds__(tess_block)
    k_tess_normals(const int* __restrict__ elems, int verts_per, const unsigned* __restrict__ vkey,
        const int* __restrict__ entry, int nentries, const float* __restrict__ positions, int nv,
        float* __restrict__ normals) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nv) return;
  int lo = 0, hi = nentries;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (vkey[mid] < (unsigned)i) lo = mid + 1; else hi = mid;
  }
  auto sum = f3{0, 0, 0};
  for (int s = lo; s < nentries && vkey[s] == (unsigned)i; s++) {
    int  e  = entry[s] / verts_per;
    auto q  = elems + (size_t)verts_per * e;
    auto p0 = tess_p(positions, q[0]), p1 = tess_p(positions, q[1]), p2 = tess_p(positions, q[2]);
    f3    n;
    float area;
    if (verts_per == 3) {
      n    = normalize(cross(p1 - p0, p2 - p0));
      area = length(cross(p1 - p0, p2 - p0)) / 2;
    } else {
      auto p3 = tess_p(positions, q[3]);
      n       = normalize(normalize(cross(p1 - p0, p3 - p0)) + normalize(cross(p3 - p2, p1 - p2)));
      area    = length(cross(p1 - p0, p3 - p0)) / 2 + length(cross(p3 - p2, p1 - p2)) / 2;
    }
    sum = sum + n * area;
  }
  sum = normalize(sum);
  normals[3 * (size_t)i] = sum.x, normals[3 * (size_t)i + 1] = sum.y, normals[3 * (size_t)i + 2] = sum.z;
}

// ---- displacement (yocto_pathtrace.cpp:1256-1263) ---------------------------------------------
// `sc` carries only the one texture (record 0) and the sRGB table
__global__ void __launch_bounds__(tess_block)
    k_tess_displace(const DScene sc, int byte_texture, float displacement,
        const float* __restrict__ normals, const float* __restrict__ texcoords, int nv,
        float* __restrict__ positions) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nv) return;
  auto uv   = f2{texcoords[2 * (size_t)i], texcoords[2 * (size_t)i + 1]};
  auto tx   = eval_texture_rec(sc, ld4(sc.textures), uv, true);
  auto disp = (tx.x + tx.y + tx.z) / 3;
  if (byte_texture) disp -= 0.5f;
  auto n = tess_p(normals, i);
  auto p = tess_p(positions, i) + n * displacement * disp;
  positions[3 * (size_t)i] = p.x, positions[3 * (size_t)i + 1] = p.y, positions[3 * (size_t)i + 2] = p.z;
}

}  // namespace vpt
