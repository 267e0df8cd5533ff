// bvh.cuh — two-level BVH traversal over the flattened layout of dscene.cuh.
//
// Reproduces the reference's visit order and arithmetic so that first hits
// are bit-identical (yocto_bvh.cpp:699-881, yocto_geometry.h:786-868):
//   * a node's box is tested when it is popped, against the current tmax,
//     with the slab test's `t1 *= 1.00000024f` and ternary min/max;
//   * children are visited near-first by the sign of 1/d on the split axis
//     (the second-pushed child is popped first);
//   * leaf primitives are tested in stored order and `t <= tmax` accepts, so
//     the later of two equal-distance primitives wins;
//   * instance rays are transformed by the host-precomputed non-rigid inverse
//     frame and their direction is NOT renormalised (t is shared across spaces).
// The traversal stack holds node indices only; the shape-level walk runs on the
// same stack above the scene-level entries. The near child is kept in a
// register instead of a push/pop round trip (same order, fewer stack accesses).
#pragma once

#include "dscene.cuh"

namespace vpt {

// Per-thread traversal stack. On the device it is a column of a shared-memory
// array (stride = threads per block, so lanes hit distinct banks); the host
// build of this header uses a plain array.
#if defined(VPT_STACK_CHECK)
inline int vpt_stack_high_water = -1;  // tests/hostsim only
#endif
struct stack_t {
  int* base;
  int  stride;
  VPT_FN int  get(int i) const { return base[i * stride]; }
  VPT_FN void set(int i, int v) const {
#if defined(VPT_STACK_CHECK)
    if (i > vpt_stack_high_water) vpt_stack_high_water = i;
#endif
    base[i * stride] = v;
  }
};

constexpr int vpt_stack_depth = 64;  // uploader rejects deeper trees

VPT_FN bool intersect_box(f3 o, f3 dinv, float tmin, float tmax, f4 n0, f4 n1) {
  VPT_WORK(box_tests, 1);
  auto bmin   = f3{n0.x, n0.y, n0.z};
  auto bmax   = f3{n0.w, n1.x, n1.y};
  auto it_min = (bmin - o) * dinv;
  auto it_max = (bmax - o) * dinv;
  auto lo     = vmin(it_min, it_max);
  auto hi     = vmax(it_min, it_max);
  auto t0     = vmax(max3(lo), tmin);
  auto t1     = vmin(min3(hi), tmax);
  t1 *= 1.00000024f;
  return t0 <= t1;
}

// Moller-Trumbore, two-sided (yocto_geometry.h:786-819)
VPT_FN bool intersect_triangle(f3 o, f3 d, float tmin, float tmax, f3 p0, f3 p1, f3 p2,
    f2& uv, float& dist) {
  VPT_WORK(tri_tests, 1);
  auto edge1 = p1 - p0;
  auto edge2 = p2 - p0;
  auto pvec  = cross(d, edge2);
  auto det   = dot(edge1, pvec);
  if (det == 0) return false;
  auto inv_det = 1.0f / det;
  auto tvec    = o - p0;
  auto u       = dot(tvec, pvec) * inv_det;
  if (u < 0 || u > 1) return false;
  auto qvec = cross(tvec, edge1);
  auto v    = dot(d, qvec) * inv_det;
  if (v < 0 || u + v > 1) return false;
  auto t = dot(edge2, qvec) * inv_det;
  if (t < tmin || t > tmax) return false;
  uv   = {u, v};
  dist = t;
  return true;
}

// (p0,p1,p3) then (p2,p3,p1) with uv -> 1-uv (yocto_geometry.h:822-839)
VPT_FN bool intersect_quad(f3 o, f3 d, float tmin, float tmax, f3 p0, f3 p1, f3 p2, f3 p3,
    f2& uv, float& dist) {
  if (p2 == p3) return intersect_triangle(o, d, tmin, tmax, p0, p1, p3, uv, dist);
  auto hit = false;
  if (intersect_triangle(o, d, tmin, tmax, p0, p1, p3, uv, dist)) {
    hit  = true;
    tmax = dist;
  }
  if (intersect_triangle(o, d, tmin, tmax, p2, p3, p1, uv, dist)) {
    hit = true;
    uv  = {1 - uv.x, 1 - uv.y};
  }
  return hit;
}

// Shape-level walk in instance space; uses stack slots [base, ...).
#if defined(VPT_NOINLINE_TRAVERSAL)
VPT_NOINLINE
#else
VPT_FN
#endif
bool intersect_shape(const DScene& sc, int root, int kind, f3 o, f3 d, float tmin,
    float tmax, stack_t st, int base, int& element, f2& uv, float& distance) {
  if (root < 0) return false;
  auto hit  = false;
  auto dinv = f3{1 / d.x, 1 / d.y, 1 / d.z};
  int  neg[3] = {dinv.x < 0 ? 1 : 0, dinv.y < 0 ? 1 : 0, dinv.z < 0 ? 1 : 0};
  // single-exit loop: push root, pop until the stack is back at `base`
  int cur = base;
  st.set(cur++, root);
  VPT_GUARD_DECL(guard);
  while (cur != base) {
    VPT_GUARD_STEP(1, guard, 30000);
    int  ni = st.get(--cur);
    auto n0 = ld4(sc.nodes + 2 * ni), n1 = ld4(sc.nodes + 2 * ni + 1);
    if (intersect_box(o, dinv, tmin, tmax, n0, n1)) {
      auto start = f2i(n1.z), meta = f2i(n1.w);
      if (meta >> 24) {
        auto axis = (meta >> 16) & 3;
        auto sgn  = axis == 0 ? neg[0] : (axis == 1 ? neg[1] : neg[2]);
        // reference order: push (c0, c1) if the sign is set, else (c1, c0); the
        // last pushed child is visited first
        st.set(cur++, sgn ? start : start + 1);
        st.set(cur++, sgn ? start + 1 : start);
      } else {
        auto num = meta & 0xffff;
        if (kind == kind_quads) {
          for (auto i = 0; i < num; i++) {
            auto p = sc.prims + start + 4 * i;
            auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2), e = ld4(p + 3);
            if (intersect_quad(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), xyz(e), uv, distance)) {
              hit     = true;
              element = f2i(a.w);
              tmax    = distance;
            }
          }
        } else {
          for (auto i = 0; i < num; i++) {
            auto p = sc.prims + start + 3 * i;
            auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2);
            if (intersect_triangle(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), uv, distance)) {
              hit     = true;
              element = f2i(a.w);
              tmax    = distance;
            }
          }
        }
      }
    }
  }
#if defined(VPT_GUARD) && defined(__CUDA_ARCH__)
  VPT_COUNT_ADD(5, guard >> 4);
  VPT_COUNT_MAX(8, guard);
#endif
  return hit;
}

// Instance-targeted query (yocto_bvh.cpp:874-881), used by the light pdf.
VPT_FN hit_t intersect_instance(
    const DScene& sc, int instance, const ray3& ray, stack_t st, int base = 0) {
  auto h  = hit_t{instance, -1, {0, 0}, 0, false};
  VPT_WORK(instance_queries, 1);
  VPT_WORK(inst_transforms, 1);
  auto t  = sc.inst_trav + 4 * instance;
  auto c0 = ld4(t), c1 = ld4(t + 1), c2 = ld4(t + 2), c3 = ld4(t + 3);
  auto inv = frame3{xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
  auto o   = transform_point(inv, ray.o);
  auto d   = transform_vector(inv, ray.d);
  h.hit    = intersect_shape(sc, f2i(c0.w), f2i(c1.w), o, d, ray.tmin, ray.tmax, st, base,
         h.element, h.uv, h.distance);
  return h;
}

// Closest hit over the whole scene (yocto_bvh.cpp:800-871).
VPT_FN hit_t intersect_scene(const DScene& sc, const ray3& ray, stack_t st) {
  auto h = hit_t{-1, -1, {0, 0}, 0, false};
  VPT_WORK(scene_queries, 1);
  if (sc.scene_root < 0) return h;
  auto o = ray.o, d = ray.d;
  auto tmin = ray.tmin, tmax = ray.tmax;
  auto dinv = f3{1 / d.x, 1 / d.y, 1 / d.z};
  int  neg[3] = {dinv.x < 0 ? 1 : 0, dinv.y < 0 ? 1 : 0, dinv.z < 0 ? 1 : 0};
  int cur = 0;
  st.set(cur++, sc.scene_root);
  VPT_GUARD_DECL(guard);
  while (cur != 0) {
    VPT_GUARD_STEP(2, guard, 30000);
    int  ni = st.get(--cur);
    auto n0 = ld4(sc.nodes + 2 * ni), n1 = ld4(sc.nodes + 2 * ni + 1);
    if (intersect_box(o, dinv, tmin, tmax, n0, n1)) {
      auto start = f2i(n1.z), meta = f2i(n1.w);
      if (meta >> 24) {
        auto axis = (meta >> 16) & 3;
        auto sgn  = axis == 0 ? neg[0] : (axis == 1 ? neg[1] : neg[2]);
        st.set(cur++, sgn ? start : start + 1);
        st.set(cur++, sgn ? start + 1 : start);
      } else {
        auto num = meta & 0xffff;
        for (auto i = 0; i < num; i++) {
          auto inst = ld1(sc.scene_prims + start + i);
          VPT_WORK(inst_transforms, 1);
          auto t    = sc.inst_trav + 4 * inst;
          auto c0 = ld4(t), c1 = ld4(t + 1), c2 = ld4(t + 2), c3 = ld4(t + 3);
          auto inv = frame3{xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
          auto io  = transform_point(inv, o);
          auto id  = transform_vector(inv, d);
          if (intersect_shape(sc, f2i(c0.w), f2i(c1.w), io, id, tmin, tmax, st, cur, h.element,
                  h.uv, h.distance)) {
            h.hit      = true;
            h.instance = inst;
            tmax       = h.distance;
          }
        }
      }
    }
  }
  return h;
}

}  // namespace vpt
