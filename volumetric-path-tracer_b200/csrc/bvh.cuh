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
//
// Both levels run in ONE loop over ONE stack ("while-while" form): every lane
// first pops entries until it holds a leaf, then the warp processes leaves
// together. Stack entries are node indices (>= 0), instance entries (~instance,
// pushed in reverse so they pop in leaf order) and a sentinel that restores the
// world-space ray when an instance's subtree is exhausted. Per lane the visit
// order is exactly the nested two-level walk's; only the SIMT schedule differs
// (ncu r01: the nested form ran 2.8 of 32 lanes, leaf tests 1.2 of 32).
// Loops are single-exit on purpose: multi-exit forms of this loop hung on
// sm_100a ("single-exit loops").
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

// `plain` = all three 1/d are finite. Then no product below can be 0 * inf, nothing is NaN, and
// the reference's ternary min / max agree with fminf / fmaxf except for the sign of a zero,
// which cannot change `t0 <= t1` because t0 >= tmin > 0: the device takes the 1-instruction
// FMNMX form (12 instead of 24 instructions for the selects). Rays with a zero direction
// component keep the ternaries, whose NaN behaviour the reference depends on.
VPT_FN bool intersect_box(f3 o, f3 dinv, float tmin, float tmax, f4 n0, f4 n1, bool plain) {
  VPT_WORK(box_tests, 1);
  auto bmin   = f3{n0.x, n0.y, n0.z};
  auto bmax   = f3{n0.w, n1.x, n1.y};
  auto it_min = (bmin - o) * dinv;
  auto it_max = (bmax - o) * dinv;
#if defined(__CUDA_ARCH__)
  if (plain) {
    auto t0 = fmaxf(fmaxf(fmaxf(fminf(it_min.x, it_max.x), fminf(it_min.y, it_max.y)),
                        fminf(it_min.z, it_max.z)), tmin);
    auto t1 = fminf(fminf(fminf(fmaxf(it_min.x, it_max.x), fmaxf(it_min.y, it_max.y)),
                        fmaxf(it_min.z, it_max.z)), tmax);
    t1 *= 1.00000024f;
    return t0 <= t1;
  }
#endif
  auto lo = vmin(it_min, it_max);
  auto hi = vmax(it_min, it_max);
  auto t0 = vmax(max3(lo), tmin);
  auto t1 = vmin(min3(hi), tmax);
  t1 *= 1.00000024f;
  return t0 <= t1;
}
VPT_FN bool all_finite(f3 v) { return vfinite(v.x) && vfinite(v.y) && vfinite(v.z); }

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

constexpr int stack_sentinel = (int)0x80000000;
#if !defined(VPT_TRAV_POPS)
#define VPT_TRAV_POPS 8
#endif
constexpr int trav_pops = VPT_TRAV_POPS;

// One ray's walk as resumable state, so that a plain loop (traverse) and the persistent
// extend kernel (wavefront.cuh, which refills finished lanes between rounds) run the same code.
struct trav_t {
  f3    o, d, dinv;  // ray in the current space (world or the entered instance's)
  f3    wdinv;       // 1 / d of the world-space ray, kept so leaving an instance costs no divisions
  float tmin, tmax;
  bool  plain, wplain;  // all of dinv finite (see intersect_box), for the current / world ray
  int   kind;        // -1: scene level, else the element kind of the entered shape
  int   inst, cur;
  hit_t h;

  // `first`: a scene node index, or ~instance for the instance-targeted query
  VPT_FN void begin(int first, const ray3& ray, stack_t st) {
    h = hit_t{-1, -1, {0, 0}, 0, false};
    o = ray.o, d = ray.d;
    tmin = ray.tmin, tmax = ray.tmax;
    dinv  = f3{1 / d.x, 1 / d.y, 1 / d.z};
    plain = all_finite(dinv);
    wdinv = dinv, wplain = plain;
    kind = -1, inst = -1, cur = 0;
    st.set(cur++, first);
  }
  VPT_FN bool done() const { return cur == 0; }

  // One round: pop until this lane holds a leaf, its stack is empty or it has popped
  // `trav_pops` entries, then process the leaf if it has one. The bound keeps lanes that
  // already hold a leaf from waiting on the lane with the longest descent (pops between
  // leaves average ~10 and are roughly geometric, so the maximum over a warp was ~22 per
  // round in ncu r01 while the mean lane needed 10); it changes the SIMT schedule only.
  VPT_FN void round(const DScene& sc, const ray3& ray, stack_t st) {
    int lstart = 0, lnum = 0;
    int pops = trav_pops;
    while (cur != 0 && lnum == 0 && pops != 0) {
      pops--;
      int e = st.get(--cur);
      if (e >= 0) {
        auto n0 = ld4(sc.nodes + 2 * e), n1 = ld4(sc.nodes + 2 * e + 1);
        if (intersect_box(o, dinv, tmin, tmax, n0, n1, plain)) {
          auto start = f2i(n1.z), meta = f2i(n1.w);
          if (meta >> 24) {
            auto axis = (meta >> 16) & 3;
            auto sgn  = (axis == 0 ? dinv.x : (axis == 1 ? dinv.y : dinv.z)) < 0;
            // reference order: push (c0, c1) if the sign is set, else (c1, c0);
            // the last pushed child is visited first
            st.set(cur++, sgn ? start : start + 1);
            st.set(cur++, sgn ? start + 1 : start);
          } else {
            lstart = start, lnum = meta & 0xffff;
          }
        }
      } else if (e == stack_sentinel) {
        o = ray.o, d = ray.d;
        dinv = wdinv, plain = wplain;
        kind = -1;
      } else {
        inst = ~e;
        VPT_WORK(inst_transforms, 1);
        auto t  = sc.inst_trav + 4 * inst;
        auto c0 = ld4(t), c1 = ld4(t + 1), c2 = ld4(t + 2), c3 = ld4(t + 3);
        auto inv = frame3{xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
        auto root = f2i(c0.w);
        if (root >= 0) {
          o     = transform_point(inv, ray.o);
          d     = transform_vector(inv, ray.d);
          dinv  = f3{1 / d.x, 1 / d.y, 1 / d.z};
          plain = all_finite(dinv);
          kind  = f2i(c1.w);
          st.set(cur++, stack_sentinel);
          st.set(cur++, root);
        }
      }
    }
    // leaves: primitives in stored order; a scene leaf queues its instances
    if (kind == kind_quads) {
      for (auto i = 0; i < lnum; i++) {
        auto p = sc.prims + lstart + 4 * i;
        auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2), q = ld4(p + 3);
        if (intersect_quad(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), xyz(q), h.uv, h.distance)) {
          h.hit = true, h.instance = inst, h.element = f2i(a.w);
          tmax = h.distance;
        }
      }
    } else if (kind >= 0) {
      for (auto i = 0; i < lnum; i++) {
        auto p = sc.prims + lstart + 3 * i;
        auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2);
        if (intersect_triangle(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), h.uv, h.distance)) {
          h.hit = true, h.instance = inst, h.element = f2i(a.w);
          tmax = h.distance;
        }
      }
    } else {
      for (auto i = lnum - 1; i >= 0; i--) st.set(cur++, ~ld1(sc.scene_prims + lstart + i));
    }
  }
};

// Closest hit below `first`. Uses stack slots [0, ...).
VPT_FN hit_t traverse(const DScene& sc, int first, const ray3& ray, stack_t st) {
  trav_t t;
  t.begin(first, ray, st);
  VPT_GUARD_DECL(guard);
  while (!t.done()) {
    VPT_GUARD_STEP(2, guard, 30000);
    t.round(sc, ray, st);
  }
  return t.h;
}

// Instance-targeted query (yocto_bvh.cpp:874-881), used by the light pdf.
VPT_FN hit_t intersect_instance(const DScene& sc, int instance, const ray3& ray, stack_t st) {
  VPT_WORK(instance_queries, 1);
#if defined(VPT_WORK_COUNTERS) && !defined(__CUDACC__)
  // share of the box / primitive tests spent in light-pdf queries (run hostsim with --threads 1)
  long long b0 = g_work.box_tests, t0 = g_work.tri_tests;
#endif
  auto h     = traverse(sc, ~instance, ray, st);
  h.instance = instance;
#if defined(VPT_WORK_COUNTERS) && !defined(__CUDACC__)
  g_work.iq_box_tests += g_work.box_tests - b0, g_work.iq_tri_tests += g_work.tri_tests - t0;
#endif
  return h;
}

// Closest hit over the whole scene (yocto_bvh.cpp:800-871).
VPT_FN hit_t intersect_scene(const DScene& sc, const ray3& ray, stack_t st) {
  VPT_WORK(scene_queries, 1);
  if (sc.scene_root < 0) return hit_t{-1, -1, {0, 0}, 0, false};
  return traverse(sc, sc.scene_root, ray, st);
}

}  // namespace vpt
