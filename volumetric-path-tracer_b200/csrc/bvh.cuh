// bvh.cuh — two-level BVH traversal over the flattened layout of dscene.cuh.
//
// Reproduces the reference's visit order and arithmetic so that first hits
// are bit-identical (yocto_bvh.cpp:699-881, yocto_geometry.h:786-868):
//   * children are visited near-first by the sign of 1/d on the split axis
//     (the second-pushed child is popped first);
//   * leaf primitives are tested in stored order and `t <= tmax` accepts, so
//     the later of two equal-distance primitives wins;
//   * instance rays are transformed by the host-precomputed non-rigid inverse
//     frame and their direction is NOT renormalised (t is shared across spaces).
//
// Box tests. The reference tests a node's box when the node is POPPED, against
// the tmax of that moment: pass_pop = t0 <= min(hi, tmax_pop) * k, with
// t0 = max(lo, tmin), lo / hi the slab entry / exit and k = 1.00000024f. Here both
// children of a node are tested when the PARENT is expanded (one 64-byte record
// holds the two child boxes), failing children are never pushed, and a passing
// child is pushed with its t0. Because tmax only shrinks and float multiplication
// by k is monotone,
//     pass_pop  ==  pass_push  &&  t0 <= tmax_pop * k
// (if hi < tmax_pop both tests clamp to hi and agree; else the pop test IS
// t0 <= tmax_pop * k; a NaN t0 fails both, a NaN hi clamps to tmax in both), so
// the pop-time re-check is one multiply and one compare and the set and order of
// visited leaves are exactly the reference's. Root boxes (scene, instance) are
// tested explicitly when they are entered, as the reference does.
//
// Wide nodes. A ray whose origin and 1/d are finite ("wide" below) walks the 4-slot collapse
// of the same tree (dscene.cuh `wnodes`): one 128-byte record per visit, four slab tests, the
// passing slots pushed in the reference's visit order for the ray's sign octant. It reaches
// exactly the leaves the reference reaches, in the same order: for such a ray every slab term
// (b - o) * (1/d) is a monotone function of the box coordinate b (float subtraction and
// multiplication round monotonically, nothing is NaN), a descendant's box lies inside its
// ancestors' boxes coordinate by coordinate (boxes are exact float min / max unions), hence
// t0(child) >= t0(ancestor) and hi(child) <= hi(ancestor), and tmax only shrinks: a node that
// passes its own test at its pop time implies that every ancestor passed at its (earlier) pop
// time. So the intermediate boxes the collapse drops never decide anything, and the per-slot
// test + pop-time re-check is the reference's test of that node. For the same reason
// min(it_min, it_max) is the term of the box face on the near side of the sign of 1/d, so the
// near / far faces are picked by ADDRESS (which f4 of the record) and the six selects go.
// Rays with a zero direction component (1/d infinite -> NaN products, ternary min / max
// semantics) keep the binary records and the code above.
//
// Instances. The reference has no per-instance box test: every instance of a
// visited scene leaf gets its ray transformed (3 IEEE divisions for 1/d) and its
// root box tested in object space. Here a scene leaf first tests the ray against
// the instance's WORLD box, enlarged by 1e-4 of (extent + |coordinates|) — three
// orders of magnitude above the float error of either space — and skips the
// instance if it misses: such a ray passes no primitive of the instance at less
// than that margin, so the reference finds nothing there either (checked on every
// culled instance by tests/hostsim --cull-check). Rays with a zero direction
// component or an origin farther than 100x the box scale are never culled.
//
// Both levels run in ONE loop over ONE stack ("while-while" form). Per round a
// lane pops entries until it holds a leaf or an instance to enter, then the warp
// enters instances together and tests leaves together (ncu r01: with the instance
// entry inside the pop loop it ran 5.5 of 32 lanes, leaf tests 4.5 of 32).
// Stack entries are (ref, t0): ref >= 0 is a child-pair record, ref < 0 carries
// ~((payload << 2) | tag) with tag 0 = leaf {first primitive record << 3 | count},
// 1 = instance, 2 = sentinel that restores the world-space ray.
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
struct alignas(8) int2_t {
  int x, y;
};
struct stack_t {
  int* base;
  int  stride;
  // entry i = {ref, t0} as one 8-byte word at column offset 2 i stride (one LDS.64 / STS.64)
  struct entry_t {
    int   ref;
    float t0;
  };
  VPT_FN entry_t get(int i) const {
    auto v = *reinterpret_cast<const int2_t*>(base + 2 * i * stride);
    return {v.x, i2f(v.y)};
  }
  VPT_FN int   ref(int i) const { return get(i).ref; }
  VPT_FN float t0(int i) const { return get(i).t0; }
  VPT_FN void  push(int i, int r, float t) const {
#if defined(VPT_STACK_CHECK)
    if (i > vpt_stack_high_water) vpt_stack_high_water = i;
#endif
    *reinterpret_cast<int2_t*>(base + 2 * i * stride) = int2_t{r, f2i(t)};
  }
};
// The pooled extend kernel's stack (wavefront.cuh, k_wf_extend_pool): entries [0, limit) in the
// ray's shared-memory column, deeper ones in a global spill column (the uploader's depth bound
// is a worst case; walks rarely come near it).
struct pool_stack_t {
  int* base;
  int  stride;
  int* over;
  int  ostride, limit;
  VPT_FN stack_t::entry_t get(int i) const {
    int2_t v;
    if (i < limit) v = *reinterpret_cast<const int2_t*>(base + 2 * i * stride);
    else v = *reinterpret_cast<const int2_t*>(over + 2 * (size_t)(i - limit) * ostride);
    return {v.x, i2f(v.y)};
  }
  VPT_FN void push(int i, int r, float t) const {
    if (i < limit) *reinterpret_cast<int2_t*>(base + 2 * i * stride) = int2_t{r, f2i(t)};
    else *reinterpret_cast<int2_t*>(over + 2 * (size_t)(i - limit) * ostride) = int2_t{r, f2i(t)};
  }
};
constexpr int vpt_stack_words = 2;  // ints per entry (sizes the shared-memory columns)

constexpr int vpt_stack_depth = 64;  // uploader rejects deeper trees

// `plain` = all three 1/d are finite. Then no product below can be 0 * inf, nothing is NaN, and
// the reference's ternary min / max agree with fminf / fmaxf except for the sign of a zero,
// which cannot change `t0 <= t1` because t0 >= tmin > 0: the device takes the 1-instruction
// FMNMX form (12 instead of 24 instructions for the selects). Rays with a zero direction
// component keep the ternaries, whose NaN behaviour the reference depends on.
// `t0` (entry distance, independent of tmax) feeds the pop-time re-check.
constexpr float box_slack = 1.00000024f;
// max / min of two non-NaN floats (device: one FMNMX; equals the reference's ternaries up to
// the sign of a zero, which no comparison below can see)
VPT_FN float fmax2(float a, float b) {
#if defined(__CUDA_ARCH__)
  return fmaxf(a, b);
#else
  return a > b ? a : b;
#endif
}
VPT_FN float fmin2(float a, float b) {
#if defined(__CUDA_ARCH__)
  return fminf(a, b);
#else
  return a < b ? a : b;
#endif
}
VPT_FN int popc4(unsigned v) {
#if defined(__CUDA_ARCH__)
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}
VPT_FN bool intersect_box(f3 o, f3 dinv, float tmin, float tmax, f4 n0, f4 n1, bool plain,
    float& t0) {
  VPT_WORK(box_tests, 1);
  auto bmin   = f3{n0.x, n0.y, n0.z};
  auto bmax   = f3{n0.w, n1.x, n1.y};
  auto it_min = (bmin - o) * dinv;
  auto it_max = (bmax - o) * dinv;
#if defined(__CUDA_ARCH__)
  if (plain) {
    t0 = fmaxf(fmaxf(fmaxf(fminf(it_min.x, it_max.x), fminf(it_min.y, it_max.y)),
                   fminf(it_min.z, it_max.z)), tmin);
    auto t1 = fminf(fminf(fminf(fmaxf(it_min.x, it_max.x), fmaxf(it_min.y, it_max.y)),
                        fmaxf(it_min.z, it_max.z)), tmax);
    t1 *= box_slack;
    return t0 <= t1;
  }
#endif
  auto lo = vmin(it_min, it_max);
  auto hi = vmax(it_min, it_max);
  t0      = vmax(max3(lo), tmin);
  auto t1 = vmin(min3(hi), tmax);
  t1 *= box_slack;
  return t0 <= t1;
}

// Conservative world-space test of an instance's enlarged box (see the header). Only for
// `plain` rays; a NaN anywhere reports a hit.
VPT_FN bool misses_instance_box(f3 o, f3 dinv, float tmin, float tmax, f4 b0, f4 b1) {
  VPT_WORK(cull_tests, 1);
  auto it_min = (f3{b0.x, b0.y, b0.z} - o) * dinv;
  auto it_max = (f3{b0.w, b1.x, b1.y} - o) * dinv;
  auto t0 = fmaxf(fmaxf(fmaxf(fminf(it_min.x, it_max.x), fminf(it_min.y, it_max.y)),
                      fminf(it_min.z, it_max.z)), tmin);
  auto t1 = fminf(fminf(fminf(fmaxf(it_min.x, it_max.x), fmaxf(it_min.y, it_max.y)),
                      fmaxf(it_min.z, it_max.z)), tmax);
  return t0 > t1 * 1.001f;
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

#if defined(VPT_CULL_CHECK)
// tests/hostsim --cull-check: walks every culled instance as the reference would
inline bool vpt_cull_check_on = false;
inline void vpt_cull_check(const DScene& sc, int instance, const ray3& ray, float tmax);
#endif

constexpr int ref_leaf = 0, ref_instance = 1, ref_sentinel = 2, ref_empty = 3;
constexpr int wide_none = (int)0x80000001;  // "this BVH has no wide nodes" (never a valid ref)
VPT_FN constexpr int make_ref(int payload, int tag) { return ~((payload << 2) | tag); }
constexpr int leaf_start_limit = 1 << 26;  // f4 index of a leaf's first primitive record
#if !defined(VPT_TRAV_POPS)
#define VPT_TRAV_POPS 4
#endif
constexpr int trav_pops = VPT_TRAV_POPS;

// One ray's walk as resumable state, so that a plain loop (traverse) and the persistent
// extend kernel (wavefront.cuh, which refills finished lanes between rounds) run the same code.
struct trav_t {
  f3    o, d, dinv;  // ray in the current space (world or the entered instance's)
  f3    wdinv;       // 1 / d of the world-space ray, kept so leaving an instance costs no divisions
  float tmin, tmax;
  bool  plain, wplain;  // all of dinv finite (see intersect_box), for the current / world ray
  bool  wide, wwide;    // this space's walk uses the wide nodes (origin and dinv finite)
  int   oct, woct;      // sign octant of dinv: bit k set if dinv[k] < 0
  bool  cull;           // world-space instance culling allowed for this ray
  float omax;           // largest |coordinate| of the world-space origin
  int   kind;        // -1: scene level, else the element kind of the entered shape
  int   inst, cur;
  hit_t h;

  // `first`: < 0 for the whole scene, else the instance of the instance-targeted query
  template <class ST>
  VPT_FN void begin(const DScene& sc, int first, const ray3& ray, ST st) {
    h = hit_t{-1, -1, {0, 0}, 0, false};
    o = ray.o, d = ray.d;
    tmin = ray.tmin, tmax = ray.tmax;
    dinv  = f3{1 / d.x, 1 / d.y, 1 / d.z};
    plain = all_finite(dinv);
    wdinv = dinv, wplain = plain;
    omax  = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
    cull  = plain && vfinite(omax);
    oct   = octant(dinv);
    wide  = plain && all_finite(o) && sc.scene_wroot != wide_none;
    wwide = wide, woct = oct;
    kind = -1, inst = -1, cur = 0;
    if (first >= 0) {
      st.push(cur++, make_ref(first, ref_instance), 0);
    } else if (sc.num_instances > 0) {
      float t0;
      if (intersect_box(o, dinv, tmin, tmax, sc.scene_box0, sc.scene_box1, plain, t0))
        st.push(cur++, wide ? sc.scene_wroot : sc.scene_root, t0);
    }
  }
  VPT_FN bool done() const { return cur == 0; }
  static VPT_FN int octant(f3 di) {
    return (di.x < 0 ? 1 : 0) | (di.y < 0 ? 2 : 0) | (di.z < 0 ? 4 : 0);
  }

  // One visit of a wide node (see the header): four slab tests, pushes in visit order.
  template <class ST>
  VPT_FN void visit_wide(const DScene& sc, int e, ST st) {
    VPT_WORK(wide_visits, 1);
    auto p  = sc.wnodes + 8 * (size_t)e;
    int  sx = oct & 1, sy = (oct >> 1) & 1, sz = oct >> 2;
    auto nx = ld4(p + 3 * sx), fx = ld4(p + 3 - 3 * sx);
    auto ny = ld4(p + 1 + 3 * sy), fy = ld4(p + 4 - 3 * sy);
    auto nz = ld4(p + 2 + 3 * sz), fz = ld4(p + 5 - 3 * sz);
    auto rf = ld4(p + 6);
    auto ow = (unsigned)ld1(reinterpret_cast<const int*>(p + 7) + (oct >> 1)) >> ((oct & 1) * 16);
    unsigned m = 0;
#if defined(VPT_FAST)
    // statistical build: (b - o) * dinv as one FFMA per slab term, b * dinv + (-o * dinv)
    float ox = -o.x * dinv.x, oy = -o.y * dinv.y, oz = -o.z * dinv.z;
#define VPT_SLAB(b, d, od) fmaf(b, d, od)
#else
    const float ox = o.x, oy = o.y, oz = o.z;
#define VPT_SLAB(b, d, od) ((b - od) * d)
#endif
#define VPT_WIDE_SLOT(k, c)                                                                     \
  float lo##k = fmax2(fmax2(fmax2(VPT_SLAB(nx.c, dinv.x, ox), VPT_SLAB(ny.c, dinv.y, oy)),     \
                          VPT_SLAB(nz.c, dinv.z, oz)), tmin);                                 \
  float hi##k = fmin2(fmin2(fmin2(VPT_SLAB(fx.c, dinv.x, ox), VPT_SLAB(fy.c, dinv.y, oy)),     \
                          VPT_SLAB(fz.c, dinv.z, oz)), tmax) * box_slack;                     \
  int      r##k = f2i(rf.c);                                                                    \
  unsigned h##k = (ow >> (4 * k)) & 15u;                                                        \
  bool     p##k = lo##k <= hi##k && r##k != make_ref(0, ref_empty);                             \
  m |= p##k ? h##k : 0u;
    VPT_WIDE_SLOT(0, x)
    VPT_WIDE_SLOT(1, y)
    VPT_WIDE_SLOT(2, z)
    VPT_WIDE_SLOT(3, w)
#undef VPT_WIDE_SLOT
#undef VPT_SLAB
    VPT_WORK(wide_slot_tests, 4);
    // the first-visited slot (rank 0) goes on top: position = passing slots of higher rank
    if (p0) st.push(cur + popc4(m & (0u - (h0 << 1))), r0, lo0);
    if (p1) st.push(cur + popc4(m & (0u - (h1 << 1))), r1, lo1);
    if (p2) st.push(cur + popc4(m & (0u - (h2 << 1))), r2, lo2);
    if (p3) st.push(cur + popc4(m & (0u - (h3 << 1))), r3, lo3);
    cur += popc4(m);
  }

  // One round: pop until this lane holds a leaf or an instance to enter, its stack is empty
  // or it has popped `trav_pops` entries; then enter the instance / process the leaf. The
  // bound keeps lanes that already hold work from waiting on the lane with the longest
  // descent; it changes the SIMT schedule only.
  template <class ST>
  VPT_FN void round(const DScene& sc, const ray3& ray, ST st) {
    int leaf = descend(sc, ray, st);
    if (leaf != 0) test_leaf(sc, leaf);
  }

  // The four operations of a walk, shared by the bounded descent below and by the phase
  // machine (fetch / run_*) that the extend kernel elects between.
  // A binary node: both child boxes, pushed in the reference's order (yocto_bvh.cpp:829-846)
  template <class ST>
  VPT_FN void visit_binary(const DScene& sc, int e, ST st) {
    VPT_WORK(node_visits, 1);
    auto p  = sc.nodes + 4 * (size_t)e;
    auto a0 = ld4(p), a1 = ld4(p + 1), b0 = ld4(p + 2), b1 = ld4(p + 3);
    float ta, tb;
    bool  pa = intersect_box(o, dinv, tmin, tmax, a0, a1, plain, ta);
    bool  pb = intersect_box(o, dinv, tmin, tmax, b0, b1, plain, tb);
    auto  axis = f2i(a1.w);
    auto  sgn  = (axis == 0 ? dinv.x : (axis == 1 ? dinv.y : dinv.z)) < 0;
    // reference order: push (c0, c1) if the sign is set, else (c1, c0);
    // the last pushed child is visited first
    int   r1 = f2i(sgn ? a1.z : b1.z), r2 = f2i(sgn ? b1.z : a1.z);
    float t1 = sgn ? ta : tb, t2 = sgn ? tb : ta;
    bool  p1 = sgn ? pa : pb, p2 = sgn ? pb : pa;
    if (p1) st.push(cur++, r1, t1);
    if (p2) st.push(cur++, r2, t2);
  }
  template <class ST>
  VPT_FN void visit_node(const DScene& sc, int e, ST st) {
    if (wide) visit_wide(sc, e, st);
    else visit_binary(sc, e, st);
  }
  // back to the world-space ray (a sentinel was popped)
  VPT_FN void leave_instance(const ray3& ray) {
    o = ray.o, d = ray.d;
    dinv = wdinv, plain = wplain;
    wide = wwide, oct = woct;
    kind = -1;
  }
  // transform the ray, test the shape's root box (yocto_bvh.cpp:854-858, :714-722)
  template <class ST>
  VPT_FN void enter_instance(const DScene& sc, const ray3& ray, int enter, ST st) {
    VPT_WORK(inst_transforms, 1);
    auto t  = sc.inst_trav + 4 * enter;
    auto c0 = ld4(t), c1 = ld4(t + 1), c2 = ld4(t + 2), c3 = ld4(t + 3);
    if (f2i(c0.w) >= 0) {
      auto inv = frame3{xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
      auto r0 = ld4(sc.inst_root + 2 * enter), r1 = ld4(sc.inst_root + 2 * enter + 1);
      auto lo = transform_point(inv, ray.o);
      auto ld = transform_vector(inv, ray.d);
      auto li = f3{1 / ld.x, 1 / ld.y, 1 / ld.z};
      auto lp = all_finite(li);
      float t0;
      if (intersect_box(lo, li, tmin, tmax, r0, r1, lp, t0)) {
        o = lo, d = ld, dinv = li, plain = lp;
        wide = lp && all_finite(lo) && f2i(r1.w) != wide_none;
        oct  = octant(li);
        inst = enter;
        kind = f2i(c1.w);
        st.push(cur++, make_ref(0, ref_sentinel), 0);
        st.push(cur++, wide ? f2i(r1.w) : f2i(r1.z), t0);
      }
    }
  }
  // a scene leaf queues its instances (reverse order, so that they pop in leaf order)
  template <class ST>
  VPT_FN void queue_instances(const DScene& sc, const ray3& ray, int leaf, ST st) {
    int lstart = leaf >> 3, lnum = leaf & 7;
    for (auto i = lnum - 1; i >= 0; i--) {
      auto p  = sc.scene_leaves + 2 * (lstart + i);
      auto b0 = ld4(p), b1 = ld4(p + 1);
      bool skip = cull && omax <= b1.w && misses_instance_box(o, dinv, tmin, tmax, b0, b1);
#if defined(VPT_CULL_CHECK)
      if (skip && vpt_cull_check_on) vpt_cull_check(sc, f2i(b1.z), ray, tmax);
#endif
      if (!skip) st.push(cur++, make_ref(f2i(b1.z), ref_instance), 0);
    }
  }

  // The descent of one round. Returns the mesh leaf this lane now holds (0: none); the
  // caller tests its primitives, alone (test_leaf) or together with the other lanes of the
  // warp (wavefront.cuh, wf_test_leaves_warp).
  template <class ST>
  VPT_FN int descend(const DScene& sc, const ray3& ray, ST st) {
    int leaf = 0, enter = -1;
    int pops = trav_pops;
    VPT_WORK(rounds, 1);
    // The ray (o, d, dinv, plain, kind) is invariant inside the pop loop: leaving an instance
    // (sentinel) ends it and is handled outside, so the loop carries only the stack state
    // (with the restore inside it, ptxas moved ~20 registers of ray state around every pop).
    bool again = true;
    while (again) {
    while (cur != 0 && leaf == 0 && enter == -1 && pops != 0) {
      pops--;
      VPT_WORK(pops, 1);
      auto  en = st.get(--cur);
      int   e  = en.ref;
      float te = en.t0;
      if (e >= 0) {
        if (!(te <= tmax * box_slack)) {
          VPT_WORK(stale_node_pops, 1);
        } else {
          visit_node(sc, e, st);
        }
      } else {
        int x = ~e, tag = x & 3;
        if (tag == ref_leaf) {
          if (te <= tmax * box_slack) leaf = x >> 2;
          else VPT_WORK(stale_pops, 1);
          VPT_WORK(leaf_pops, 1);
        } else if (tag == ref_instance) {
          enter = x >> 2;
        } else {
          enter = -2;
        }
      }
    }
    again = false;
    if (enter == -2) {
      leave_instance(ray);
      enter = -1;
      again = cur != 0 && pops != 0;
    }
    }
    if (enter >= 0) enter_instance(sc, ray, enter, st);
    if (leaf != 0 && kind < 0) {
      queue_instances(sc, ray, leaf, st);
      leaf = 0;
    }
    return leaf;
  }

  // ---- The same walk as a phase machine (k_wf_extend_persistent) ----
  // fetch() pops until the lane holds one of the three expensive operations - a node visit, an
  // instance entry, a mesh leaf - or its stack is empty or `pops` entries are gone; the cheap
  // ones (stale entries, sentinels, scene leaves) are done on the way. The kernel then runs
  // ONE kind of operation for all the lanes of the warp that hold it (run_node / run_instance
  // / test_leaf). Per ray this is the sequence of operations of descend(): only the SIMT
  // schedule differs, hits are bit-identical.
  static constexpr int op_none = 0, op_node = 1, op_leaf = 2, op_instance = 3;
  template <class ST>
  VPT_FN int fetch(const DScene& sc, const ray3& ray, ST st, int pops, int& arg) {
    while (cur != 0 && pops != 0) {
      pops--;
      VPT_WORK(pops, 1);
      auto  en = st.get(--cur);
      int   e  = en.ref;
      float te = en.t0;
      if (e >= 0) {
        if (te <= tmax * box_slack) {
          arg = e;
          return op_node;
        }
        VPT_WORK(stale_node_pops, 1);
      } else {
        int x = ~e, tag = x & 3;
        if (tag == ref_leaf) {
          VPT_WORK(leaf_pops, 1);
          if (te <= tmax * box_slack) {
            if (kind >= 0) {
              arg = x >> 2;
              return op_leaf;
            }
            queue_instances(sc, ray, x >> 2, st);
          } else {
            VPT_WORK(stale_pops, 1);
          }
        } else if (tag == ref_instance) {
          arg = x >> 2;
          return op_instance;
        } else {
          leave_instance(ray);
        }
      }
    }
    return op_none;
  }

  // mesh leaf: primitives in stored order, `t <= tmax` accepts
  VPT_FN void test_leaf(const DScene& sc, int leaf) {
    int lstart = leaf >> 3, lnum = leaf & 7;
    if (kind == kind_quads) {
      for (auto i = 0; i < lnum; i++) {
        auto p = sc.prims + lstart + 4 * i;
        auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2), q = ld4(p + 3);
        if (intersect_quad(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), xyz(q), h.uv, h.distance)) {
          h.hit = true, h.instance = inst, h.element = f2i(a.w);
          tmax = h.distance;
        }
      }
    } else {
      for (auto i = 0; i < lnum; i++) {
        auto p = sc.prims + lstart + 3 * i;
        auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2);
        if (intersect_triangle(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), h.uv, h.distance)) {
          h.hit = true, h.instance = inst, h.element = f2i(a.w);
          tmax = h.distance;
        }
      }
    }
  }
};

// Closest hit of the scene (`first` < 0) or of one instance. Uses stack slots [0, ...).
VPT_FN hit_t traverse(const DScene& sc, int first, const ray3& ray, stack_t st) {
  trav_t t;
  t.begin(sc, first, ray, st);
  VPT_GUARD_DECL(guard);
  while (!t.done()) {
    VPT_GUARD_STEP(2, guard, 30000);
    t.round(sc, ray, st);
  }
  return t.h;
}

#if defined(VPT_CULL_CHECK)
inline void vpt_cull_check(const DScene& sc, int instance, const ray3& ray, float tmax) {
  alignas(8) int mem[vpt_stack_depth * vpt_stack_words];
  auto r = ray;
  r.tmax = tmax;
  VPT_WORK(culled, 1);
  if (traverse(sc, instance, r, stack_t{mem, 1}).hit) VPT_WORK(cull_wrong, 1);
}
#endif

// Instance-targeted query (yocto_bvh.cpp:874-881), used by the light pdf.
VPT_FN hit_t intersect_instance(const DScene& sc, int instance, const ray3& ray, stack_t st) {
  VPT_WORK(instance_queries, 1);
#if defined(VPT_WORK_COUNTERS) && !defined(__CUDACC__)
  // share of the box / primitive tests spent in light-pdf queries (run hostsim with --threads 1)
  long long b0 = g_work.box_tests, t0 = g_work.tri_tests;
#endif
  auto h     = traverse(sc, instance, ray, st);
  h.instance = instance;
#if defined(VPT_WORK_COUNTERS) && !defined(__CUDACC__)
  g_work.iq_box_tests += g_work.box_tests - b0, g_work.iq_tri_tests += g_work.tri_tests - t0;
#endif
  return h;
}

// Closest hit over the whole scene (yocto_bvh.cpp:800-871).
#if defined(VPT_RAY_SINK) && !defined(__CUDACC__)
}  // namespace vpt
#include <vector>
namespace vpt {
inline std::vector<ray3>* vpt_ray_sink = nullptr;  // tests/hostsim/warpsim: every scene query
#endif
VPT_FN hit_t intersect_scene(const DScene& sc, const ray3& ray, stack_t st) {
  VPT_WORK(scene_queries, 1);
#if defined(VPT_RAY_SINK) && !defined(__CUDACC__)
  if (vpt_ray_sink) vpt_ray_sink->push_back(ray);
#endif
  return traverse(sc, -1, ray, st);
}

}  // namespace vpt
