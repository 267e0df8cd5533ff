// sdf.cuh — implicit surfaces on the device: analytic distance functions as
// typed functors, voxel-grid SDFs with a box proxy and float trilinear
// interpolation, scene-wide minimum, sphere tracing and tetrahedron
// finite-difference normals. Restates yocto_sdfs.h:43-95, yocto_sdfs.cpp:7-128
// and yocto_pathtrace.cpp:259-307. Note that world -> local uses the FORWARD
// frame (yocto_sdfs.cpp:13, :20), as the reference does.
#pragma once

#include "dscene.cuh"

namespace vpt {

enum : int { sdf_bbox = 0, sdf_box, sdf_capped_cone, sdf_plane, sdf_sphere, sdf_torus };

VPT_FN float sd_box(f3 p, f3 b) {
  auto d = vabs(p) - b;
  return vmin(vmax(d.x, vmax(d.y, d.z)), 0.0f) + length(vmax(d, 0.0f));
}
VPT_FN float sd_bbox(f3 p, f3 b, float e) {
  p      = vabs(p) - b;
  auto q = vabs(p + e) - e;
  return vmin(vmin(length(vmax(f3{p.x, q.y, q.z}, 0.0f)) + vmin(vmax(p.x, vmax(q.y, q.z)), 0.0f),
                  length(vmax(f3{q.x, p.y, q.z}, 0.0f)) + vmin(vmax(q.x, vmax(p.y, q.z)), 0.0f)),
      length(vmax(f3{q.x, q.y, p.z}, 0.0f)) + vmin(vmax(q.x, vmax(q.y, p.z)), 0.0f));
}
VPT_FN float sd_torus(f3 p, float r1, float r2) {
  return length(f2{length(f2{p.x, p.z}) - r1, p.y}) - r2;
}
VPT_FN float sd_capped_cone(f3 p, float h, float r1, float r2) {
  auto q  = f2{length(f2{p.x, p.z}), p.y};
  auto k1 = f2{r2, h};
  auto k2 = f2{r2 - r1, 2.0f * h};
  auto ca = f2{q.x - vmin(q.x, (q.y < 0.0f) ? r1 : r2), vabs(q.y) - h};
  auto cb = q - k1 + k2 * vclamp(dot(k1 - q, k2) / dot(k2, k2), 0.0f, 1.0f);
  auto s  = (cb.x < 0.0f && ca.y < 0.0f) ? -1.0f : 1.0f;
  return s * vsqrt(vmin(dot(ca, ca), dot(cb, cb)));
}

struct sdf_fn {
  frame3 frame;
  int    type, material;
  f4     params;
  f3     whd;
};
VPT_FN sdf_fn ld_sdf(const DScene& sc, int id) {
  auto p  = sc.sdfs + 10 * id;
  auto c0 = ld4(p), c1 = ld4(p + 1), c2 = ld4(p + 2), c3 = ld4(p + 3);
  auto s  = sdf_fn{};
  s.frame = {xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
  s.type = f2i(c0.w), s.material = f2i(c1.w);
  s.params = ld4(p + 8);
  s.whd    = xyz(ld4(p + 9));
  return s;
}
// the closures of yocto_sceneio.cpp:3690-3726 as a typed switch; q is local
VPT_FN float eval_sdf_fn(const sdf_fn& s, f3 q) {
  VPT_WORK(sdf_fn_evals, 1);
  switch (s.type) {
    case sdf_bbox: return sd_bbox(q, f3{s.params.y, s.params.z, s.params.w}, s.params.x);
    case sdf_box: return sd_box(q - (s.whd * 0.5f), s.whd * 0.5f);
    case sdf_capped_cone: return sd_capped_cone(q, s.params.x, s.params.y, s.params.z);
    case sdf_plane: return q.y;
    case sdf_sphere: return length(q) - s.params.x;
    default: return sd_torus(q, s.params.x, s.params.y);
  }
}

constexpr int vol_inst_f4 = 6;  // f4 per vol_instances record (dscene.cuh)
struct vol_inst {
  frame3 frame;
  float  scalef;
  int    volume, material;
  f3     half, size;  // of the proxy box, see proxy_box_size
};
// bbox_size of eval_sdf (yocto_sdfs.cpp:36-40), float for float: `(o + res whd scalef) - o`
// keeps the reference's round trip through the frame origin. Constant per instance, so the
// uploader evaluates it once (same expressions, no FMA contraction on either side).
VPT_FN f3 proxy_box_size(f3 origin, int w, int h, int d, float res, float scalef) {
  auto grid_res = f3{(float)w, (float)h, (float)d};
  auto bbox_max = origin + (res * grid_res) * scalef;
  return bbox_max - origin;
}
VPT_FN vol_inst ld_vol_instance(const DScene& sc, int id) {
  auto p  = sc.vol_instances + vol_inst_f4 * id;
  auto c0 = ld4(p), c1 = ld4(p + 1), c2 = ld4(p + 2), c3 = ld4(p + 3);
  return {{xyz(c0), xyz(c1), xyz(c2), xyz(c3)}, c0.w, f2i(c1.w), f2i(c2.w), xyz(ld4(p + 4)),
      xyz(ld4(p + 5))};
}

// eval_volume (yocto_sdfs.cpp:92-127): clamped trilinear, 8 scalar loads
VPT_FN float eval_volume(const DScene& sc, int w, int h, int d, int offset, f3 uvw) {
  VPT_WORK(trilinear, 1);
  if (w * h * d == 0) return 0;
  auto s = vclamp((uvw.x + 1.0f) * 0.5f, 0.0f, 1.0f) * (w - 1);
  auto t = vclamp((uvw.y + 1.0f) * 0.5f, 0.0f, 1.0f) * (h - 1);
  auto r = vclamp((uvw.z + 1.0f) * 0.5f, 0.0f, 1.0f) * (d - 1);
  auto i = vclamp((int)s, 0, w - 1), j = vclamp((int)t, 0, h - 1), k = vclamp((int)r, 0, d - 1);
  auto ii = vmin(i + 1, w - 1), jj = vmin(j + 1, h - 1), kk = vmin(k + 1, d - 1);
  auto u = s - i, v = t - j, ww = r - k;
  auto at = [&](int x, int y, int z) {
    return ld1(sc.voxels + offset + x + y * w + z * w * h);
  };
  return at(i, j, k) * (1 - u) * (1 - v) * (1 - ww) + at(ii, j, k) * u * (1 - v) * (1 - ww) +
         at(i, jj, k) * (1 - u) * v * (1 - ww) + at(i, j, kk) * (1 - u) * (1 - v) * ww +
         at(i, jj, kk) * (1 - u) * v * ww + at(ii, j, kk) * u * (1 - v) * ww +
         at(ii, jj, k) * u * v * (1 - ww) + at(ii, jj, kk) * u * v * ww;
}

// eval_sdf(volume, instance, p, t) (yocto_sdfs.cpp:30-49); p is already local
VPT_FN float eval_sdf_grid(const DScene& sc, const vol_inst& in, f3 p, float t) {
  VPT_WORK(sdf_grid_evals, 1);
  auto bbox_dist = sd_box(p - in.half, in.half);
  if (bbox_dist < flt_eps * t) {
    auto v0 = ld4(sc.volumes + 2 * in.volume), v1 = ld4(sc.volumes + 2 * in.volume + 1);
    int  w = f2i(v0.x), h = f2i(v0.y), d = f2i(v0.z);
    auto uvw = p * 2.f / (in.size)-1;
    return eval_volume(sc, w, h, d, f2i(v1.x), uvw) * in.scalef;
  }
  return bbox_dist;
}

struct sdf_result {
  float result;
  int   instance, sdf;
};

// eval_sdf_scene (yocto_sdfs.cpp:7-26): linear scan, grids first
VPT_FN sdf_result eval_sdf_scene(const DScene& sc, f3 p, float t) {
  auto res = sdf_result{flt_max, invalidid, invalidid};
  VPT_WORK(sdf_scene_evals, 1);
  for (auto idx = 0; idx < sc.num_vol_instances; idx++) {
    auto in = ld_vol_instance(sc, idx);
    auto v  = eval_sdf_grid(sc, in, transform_point(in.frame, p), t);
    if (v < res.result) res = {v, idx, invalidid};
  }
  for (auto idx = 0; idx < sc.num_sdfs; idx++) {
    auto s = ld_sdf(sc, idx);
    auto v = eval_sdf_fn(s, transform_point(s.frame, p));
    if (v < res.result) res = {v, invalidid, idx};
  }
  return res;
}

// The same minimum without evaluating distance fields that cannot be it.
//
// The reference's scan returns (min v_k, the FIRST k that attains it). Any evaluation order
// gives that result if the running best is updated lexicographically (smaller v, or equal v
// and smaller k), and a field may be left out altogether if its value is known to be
// STRICTLY above the running best: it can then neither be the minimum nor tie with it.
// dscene.cuh `sdf_bounds` gives such a bound per field k: a world-space box W_k that contains
// the pre-image of the field's local bounding box, and s_k, a lower bound of the smallest
// singular value of the frame's linear part, so that for a point p outside W_k
//     v_k(p) >= dist_local(frame_k p, box_k) >= s_k dist(p, W_k)
// (all six analytic fields are exact distances to a shape inside box_k; a grid instance returns
// the exact distance to its proxy box when that distance is >= flt_eps t, which the grid test
// below also demands, and its voxels otherwise). W_k is enlarged by 1e-4 of its scale, s_k
// shrunk by 1e-3 and the comparison carries a 1 % margin on the squares: four orders of
// magnitude above the float error of either side. Planes and singular frames have s_k = 0
// (never skipped). tests/hostsim --sdf-cull-check evaluates every skipped field anyway.
// `seed`: field to evaluate first (the previous march step's nearest one), -1 for none;
// fields are numbered grids first, as the reference scans them.
// MEASURED ON B200 AND LEFT OFF (VPT_SDF_CULL: 0 = the plain scan, default; 1 = bounds in scan
// order; 2 = seeded): a lane skips 3.4 of 6 fields per evaluation (hostsim), but a warp walks a
// field's code whenever ANY of its lanes needs it, and the plain scan is warp-uniform (same
// field, same type for all 32 lanes) while the seed makes lanes start on different types:
// 07_sdfunction 1920x800 173.6 (plain) / 143.8 (1) / 119.2 (2) Mpaths/s, 06_gridsdf 117.9 /
// 98.7 / 90.3. Kept as a checked restatement of the exactness argument (CPU suite).
#if !defined(VPT_SDF_CULL)
#define VPT_SDF_CULL 0
#endif
#if defined(VPT_SDF_CULL_CHECK)
inline bool vpt_sdf_cull_check_on = false;
inline int  vpt_sdf_cull_mode     = VPT_SDF_CULL;
#endif
VPT_FN float eval_sdf_field(const DScene& sc, int k, f3 p, float t) {
  if (k < sc.num_vol_instances) {
    auto in = ld_vol_instance(sc, k);
    return eval_sdf_grid(sc, in, transform_point(in.frame, p), t);
  }
  auto s = ld_sdf(sc, k - sc.num_vol_instances);
  return eval_sdf_fn(s, transform_point(s.frame, p));
}
VPT_FN sdf_result eval_sdf_scene_culled(const DScene& sc, f3 p, float t, int seed, int& nearest) {
  VPT_WORK(sdf_scene_evals, 1);
  int   n     = sc.num_vol_instances + sc.num_sdfs;
  float best  = flt_max;
  int   bestk = -1;
  if (seed >= 0) {
    best  = eval_sdf_field(sc, seed, p, t);
    bestk = seed;
    // the reference's scan only ever takes a value that is < flt_max (never NaN / inf)
    if (!(best < flt_max)) best = flt_max, bestk = -1;
  }
  for (auto k = 0; k < n; k++) {
    if (k == seed) continue;
    auto b0 = ld4(sc.sdf_bounds + 2 * k), b1 = ld4(sc.sdf_bounds + 2 * k + 1);
    auto dx = vmax(vmax(b0.x - p.x, p.x - b1.x), 0.0f);
    auto dy = vmax(vmax(b0.y - p.y, p.y - b1.y), 0.0f);
    auto dz = vmax(vmax(b0.z - p.z, p.z - b1.z), 0.0f);
    auto lb2  = (dx * dx + dy * dy + dz * dz) * b0.w;  // (s_k dist(p, W_k))^2, s_k^2 in b0.w
    auto et   = flt_eps * t;
    bool skip = best > 0 && lb2 > best * best * 1.01f && (b1.w == 0 || lb2 > et * et * 1.01f);
    if (skip) {
      VPT_WORK(sdf_skips, 1);
#if defined(VPT_SDF_CULL_CHECK)
      if (vpt_sdf_cull_check_on && !(eval_sdf_field(sc, k, p, t) > best)) VPT_WORK(sdf_skips_wrong, 1);
#endif
      continue;
    }
    auto v = eval_sdf_field(sc, k, p, t);
    if (v < best || (v == best && k < bestk)) best = v, bestk = k;
  }
  nearest = bestk;
  if (bestk < 0 || !(best < flt_max)) {
    // the reference's scan keeps (flt_max, none) unless some field is < flt_max
    nearest = -1;
    return {flt_max, invalidid, invalidid};
  }
  if (bestk < sc.num_vol_instances) return {best, bestk, invalidid};
  return {best, invalidid, bestk - sc.num_vol_instances};
}

struct spheretrace_result {
  bool  hit;
  float dist;
  int   instance, sdf;
};

// spheretrace over the whole scene (yocto_pathtrace.cpp:289-307) as resumable state: the
// persistent extend kernel advances every lane a few steps per round and refills the lanes
// whose march is over (trip counts run from 1 to maxiter, so a fixed ray-per-lane launch
// leaves most lanes idle: ncu r01 10 of 32).
struct sphere_t {
  float              t;
  int                i;
  int                nearest;  // the field that was nearest at the previous step (-1: none)
  bool               finished;
  spheretrace_result r;
  VPT_FN void begin(const ray3& ray) {
    t = ray.tmin, i = 0, finished = false, nearest = -1;
    r = {false, flt_max, invalidid, invalidid};
  }
  VPT_FN bool done() const { return finished; }
  VPT_FN void step(const DScene& sc, const ray3& ray, int maxiter) {
    if (i < maxiter && t < ray.tmax) {
      auto p   = ray_point(ray, t);
#if defined(VPT_SDF_CULL_CHECK)
      const int mode = vpt_sdf_cull_mode;  // tests/hostsim picks it at run time
#else
      constexpr int mode = VPT_SDF_CULL;
#endif
      sdf_result res;
      if (mode == 2)
        res = eval_sdf_scene_culled(sc, p, t, nearest, nearest);
      else if (mode == 1)
        res = eval_sdf_scene_culled(sc, p, t, -1, nearest);
      else
        res = eval_sdf_scene(sc, p, t);
      if (vabs(res.result) < (flt_eps * t)) {
        r        = {true, t, res.instance, res.sdf};
        finished = true;
      } else {
        t += res.result;
        ++i;
      }
    } else {
      finished = true;
    }
  }
};
VPT_FN spheretrace_result spheretrace(const DScene& sc, const ray3& ray, int maxiter) {
  sphere_t s;
  s.begin(ray);
  while (!s.done()) s.step(sc, ray, maxiter);
  return s.r;
}

// spheretrace against one analytic sdf (yocto_pathtrace.cpp:267-286), light pdf
VPT_FN spheretrace_result spheretrace_one(
    const DScene& sc, const ray3& ray, int sdf_handle, int maxiter) {
  auto t = ray.tmin;
  auto s = ld_sdf(sc, sdf_handle);
  for (auto i = 0; i < maxiter && t < ray.tmax; ++i) {
    auto p   = ray_point(ray, t);
    auto res = eval_sdf_fn(s, transform_point(s.frame, p));
    if (vabs(res) < (flt_eps * t)) return {true, t, invalidid, sdf_handle};
    t += res;
  }
  return {false, flt_max, invalidid, invalidid};
}

// tetrahedron finite differences (yocto_sdfs.cpp:67-89)
VPT_FN f3 eval_sdf_normal_fn(const sdf_fn& s, f3 p, float t) {
  auto h  = flt_eps * t;
  auto p1 = transform_point(s.frame, p + f3{1, -1, -1} * h);
  auto p2 = transform_point(s.frame, p + f3{-1, -1, 1} * h);
  auto p3 = transform_point(s.frame, p + f3{-1, 1, -1} * h);
  auto p4 = transform_point(s.frame, p + f3{1, 1, 1} * h);
  return normalize(f3{1, -1, -1} * eval_sdf_fn(s, p1) + f3{-1, -1, 1} * eval_sdf_fn(s, p2) +
                   f3{-1, 1, -1} * eval_sdf_fn(s, p3) + f3{1, 1, 1} * eval_sdf_fn(s, p4));
}
VPT_FN f3 eval_sdf_normal_grid(const DScene& sc, const vol_inst& in, f3 p, float t) {
  auto h  = flt_eps * t;
  auto p1 = transform_point(in.frame, p + f3{1, -1, -1} * h);
  auto p2 = transform_point(in.frame, p + f3{-1, -1, 1} * h);
  auto p3 = transform_point(in.frame, p + f3{-1, 1, -1} * h);
  auto p4 = transform_point(in.frame, p + f3{1, 1, 1} * h);
  return normalize(
      f3{1, -1, -1} * eval_sdf_grid(sc, in, p1, t) + f3{-1, -1, 1} * eval_sdf_grid(sc, in, p2, t) +
      f3{-1, 1, -1} * eval_sdf_grid(sc, in, p3, t) + f3{1, 1, 1} * eval_sdf_grid(sc, in, p4, t));
}

}  // namespace vpt
