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

struct vol_inst {
  frame3 frame;
  float  scalef;
  int    volume, material;
};
VPT_FN vol_inst ld_vol_instance(const DScene& sc, int id) {
  auto p  = sc.vol_instances + 4 * id;
  auto c0 = ld4(p), c1 = ld4(p + 1), c2 = ld4(p + 2), c3 = ld4(p + 3);
  return {{xyz(c0), xyz(c1), xyz(c2), xyz(c3)}, c0.w, f2i(c1.w), f2i(c2.w)};
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
  auto v0 = ld4(sc.volumes + 2 * in.volume), v1 = ld4(sc.volumes + 2 * in.volume + 1);
  int  w = f2i(v0.x), h = f2i(v0.y), d = f2i(v0.z);
  auto res       = v0.w;
  auto grid_res  = f3{(float)w, (float)h, (float)d};
  auto origin    = in.frame.o;
  auto bbox_max  = origin + (res * grid_res) * in.scalef;
  auto bbox_size = (bbox_max - origin);
  auto bbox_dist = sd_box(p - (bbox_size * 0.5f), (bbox_size * 0.5f));
  if (bbox_dist < flt_eps * t) {
    auto uvw = p * 2.f / (bbox_size)-1;
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
  bool               finished;
  spheretrace_result r;
  VPT_FN void begin(const ray3& ray) {
    t = ray.tmin, i = 0, finished = false;
    r = {false, flt_max, invalidid, invalidid};
  }
  VPT_FN bool done() const { return finished; }
  VPT_FN void step(const DScene& sc, const ray3& ray, int maxiter) {
    if (i < maxiter && t < ray.tmax) {
      auto p   = ray_point(ray, t);
      auto res = eval_sdf_scene(sc, p, t);
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
