// lights.cuh — light direction sampling and its solid-angle pdf.
// Restates sample_lights / sample_lights_pdf (yocto_pathtrace.cpp:312-421) and
// the discrete-CDF helpers of yocto_sampling.h:368-395. There is no shadow ray
// in this renderer: sampling only picks a direction; the pdf re-intersects each
// mesh light's instance up to 100 times along that direction (instance-targeted
// traversals), marches the analytic sdf of each sdf light, and looks up the
// environment texel probability.
#pragma once

#include "bsdf.cuh"
#include "bvh.cuh"
#include "sdf.cuh"

namespace vpt {

// sample_discrete: clamp then std::upper_bound (same bisection as libstdc++).
// `sorted` (lights record, set by the uploader / k_light_sorted when the running sums never
// decrease): the upper bound of a sorted array does not depend on the probe order, so the
// device first narrows the range 16-fold per step with 15 INDEPENDENT loads (one memory
// latency per step instead of four; ncu r02c: the bisection's 17 dependent loads over the
// 131 k-texel environment CDF were 12 % of k_wf_shade's stall samples at 5 lanes), then bisects.
// An array that is not sorted keeps the reference's probe sequence.
VPT_FN int sample_discrete(const float* cdf, int n, float r, bool sorted = false) {
  auto last = ld1(cdf + n - 1);
  r         = vclamp(r * last, (float)0, last - (float)0.00001);
  int first = 0, len = n;
  VPT_GUARD_DECL(guard);
#if defined(__CUDA_ARCH__)
  while (sorted && len > 32) {
    int step = len >> 4;  // >= 2; pivots p_k = first + k step, k = 1 .. 15
    int c    = 0;         // pivots with cdf[p_k] <= r: exactly p_1 .. p_c
#pragma unroll
    for (int k = 1; k < 16; k++) c += r < ld1(cdf + first + k * step) ? 0 : 1;
    int lo = c > 0 ? first + c * step + 1 : first;
    int hi = c < 15 ? first + (c + 1) * step : first + len;  // the bound is in [lo, hi]
    first = lo, len = hi - lo;
  }
#endif
  while (len > 0) {
    VPT_GUARD_STEP(4, guard, 64);
    auto half = len >> 1;
    auto mid  = first + half;
    if (r < ld1(cdf + mid)) {
      len = half;
    } else {
      first = mid + 1;
      len   = len - half - 1;
    }
  }
  return vclamp(first, 0, n - 1);
}
VPT_FN float sample_discrete_pdf(const float* cdf, int idx) {
  if (idx == 0) return ld1(cdf);
  return ld1(cdf + idx) - ld1(cdf + idx - 1);
}
VPT_FN int sample_uniform(int size, float r) { return vclamp((int)(r * size), 0, size - 1); }

// sample_lights (yocto_pathtrace.cpp:312-350)
VPT_CALL f3 sample_lights(const DScene& sc, f3 position, float rl, float rel, f2 ruv) {
  // no emitters: the reference indexes an empty vector here (UB on the host); on the device an
  // out-of-range read can poison the context, so the direction is zero, which ends the path
  // exactly as `incoming == 0` does (yocto_pathtrace.cpp:736)
  if (sc.num_lights <= 0) return {0, 0, 0};
  auto light_id = sample_uniform(sc.num_lights, rl);
  auto l0 = ld4(sc.lights + 2 * light_id), l1 = ld4(sc.lights + 2 * light_id + 1);
  if (l0.x != invalidid) {
    auto in      = ld_instance(sc, l0.x);
    auto element = sample_discrete(sc.cdf + l0.w, l1.x, rel, l1.y != 0);
    auto uv      = ruv;
    if (in.s.kind == kind_triangles) {
      auto sq = vsqrt(ruv.x);
      uv      = {1 - sq, ruv.y * sq};
    }
    auto lposition = eval_position(sc, in, element, uv);
    return normalize(lposition - position);
  } else if (l0.z != invalidid) {
    auto p       = sc.sdfs + 10 * l0.z;
    auto inv     = ld_frame(p + 4);  // inverse(sdf.frame), rigid, host-precomputed
    auto whd     = xyz(ld4(p + 9));
    auto wlightp = transform_point(inv, f3{ruv.x, ruv.y, 1} * whd);
    return normalize(wlightp - position);
  } else if (l0.y != invalidid) {
    auto e  = sc.environments + 9 * l0.y;
    auto em = ld4(e + 8);
    if (f2i(em.w) != invalidid) {
      auto tex = ld4(sc.textures + f2i(em.w));
      auto idx = sample_discrete(sc.cdf + l0.w, l1.x, rel, l1.y != 0);
      auto uv  = f2{((idx % tex.x) + 0.5f) / tex.x, ((idx / tex.x) + 0.5f) / tex.y};
      return transform_direction(ld_frame(e),
          f3{vcos(uv.x * 2 * pif) * vsin(uv.y * pif), vcos(uv.y * pif),
              vsin(uv.x * 2 * pif) * vsin(uv.y * pif)});
    } else {
      return sample_sphere(ruv);
    }
  }
  return {0, 0, 0};
}

// sample_lights_pdf (yocto_pathtrace.cpp:353-421)
VPT_FN float sample_lights_pdf(const DScene& sc, f3 position, f3 direction,
    int spheretrace_maxiter, stack_t st, unsigned long long* ray_counter = nullptr) {
  auto pdf = 0.0f;
  VPT_WORK(lightpdf_calls, 1);
  if (sc.num_lights <= 0) return 0;  // see sample_lights
  for (auto li = 0; li < sc.num_lights; li++) {
    auto l0 = ld4(sc.lights + 2 * li), l1 = ld4(sc.lights + 2 * li + 1);
    if (l0.x != invalidid) {
      auto in            = ld_instance(sc, l0.x);
      auto area          = ld1(sc.cdf + l0.w + l1.x - 1);
      auto lpdf          = 0.0f;
      auto next_position = position;
      for (auto bounce = 0; bounce < 100; bounce++) {
        auto isec = intersect_instance(sc, l0.x, make_ray(next_position, direction), st);
        if (ray_counter) (*ray_counter)++;
        if (!isec.hit) break;
        auto lposition = eval_position(sc, in, isec.element, isec.uv);
        auto lnormal   = eval_element_normal(sc, in, isec.element);
        lpdf += distance_squared(lposition, position) / (vabs(dot(lnormal, direction)) * area);
        next_position = lposition + direction * 1e-3f;
        VPT_COUNT_ADD(6, 1);
        VPT_COUNT_MAX(9, bounce);
      }
      pdf += lpdf;
    } else if (l0.z != invalidid) {
      auto ray  = make_ray(position, direction);
      auto isec = spheretrace_one(sc, ray, l0.z, spheretrace_maxiter);
      if (isec.hit) {
        auto lposition = ray_point(ray, isec.dist);
        // the reference evaluates the normal at the shading position (quirk kept)
        auto lnormal = eval_sdf_normal_fn(ld_sdf(sc, isec.sdf), position, isec.dist);
        auto area    = ld1(sc.cdf + l0.w + l1.x - 1);
        pdf += distance_squared(lposition, position) / (vabs(dot(lnormal, direction)) * area);
      }
    } else if (l0.y != invalidid) {
      auto e  = sc.environments + 9 * l0.y;
      auto em = ld4(e + 8);
      if (f2i(em.w) != invalidid) {
        auto tex      = ld4(sc.textures + f2i(em.w));
        auto wl       = transform_direction(ld_frame(e + 4), direction);
        auto texcoord = environment_texcoord(wl);
        auto i    = vclamp((int)(texcoord.x * tex.x), 0, tex.x - 1);
        auto j    = vclamp((int)(texcoord.y * tex.y), 0, tex.y - 1);
        auto prob = sample_discrete_pdf(sc.cdf + l0.w, j * tex.x + i) /
                    ld1(sc.cdf + l0.w + l1.x - 1);
        auto angle = (2 * pif / tex.x) * (pif / tex.y) * vsin(pif * (j + 0.5f) / tex.y);
        pdf += prob / angle;
      } else {
        pdf += 1 / (4 * pif);
      }
    }
  }
  pdf *= (float)1 / (float)sc.num_lights;
  return pdf;
}

}  // namespace vpt
