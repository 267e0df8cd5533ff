// shaders.cuh — the nine shaders of the reference (yocto_pathtrace.cpp:425-930)
// written as a bounce state machine so that the megakernel (one thread runs a
// whole path) and the wavefront pipeline (extend / shade / light-pdf queues)
// execute exactly the same arithmetic in the same order:
//
//   find_hit        closest hit (BVH) or sphere-trace, per shader family
//   shade_hit       everything from the hit to the choice of the next direction;
//                   returns step_lightpdf when the MIS weight still needs
//                   sample_lights_pdf for the chosen direction
//   apply_lightpdf  weight *= f / (0.5 pdf_bsdf + 0.5 pdf_light)
//   bounce_tail     weight check + russian roulette + bounce increment
//
// Random-number draw order matters for stream replay. The reference is built
// with gcc, which evaluates call arguments right to left, so e.g.
// sample_bsdfcos(..., rand1f(rng), rand2f(rng)) draws rn first, then rnl
// (SURVEY.md section 8(a), Appendix B.6). That order is hard-coded here.
#pragma once

#include "lights.cuh"
#include "rng.cuh"

namespace vpt {

enum : int {
  shader_volpathtrace = 0, shader_pathtrace, shader_naive, shader_eyelight, shader_normal,
  shader_texcoord, shader_color, shader_implicit, shader_implicit_normal
};

enum step_t : int {
  step_done = 0,     // path ended inside shade_hit
  step_skip = 1,     // opacity pass-through: new ray, bounce not consumed, no tail
  step_continue = 2, // new ray set, weight final: run bounce_tail
  step_lightpdf = 3  // new ray set, weight waits for the light pdf, then bounce_tail
};

struct path_t {
  f3    o, d;              // next ray (tmin / tmax are always the defaults)
  f3    radiance, weight;
  int   bounce;
  bool  hit;
  bool  in_volume;         // vstack depth (0 or 1, yocto_pathtrace.cpp:642-648)
  f3    vol_density, vol_scattering, vol_emission;
  float vol_anisotropy;
};

struct pending_t {  // MIS terms waiting for sample_lights_pdf
  f3    position, incoming, f;
  float bpdf;
};

// what find_hit returns: a BVH hit or a sphere-trace hit
struct isect_t {
  bool  hit;
  int   instance, element;  // BVH: instance / element; sdf: vol_instance / sdf handle
  f2    uv;
  float distance;
};

VPT_FN void path_init(path_t& p, const ray3& ray) {
  p.o = ray.o, p.d = ray.d;
  p.radiance = {0, 0, 0}, p.weight = {1, 1, 1};
  p.bounce = 0, p.hit = false, p.in_volume = false;
  p.vol_density = p.vol_scattering = p.vol_emission = {0, 0, 0};
  p.vol_anisotropy = 0;
}

template <int S>
VPT_FN constexpr bool is_sdf_shader() {
  return S == shader_implicit || S == shader_implicit_normal;
}
template <int S>
VPT_FN int max_bounces(const DParams& prm) {
  return S == shader_eyelight ? vmax(prm.bounces, 4) : prm.bounces;
}

template <int S>
VPT_FN isect_t find_hit(const DScene& sc, const DParams& prm, const path_t& p, stack_t st) {
  auto ray = make_ray(p.o, p.d);
  if (is_sdf_shader<S>()) {
    auto r = spheretrace(sc, ray, prm.spheretrace_maxiter);
    return {r.hit, r.instance, r.sdf, {0, 0}, r.dist};
  } else {
    auto h = intersect_scene(sc, ray, st);
    return {h.hit, h.instance, h.element, h.uv, h.distance};
  }
}

// the scattering decision shared by pathtrace / volpathtrace / implicit surfaces
template <int S>
VPT_FN step_t scatter_surface(const DScene& sc, const DParams& prm, path_t& p, rng_t& rng,
    const material_point& material, f3 position, f3 normal, f3 outgoing, f3& incoming,
    pending_t& pd) {
  incoming = {0, 0, 0};
  if (!is_delta(material)) {
    auto pbsdf = (S == shader_implicit && prm.noimplicit_mis) ? 1.0f : 0.5f;
    if (rand1f(rng) < pbsdf) {
      auto rn  = rand2f(rng);
      auto rnl = rand1f(rng);
      incoming = sample_bsdfcos(material, normal, outgoing, rnl, rn);
    } else {
      auto ruv = rand2f(rng);
      auto rel = rand1f(rng);
      auto rl  = rand1f(rng);
      incoming = sample_lights(sc, position, rl, rel, ruv);
    }
    if (is_zero(incoming)) return step_done;
    auto f    = eval_bsdfcos(material, normal, outgoing, incoming);
    auto bpdf = sample_bsdfcos_pdf(material, normal, outgoing, incoming);
    if (S == shader_implicit && prm.noimplicit_mis) {
      p.weight *= f / bpdf;
      return step_continue;
    }
    pd = {position, incoming, f, bpdf};
    return step_lightpdf;
  } else {
    incoming = sample_delta(material, normal, outgoing, rand1f(rng));
    p.weight *= eval_delta(material, normal, outgoing, incoming) /
                sample_delta_pdf(material, normal, outgoing, incoming);
    return step_continue;
  }
}

// Everything between the intersection and the next ray for one bounce.
template <int S>
VPT_FN step_t shade_hit(const DScene& sc, const DParams& prm, path_t& p, rng_t& rng,
    isect_t isec, pending_t& pd) {
  auto ray_d = p.d;
  if (!isec.hit) {
    // the debug shaders return (0,0,0,0) on a miss, not the environment
    if (S == shader_normal || S == shader_texcoord || S == shader_color) return step_done;
    p.radiance += p.weight * eval_environment(sc, ray_d);
    return step_done;
  }

  if (S == shader_implicit) {
    // shade_implicit (yocto_pathtrace.cpp:425-535)
    auto outgoing = -ray_d;
    auto position = p.o + ray_d * isec.distance;
    f3   normal;
    int  material_handle;
    if (isec.instance != invalidid) {
      auto in         = ld_vol_instance(sc, isec.instance);
      normal          = eval_sdf_normal_grid(sc, in, position, isec.distance);
      material_handle = in.material;
    } else {
      auto s          = ld_sdf(sc, isec.element);
      normal          = eval_sdf_normal_fn(s, position, isec.distance);
      material_handle = s.material;
    }
    auto material = eval_material_plain(ld_material(sc, material_handle));
    if (material.opacity < 1 && rand1f(rng) >= material.opacity) {
      p.o = position + ray_d * 1e-2f;
      return step_skip;
    }
    if (p.bounce == 0) p.hit = true;
    p.radiance += p.weight * eval_emission(material, normal, outgoing);
    f3   incoming;
    auto r = scatter_surface<S>(sc, prm, p, rng, material, position, normal, outgoing, incoming,
        pd);
    if (r == step_done) return r;
    p.o = position, p.d = incoming;
    return r;
  }

  // mesh shaders ---------------------------------------------------------------
  auto in = ld_instance(sc, isec.instance);
  auto md = ld_material(sc, in.material);

  if (S == shader_volpathtrace) {
    // transmittance sampling inside a volume (yocto_pathtrace.cpp:585-596)
    auto in_medium = false;
    if (p.in_volume) {
      auto rd       = rand1f(rng);
      auto rl       = rand1f(rng);
      auto distance = sample_transmittance(p.vol_density, isec.distance, rl, rd);
      p.weight *= eval_transmittance(p.vol_density, distance) /
                  sample_transmittance_pdf(p.vol_density, distance, isec.distance);
      in_medium     = distance < isec.distance;
      isec.distance = distance;
    }
    if (in_medium) {
      // volume event (yocto_pathtrace.cpp:654-673)
      auto outgoing = -ray_d;
      auto position = p.o + ray_d * isec.distance;
      // quirk kept: the position is passed where eval_emission expects a normal
      p.radiance += p.weight * (dot(position, outgoing) >= 0 ? p.vol_emission : f3{0, 0, 0});
      f3 incoming;
      if (rand1f(rng) < 0.5f) {
        auto rn = rand2f(rng);
        (void)rand1f(rng);  // rnl: drawn, unused by sample_scattering
        incoming = sample_phasefunction(p.vol_anisotropy, outgoing, rn);
      } else {
        auto ruv = rand2f(rng);
        auto rel = rand1f(rng);
        auto rl  = rand1f(rng);
        incoming = sample_lights(sc, position, rl, rel, ruv);
      }
      auto f = p.vol_density * p.vol_scattering *
               eval_phasefunction(p.vol_anisotropy, incoming, outgoing);
      auto bpdf = eval_phasefunction(p.vol_anisotropy, outgoing, incoming);
      pd        = {position, incoming, f, bpdf};
      p.o = position, p.d = incoming;
      return step_lightpdf;
    }
  }

  auto outgoing = -ray_d;
  auto position = eval_position(sc, in, isec.element, isec.uv);

  if (S == shader_normal) {
    auto normal = eval_shading_normal(sc, in, md, isec.element, isec.uv, outgoing);
    p.radiance  = normal;
    p.hit       = true;
    return step_done;
  }
  if (S == shader_texcoord) {
    auto tc    = eval_texcoord(sc, in, isec.element, isec.uv);
    p.radiance = {tc.x, tc.y, 0};
    p.hit      = true;
    return step_done;
  }
  if (S == shader_color) {
    p.radiance = eval_material(sc, in, md, isec.element, isec.uv).color;
    p.hit      = true;
    return step_done;
  }

  auto normal   = eval_shading_normal(sc, in, md, isec.element, isec.uv, outgoing);
  auto material = eval_material(sc, in, md, isec.element, isec.uv);

  // opacity pass-through does not consume a bounce
  if (material.opacity < 1 && rand1f(rng) >= material.opacity) {
    p.o = position + ray_d * 1e-2f;
    return step_skip;
  }
  if (p.bounce == 0) p.hit = true;
  p.radiance += p.weight * eval_emission(material, normal, outgoing);

  if (S == shader_eyelight) {
    // yocto_pathtrace.cpp:869-886
    p.radiance += p.weight * pif * eval_bsdfcos(material, normal, outgoing, outgoing);
    if (!is_delta(material)) return step_done;
    auto incoming = sample_delta(material, normal, outgoing, rand1f(rng));
    if (is_zero(incoming)) return step_done;
    p.weight *= eval_delta(material, normal, outgoing, incoming) /
                sample_delta_pdf(material, normal, outgoing, incoming);
    p.o = position, p.d = incoming;
    return step_continue;
  }

  if (S == shader_naive) {
    // yocto_pathtrace.cpp:802-815
    f3 incoming;
    if (material.roughness != 0) {
      auto rn  = rand2f(rng);
      auto rnl = rand1f(rng);
      incoming = sample_bsdfcos(material, normal, outgoing, rnl, rn);
      if (is_zero(incoming)) return step_done;
      p.weight *= eval_bsdfcos(material, normal, outgoing, incoming) /
                  sample_bsdfcos_pdf(material, normal, outgoing, incoming);
    } else {
      incoming = sample_delta(material, normal, outgoing, rand1f(rng));
      if (is_zero(incoming)) return step_done;
      p.weight *= eval_delta(material, normal, outgoing, incoming) /
                  sample_delta_pdf(material, normal, outgoing, incoming);
    }
    p.o = position, p.d = incoming;
    return step_continue;
  }

  // pathtrace / volpathtrace
  f3   incoming;
  auto r = scatter_surface<S>(sc, prm, p, rng, material, position, normal, outgoing, incoming, pd);
  if (r == step_done) return r;
  if (S == shader_volpathtrace) {
    // vstack update (yocto_pathtrace.cpp:641-648)
    if (is_volumetric_type(md.type) && dot(normal, outgoing) * dot(normal, incoming) < 0) {
      if (!p.in_volume) {
        p.in_volume      = true;
        p.vol_density    = material.density;
        p.vol_scattering = material.scattering;
        p.vol_emission   = material.emission;
        p.vol_anisotropy = material.scanisotropy;
      } else {
        p.in_volume = false;
      }
    }
  }
  p.o = position, p.d = incoming;
  return r;
}

VPT_FN void apply_lightpdf(path_t& p, const pending_t& pd, float lpdf) {
  p.weight *= pd.f / (0.5f * pd.bpdf + 0.5f * lpdf);
}

// weight check, russian roulette, bounce increment; false when the path ends
template <int S>
VPT_FN bool bounce_tail(const DParams& prm, path_t& p, rng_t& rng) {
  if (is_zero(p.weight) || !vfinite(p.weight)) return false;
  if (S != shader_eyelight) {
    if (p.bounce > 3) {
      auto rr_prob = vmin((float)0.99, max3(p.weight));
      if (rand1f(rng) >= rr_prob) return false;
      p.weight *= 1 / rr_prob;
    }
  }
  p.bounce += 1;
  return p.bounce < max_bounces<S>(prm);
}

// shade_implicit_normal (yocto_pathtrace.cpp:538-562)
VPT_FN f4 shade_implicit_normal(const DScene& sc, const DParams& prm, const ray3& ray) {
  auto r = spheretrace(sc, ray, prm.spheretrace_maxiter);
  if (!r.hit) return {0, 0, 0, 0};
  auto position = ray_point(ray, r.dist);
  f3   normal;
  if (r.instance != invalidid) {
    normal = eval_sdf_normal_grid(sc, ld_vol_instance(sc, r.instance), position, r.dist);
  } else {
    normal = eval_sdf_normal_fn(ld_sdf(sc, r.sdf), position, r.dist);
  }
  normal = normal * 0.5f + 0.5f;
  return {normal.x, normal.y, normal.z, 1};
}

VPT_FN f4 path_result(int shader, const path_t& p) {
  auto alpha = (shader == shader_implicit) ? 1.0f : (p.hit ? 1.0f : 0.0f);
  return {p.radiance.x, p.radiance.y, p.radiance.z, alpha};
}

// Whole path in one thread (the megakernel body).
template <int S>
VPT_FN f4 shade_path(const DScene& sc, const DParams& prm, const ray3& ray, rng_t& rng,
    stack_t st) {
  if (S == shader_implicit_normal) return shade_implicit_normal(sc, prm, ray);
  path_t p;
  path_init(p, ray);
  // single-exit loop: `alive` is the only way out
  auto alive = p.bounce < max_bounces<S>(prm);
  VPT_GUARD_DECL(guard);
  while (alive) {
    VPT_GUARD_STEP(3, guard, 1000);
    VPT_WORK(bounces, 1);
    auto      isec = find_hit<S>(sc, prm, p, st);
    pending_t pd;
    auto      r = shade_hit<S>(sc, prm, p, rng, isec, pd);
    if (r == step_lightpdf) {
      auto lpdf = sample_lights_pdf(sc, pd.position, pd.incoming, prm.spheretrace_maxiter, st);
      apply_lightpdf(p, pd, lpdf);
    }
    if (r == step_done) {
      alive = false;
    } else if (r != step_skip) {
      alive = bounce_tail<S>(prm, p, rng);
    }
  }
  return path_result(S, p);
}

// One camera sample: jitter, lens, camera ray (yocto_pathtrace.cpp:1059-1089)
VPT_FN ray3 sample_camera(const DScene& sc, const DParams& prm, int idx, rng_t& rng) {
  auto i = idx % prm.width, j = idx / prm.width;
  f2   uv;
  if (prm.samples == 1) {
    uv = {(i + 0.5f) / prm.width, (j + 0.5f) / prm.height};
  } else {
    auto u = (i + rand1f(rng)) / prm.width;
    auto v = (j + rand1f(rng)) / prm.height;
    uv     = {u, v};
  }
  return eval_camera(sc, prm.camera, uv, rand2f(rng));
}

// isfinite guard: a non-finite sample becomes (0,0,0,0) (yocto_pathtrace.cpp:1087;
// the 3-component brace-init zeroes alpha too)
VPT_FN f4 guard_radiance(f4 r) {
  if (!(vfinite(r.x) && vfinite(r.y) && vfinite(r.z) && vfinite(r.w))) return {0, 0, 0, 0};
  return r;
}

}  // namespace vpt
