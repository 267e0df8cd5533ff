// bsdf.cuh — surface lobes, volume free-flight and phase-function math.
// Restates yocto_shading.h:296-1102 (Fresnel, GGX microfacets, the matte /
// glossy / reflective / transparent / refractive / gltfpbr / passthrough
// lobes in rough and delta form, homogeneous transmittance, Henyey-Greenstein)
// and the type dispatch of yocto_pathtrace.cpp:86-255, expression for
// expression, so a path fed the same random numbers takes the same branches.
#pragma once

#include "scene_eval.cuh"

namespace vpt {

// ---- hemisphere sampling (yocto_sampling.h:273-330) ---------------------------
VPT_FN f3 sample_hemisphere_cos(f3 normal, f2 ruv) {
  auto z     = vsqrt(ruv.y);
  auto r     = vsqrt(1 - z * z);
  auto phi   = 2 * pif * ruv.x;
  auto local = f3{r * vcos(phi), r * vsin(phi), z};
  return transform_direction(basis_fromz(normal), local);
}
VPT_FN float sample_hemisphere_cos_pdf(f3 normal, f3 direction) {
  auto cosw = dot(normal, direction);
  return (cosw <= 0) ? 0 : cosw / pif;
}
VPT_FN f3 sample_sphere(f2 ruv) {
  auto z   = 2 * ruv.y - 1;
  auto r   = vsqrt(vclamp(1 - z * z, 0.0f, 1.0f));
  auto phi = 2 * pif * ruv.x;
  return {r * vcos(phi), r * vsin(phi), z};
}

// ---- Fresnel (yocto_shading.h:296-360) ---------------------------------------
VPT_FN bool same_hemisphere(f3 normal, f3 outgoing, f3 incoming) {
  return dot(normal, outgoing) * dot(normal, incoming) >= 0;
}
VPT_FN f3 fresnel_schlick(f3 specular, f3 normal, f3 outgoing) {
  if (is_zero(specular)) return {0, 0, 0};
  auto cosine = dot(normal, outgoing);
  return specular + (1 - specular) * vpow(vclamp(1 - vabs(cosine), 0.0f, 1.0f), 5.0f);
}
VPT_FN float fresnel_dielectric(float eta, f3 normal, f3 outgoing) {
  auto cosw  = vabs(dot(normal, outgoing));
  auto sin2  = 1 - cosw * cosw;
  auto eta2  = eta * eta;
  auto cos2t = 1 - sin2 / eta2;
  if (cos2t < 0) return 1;
  auto t0 = vsqrt(cos2t);
  auto t1 = eta * t0;
  auto t2 = eta * cosw;
  auto rs = (cosw - t1) / (cosw + t1);
  auto rp = (t0 - t2) / (t0 + t2);
  return (rs * rs + rp * rp) / 2;
}
VPT_FN f3 fresnel_conductor(f3 eta, f3 etak, f3 normal, f3 outgoing) {
  auto cosw = dot(normal, outgoing);
  if (cosw <= 0) return {0, 0, 0};
  cosw       = vclamp(cosw, -1.0f, 1.0f);
  auto cos2  = cosw * cosw;
  auto sin2  = vclamp(1 - cos2, 0.0f, 1.0f);
  auto eta2  = eta * eta;
  auto etak2 = etak * etak;
  auto t0       = eta2 - etak2 - sin2;
  auto a2plusb2 = vsqrt(t0 * t0 + 4 * eta2 * etak2);
  auto t1       = a2plusb2 + cos2;
  auto a        = vsqrt((a2plusb2 + t0) / 2);
  auto t2       = 2 * a * cosw;
  auto rs       = (t1 - t2) / (t1 + t2);
  auto t3 = cos2 * a2plusb2 + sin2 * sin2;
  auto t4 = t2 * sin2;
  auto rp = rs * (t3 - t4) / (t3 + t4);
  return (rp + rs) / 2;
}
VPT_FN f3 eta_to_reflectivity(f3 eta) { return ((eta - 1) * (eta - 1)) / ((eta + 1) * (eta + 1)); }
VPT_FN f3 reflectivity_to_eta(f3 reflectivity_) {
  auto reflectivity = vclamp(reflectivity_, 0.0f, 0.99f);
  return (1 + vsqrt(reflectivity)) / (1 - vsqrt(reflectivity));
}

// ---- GGX microfacets (yocto_shading.h:402-470) --------------------------------
VPT_FN float microfacet_distribution(float roughness, f3 normal, f3 halfway) {
  auto cosine = dot(normal, halfway);
  if (cosine <= 0) return 0;
  auto roughness2 = roughness * roughness;
  auto cosine2    = cosine * cosine;
  return roughness2 / (pif * (cosine2 * roughness2 + 1 - cosine2) *
                          (cosine2 * roughness2 + 1 - cosine2));
}
VPT_FN float microfacet_shadowing1(float roughness, f3 normal, f3 halfway, f3 direction) {
  auto cosine  = dot(normal, direction);
  auto cosineh = dot(halfway, direction);
  if (cosine * cosineh <= 0) return 0;
  auto roughness2 = roughness * roughness;
  auto cosine2    = cosine * cosine;
  return 2 * vabs(cosine) / (vabs(cosine) + vsqrt(cosine2 - roughness2 * cosine2 + roughness2));
}
VPT_FN float microfacet_shadowing(
    float roughness, f3 normal, f3 halfway, f3 outgoing, f3 incoming) {
  return microfacet_shadowing1(roughness, normal, halfway, outgoing) *
         microfacet_shadowing1(roughness, normal, halfway, incoming);
}
VPT_FN f3 sample_microfacet(float roughness, f3 normal, f2 rn) {
  auto phi   = 2 * pif * rn.x;
  auto theta = vatan(roughness * vsqrt(rn.y / (1 - rn.y)));
  auto local = f3{vcos(phi) * vsin(theta), vsin(phi) * vsin(theta), vcos(theta)};
  return transform_direction(basis_fromz(normal), local);
}
VPT_FN float sample_microfacet_pdf(float roughness, f3 normal, f3 halfway) {
  auto cosine = dot(normal, halfway);
  if (cosine < 0) return 0;
  return microfacet_distribution(roughness, normal, halfway) * cosine;
}

VPT_FN f3 up_normal_of(f3 normal, f3 outgoing) {
  return dot(normal, outgoing) <= 0 ? -normal : normal;
}

// ---- matte --------------------------------------------------------------------
VPT_FN f3 eval_matte(f3 color, f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return {0, 0, 0};
  return color / pif * vabs(dot(normal, incoming));
}
VPT_FN f3 sample_matte(f3 normal, f3 outgoing, f2 rn) {
  return sample_hemisphere_cos(up_normal_of(normal, outgoing), rn);
}
VPT_FN float sample_matte_pdf(f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return 0;
  return sample_hemisphere_cos_pdf(up_normal_of(normal, outgoing), incoming);
}

// ---- glossy -------------------------------------------------------------------
VPT_FN f3 eval_glossy(f3 color, float ior, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return {0, 0, 0};
  auto up_normal = up_normal_of(normal, outgoing);
  auto F1        = fresnel_dielectric(ior, up_normal, outgoing);
  auto halfway   = normalize(incoming + outgoing);
  auto F         = fresnel_dielectric(ior, halfway, incoming);
  auto D         = microfacet_distribution(roughness, up_normal, halfway);
  auto G         = microfacet_shadowing(roughness, up_normal, halfway, outgoing, incoming);
  return color * (1 - F1) / pif * vabs(dot(up_normal, incoming)) +
         f3{1, 1, 1} * F * D * G / (4 * dot(up_normal, outgoing) * dot(up_normal, incoming)) *
             vabs(dot(up_normal, incoming));
}
VPT_FN f3 sample_glossy(float ior, float roughness, f3 normal, f3 outgoing, float rnl, f2 rn) {
  auto up_normal = up_normal_of(normal, outgoing);
  if (rnl < fresnel_dielectric(ior, up_normal, outgoing)) {
    auto halfway  = sample_microfacet(roughness, up_normal, rn);
    auto incoming = reflect(outgoing, halfway);
    if (!same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
    return incoming;
  } else {
    return sample_hemisphere_cos(up_normal, rn);
  }
}
VPT_FN float sample_glossy_pdf(float ior, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return 0;
  auto up_normal = up_normal_of(normal, outgoing);
  auto halfway   = normalize(outgoing + incoming);
  auto F         = fresnel_dielectric(ior, up_normal, outgoing);
  return F * sample_microfacet_pdf(roughness, up_normal, halfway) /
             (4 * vabs(dot(outgoing, halfway))) +
         (1 - F) * sample_hemisphere_cos_pdf(up_normal, incoming);
}

// ---- reflective (rough + delta) -------------------------------------------------
VPT_FN f3 eval_reflective(f3 color, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return {0, 0, 0};
  auto up_normal = up_normal_of(normal, outgoing);
  auto halfway   = normalize(incoming + outgoing);
  auto F = fresnel_conductor(reflectivity_to_eta(color), {0, 0, 0}, halfway, incoming);
  auto D = microfacet_distribution(roughness, up_normal, halfway);
  auto G = microfacet_shadowing(roughness, up_normal, halfway, outgoing, incoming);
  return F * D * G / (4 * dot(up_normal, outgoing) * dot(up_normal, incoming)) *
         vabs(dot(up_normal, incoming));
}
VPT_FN f3 sample_reflective(float roughness, f3 normal, f3 outgoing, f2 rn) {
  auto up_normal = up_normal_of(normal, outgoing);
  auto halfway   = sample_microfacet(roughness, up_normal, rn);
  auto incoming  = reflect(outgoing, halfway);
  if (!same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
  return incoming;
}
VPT_FN float sample_reflective_pdf(float roughness, f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return 0;
  auto up_normal = up_normal_of(normal, outgoing);
  auto halfway   = normalize(outgoing + incoming);
  return sample_microfacet_pdf(roughness, up_normal, halfway) /
         (4 * vabs(dot(outgoing, halfway)));
}
VPT_FN f3 eval_reflective_delta(f3 color, f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return {0, 0, 0};
  auto up_normal = up_normal_of(normal, outgoing);
  return fresnel_conductor(reflectivity_to_eta(color), {0, 0, 0}, up_normal, outgoing);
}
VPT_FN f3 sample_reflective_delta(f3 normal, f3 outgoing) {
  return reflect(outgoing, up_normal_of(normal, outgoing));
}
VPT_FN float sample_reflective_delta_pdf(f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return 0;
  return 1;
}

// ---- gltfpbr ------------------------------------------------------------------
VPT_FN f3 eval_gltfpbr(f3 color, float ior, float roughness, float metallic, f3 normal,
    f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return {0, 0, 0};
  auto reflectivity = lerp(eta_to_reflectivity(f3{ior, ior, ior}), color, metallic);
  auto up_normal    = up_normal_of(normal, outgoing);
  auto F1           = fresnel_schlick(reflectivity, up_normal, outgoing);
  auto halfway      = normalize(incoming + outgoing);
  auto F            = fresnel_schlick(reflectivity, halfway, incoming);
  auto D            = microfacet_distribution(roughness, up_normal, halfway);
  auto G            = microfacet_shadowing(roughness, up_normal, halfway, outgoing, incoming);
  return color * (1 - metallic) * (1 - F1) / pif * vabs(dot(up_normal, incoming)) +
         F * D * G / (4 * dot(up_normal, outgoing) * dot(up_normal, incoming)) *
             vabs(dot(up_normal, incoming));
}
VPT_FN f3 sample_gltfpbr(f3 color, float ior, float roughness, float metallic, f3 normal,
    f3 outgoing, float rnl, f2 rn) {
  auto up_normal    = up_normal_of(normal, outgoing);
  auto reflectivity = lerp(eta_to_reflectivity(f3{ior, ior, ior}), color, metallic);
  if (rnl < mean(fresnel_schlick(reflectivity, up_normal, outgoing))) {
    auto halfway  = sample_microfacet(roughness, up_normal, rn);
    auto incoming = reflect(outgoing, halfway);
    if (!same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
    return incoming;
  } else {
    return sample_hemisphere_cos(up_normal, rn);
  }
}
VPT_FN float sample_gltfpbr_pdf(f3 color, float ior, float roughness, float metallic, f3 normal,
    f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) <= 0) return 0;
  auto up_normal    = up_normal_of(normal, outgoing);
  auto halfway      = normalize(outgoing + incoming);
  auto reflectivity = lerp(eta_to_reflectivity(f3{ior, ior, ior}), color, metallic);
  auto F            = mean(fresnel_schlick(reflectivity, up_normal, outgoing));
  return F * sample_microfacet_pdf(roughness, up_normal, halfway) /
             (4 * vabs(dot(outgoing, halfway))) +
         (1 - F) * sample_hemisphere_cos_pdf(up_normal, incoming);
}

// ---- transparent (thin) ---------------------------------------------------------
VPT_FN f3 eval_transparent(
    f3 color, float ior, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  auto up_normal = up_normal_of(normal, outgoing);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    auto halfway = normalize(incoming + outgoing);
    auto F       = fresnel_dielectric(ior, halfway, outgoing);
    auto D       = microfacet_distribution(roughness, up_normal, halfway);
    auto G       = microfacet_shadowing(roughness, up_normal, halfway, outgoing, incoming);
    return f3{1, 1, 1} * F * D * G / (4 * dot(up_normal, outgoing) * dot(up_normal, incoming)) *
           vabs(dot(up_normal, incoming));
  } else {
    auto reflected = reflect(-incoming, up_normal);
    auto halfway   = normalize(reflected + outgoing);
    auto F         = fresnel_dielectric(ior, halfway, outgoing);
    auto D         = microfacet_distribution(roughness, up_normal, halfway);
    auto G         = microfacet_shadowing(roughness, up_normal, halfway, outgoing, reflected);
    return color * (1 - F) * D * G / (4 * dot(up_normal, outgoing) * dot(up_normal, reflected)) *
           (vabs(dot(up_normal, reflected)));
  }
}
VPT_FN f3 sample_transparent(float ior, float roughness, f3 normal, f3 outgoing, float rnl, f2 rn) {
  auto up_normal = up_normal_of(normal, outgoing);
  auto halfway   = sample_microfacet(roughness, up_normal, rn);
  if (rnl < fresnel_dielectric(ior, halfway, outgoing)) {
    auto incoming = reflect(outgoing, halfway);
    if (!same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
    return incoming;
  } else {
    auto reflected = reflect(outgoing, halfway);
    auto incoming  = -reflect(reflected, up_normal);
    if (same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
    return incoming;
  }
}
VPT_FN float sample_transparent_pdf(
    float ior, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  auto up_normal = up_normal_of(normal, outgoing);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    auto halfway = normalize(incoming + outgoing);
    return fresnel_dielectric(ior, halfway, outgoing) *
           sample_microfacet_pdf(roughness, up_normal, halfway) /
           (4 * vabs(dot(outgoing, halfway)));
  } else {
    auto reflected = reflect(-incoming, up_normal);
    auto halfway   = normalize(reflected + outgoing);
    auto d         = (1 - fresnel_dielectric(ior, halfway, outgoing)) *
             sample_microfacet_pdf(roughness, up_normal, halfway);
    return d / (4 * vabs(dot(outgoing, halfway)));
  }
}
VPT_FN f3 eval_transparent_delta(f3 color, float ior, f3 normal, f3 outgoing, f3 incoming) {
  auto up_normal = up_normal_of(normal, outgoing);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    return f3{1, 1, 1} * fresnel_dielectric(ior, up_normal, outgoing);
  } else {
    return color * (1 - fresnel_dielectric(ior, up_normal, outgoing));
  }
}
VPT_FN f3 sample_transparent_delta(float ior, f3 normal, f3 outgoing, float rnl) {
  auto up_normal = up_normal_of(normal, outgoing);
  if (rnl < fresnel_dielectric(ior, up_normal, outgoing)) {
    return reflect(outgoing, up_normal);
  } else {
    return -outgoing;
  }
}
VPT_FN float sample_transparent_delta_pdf(float ior, f3 normal, f3 outgoing, f3 incoming) {
  auto up_normal = up_normal_of(normal, outgoing);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    return fresnel_dielectric(ior, up_normal, outgoing);
  } else {
    return 1 - fresnel_dielectric(ior, up_normal, outgoing);
  }
}

// ---- refractive (rough + delta) -------------------------------------------------
VPT_FN f3 eval_refractive(float ior, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  auto entering  = dot(normal, outgoing) >= 0;
  auto up_normal = entering ? normal : -normal;
  auto rel_ior   = entering ? ior : (1 / ior);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    auto halfway = normalize(incoming + outgoing);
    auto F       = fresnel_dielectric(rel_ior, halfway, outgoing);
    auto D       = microfacet_distribution(roughness, up_normal, halfway);
    auto G       = microfacet_shadowing(roughness, up_normal, halfway, outgoing, incoming);
    return f3{1, 1, 1} * F * D * G / vabs(4 * dot(normal, outgoing) * dot(normal, incoming)) *
           vabs(dot(normal, incoming));
  } else {
    auto halfway = -normalize(rel_ior * incoming + outgoing) * (entering ? 1.0f : -1.0f);
    auto F       = fresnel_dielectric(rel_ior, halfway, outgoing);
    auto D       = microfacet_distribution(roughness, up_normal, halfway);
    auto G       = microfacet_shadowing(roughness, up_normal, halfway, outgoing, incoming);
    auto denom   = rel_ior * dot(halfway, incoming) + dot(halfway, outgoing);
    // pow(x, 2) in the reference is an exact square under a correctly rounded powf
    return f3{1, 1, 1} *
           vabs((dot(outgoing, halfway) * dot(incoming, halfway)) /
                (dot(outgoing, normal) * dot(incoming, normal))) *
           (1 - F) * D * G / (denom * denom) * vabs(dot(normal, incoming));
  }
}
VPT_FN f3 sample_refractive(float ior, float roughness, f3 normal, f3 outgoing, float rnl, f2 rn) {
  auto entering  = dot(normal, outgoing) >= 0;
  auto up_normal = entering ? normal : -normal;
  auto halfway   = sample_microfacet(roughness, up_normal, rn);
  if (rnl < fresnel_dielectric(entering ? ior : (1 / ior), halfway, outgoing)) {
    auto incoming = reflect(outgoing, halfway);
    if (!same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
    return incoming;
  } else {
    auto incoming = refract(outgoing, halfway, entering ? (1 / ior) : ior);
    if (same_hemisphere(up_normal, outgoing, incoming)) return {0, 0, 0};
    return incoming;
  }
}
VPT_FN float sample_refractive_pdf(
    float ior, float roughness, f3 normal, f3 outgoing, f3 incoming) {
  auto entering  = dot(normal, outgoing) >= 0;
  auto up_normal = entering ? normal : -normal;
  auto rel_ior   = entering ? ior : (1 / ior);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    auto halfway = normalize(incoming + outgoing);
    return fresnel_dielectric(rel_ior, halfway, outgoing) *
           sample_microfacet_pdf(roughness, up_normal, halfway) /
           (4 * vabs(dot(outgoing, halfway)));
  } else {
    auto halfway = -normalize(rel_ior * incoming + outgoing) * (entering ? 1.0f : -1.0f);
    auto denom   = rel_ior * dot(halfway, incoming) + dot(halfway, outgoing);
    return (1 - fresnel_dielectric(rel_ior, halfway, outgoing)) *
           sample_microfacet_pdf(roughness, up_normal, halfway) * vabs(dot(halfway, incoming)) /
           (denom * denom);
  }
}
VPT_FN bool ior_is_one(float ior) { return (double)vabs(ior - 1) < 1e-3; }
VPT_FN f3 eval_refractive_delta(float ior, f3 normal, f3 outgoing, f3 incoming) {
  if (ior_is_one(ior))
    return dot(normal, incoming) * dot(normal, outgoing) <= 0 ? f3{1, 1, 1} : f3{0, 0, 0};
  auto entering  = dot(normal, outgoing) >= 0;
  auto up_normal = entering ? normal : -normal;
  auto rel_ior   = entering ? ior : (1 / ior);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    return f3{1, 1, 1} * fresnel_dielectric(rel_ior, up_normal, outgoing);
  } else {
    return f3{1, 1, 1} * (1 / (rel_ior * rel_ior)) *
           (1 - fresnel_dielectric(rel_ior, up_normal, outgoing));
  }
}
VPT_FN f3 sample_refractive_delta(float ior, f3 normal, f3 outgoing, float rnl) {
  if (ior_is_one(ior)) return -outgoing;
  auto entering  = dot(normal, outgoing) >= 0;
  auto up_normal = entering ? normal : -normal;
  auto rel_ior   = entering ? ior : (1 / ior);
  if (rnl < fresnel_dielectric(rel_ior, up_normal, outgoing)) {
    return reflect(outgoing, up_normal);
  } else {
    return refract(outgoing, up_normal, 1 / rel_ior);
  }
}
VPT_FN float sample_refractive_delta_pdf(float ior, f3 normal, f3 outgoing, f3 incoming) {
  if (ior_is_one(ior)) return dot(normal, incoming) * dot(normal, outgoing) < 0 ? 1.0f : 0.0f;
  auto entering  = dot(normal, outgoing) >= 0;
  auto up_normal = entering ? normal : -normal;
  auto rel_ior   = entering ? ior : (1 / ior);
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) {
    return fresnel_dielectric(rel_ior, up_normal, outgoing);
  } else {
    return (1 - fresnel_dielectric(rel_ior, up_normal, outgoing));
  }
}

// ---- passthrough (volumetric boundary) ------------------------------------------
VPT_FN f3 eval_passthrough(f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) return {0, 0, 0};
  return {1, 1, 1};
}
VPT_FN float sample_passthrough_pdf(f3 normal, f3 outgoing, f3 incoming) {
  if (dot(normal, incoming) * dot(normal, outgoing) >= 0) return 0;
  return 1;
}

// ---- dispatch (yocto_pathtrace.cpp:86-236) --------------------------------------
VPT_FN f3 eval_emission(const material_point& m, f3 normal, f3 outgoing) {
  return dot(normal, outgoing) >= 0 ? m.emission : f3{0, 0, 0};
}

VPT_CALL f3 eval_bsdfcos(const material_point& m, f3 normal, f3 outgoing, f3 incoming) {
  if (m.roughness == 0) return {0, 0, 0};
  switch (m.type) {
    case mat_matte: return eval_matte(m.color, normal, outgoing, incoming);
    case mat_glossy: return eval_glossy(m.color, m.ior, m.roughness, normal, outgoing, incoming);
    case mat_reflective: return eval_reflective(m.color, m.roughness, normal, outgoing, incoming);
    case mat_transparent:
      return eval_transparent(m.color, m.ior, m.roughness, normal, outgoing, incoming);
    case mat_refractive:
    case mat_subsurface: return eval_refractive(m.ior, m.roughness, normal, outgoing, incoming);
    case mat_gltfpbr:
      return eval_gltfpbr(m.color, m.ior, m.roughness, m.metallic, normal, outgoing, incoming);
    default: return {0, 0, 0};
  }
}
VPT_CALL f3 eval_delta(const material_point& m, f3 normal, f3 outgoing, f3 incoming) {
  if (m.roughness != 0) return {0, 0, 0};
  switch (m.type) {
    case mat_reflective: return eval_reflective_delta(m.color, normal, outgoing, incoming);
    case mat_transparent: return eval_transparent_delta(m.color, m.ior, normal, outgoing, incoming);
    case mat_refractive: return eval_refractive_delta(m.ior, normal, outgoing, incoming);
    case mat_volumetric: return eval_passthrough(normal, outgoing, incoming);
    default: return {0, 0, 0};
  }
}
VPT_CALL f3 sample_bsdfcos(const material_point& m, f3 normal, f3 outgoing, float rnl, f2 rn) {
  if (m.roughness == 0) return {0, 0, 0};
  switch (m.type) {
    case mat_matte: return sample_matte(normal, outgoing, rn);
    case mat_glossy: return sample_glossy(m.ior, m.roughness, normal, outgoing, rnl, rn);
    case mat_reflective: return sample_reflective(m.roughness, normal, outgoing, rn);
    case mat_transparent: return sample_transparent(m.ior, m.roughness, normal, outgoing, rnl, rn);
    case mat_refractive:
    case mat_subsurface: return sample_refractive(m.ior, m.roughness, normal, outgoing, rnl, rn);
    case mat_gltfpbr:
      return sample_gltfpbr(m.color, m.ior, m.roughness, m.metallic, normal, outgoing, rnl, rn);
    default: return {0, 0, 0};
  }
}
VPT_CALL f3 sample_delta(const material_point& m, f3 normal, f3 outgoing, float rnl) {
  if (m.roughness != 0) return {0, 0, 0};
  switch (m.type) {
    case mat_reflective: return sample_reflective_delta(normal, outgoing);
    case mat_transparent: return sample_transparent_delta(m.ior, normal, outgoing, rnl);
    case mat_refractive: return sample_refractive_delta(m.ior, normal, outgoing, rnl);
    case mat_volumetric: return -outgoing;
    default: return {0, 0, 0};
  }
}
VPT_CALL float sample_bsdfcos_pdf(const material_point& m, f3 normal, f3 outgoing, f3 incoming) {
  if (m.roughness == 0) return 0;
  switch (m.type) {
    case mat_matte: return sample_matte_pdf(normal, outgoing, incoming);
    case mat_glossy: return sample_glossy_pdf(m.ior, m.roughness, normal, outgoing, incoming);
    case mat_reflective: return sample_reflective_pdf(m.roughness, normal, outgoing, incoming);
    case mat_transparent:
      return sample_transparent_pdf(m.ior, m.roughness, normal, outgoing, incoming);
    case mat_refractive:
    case mat_subsurface:
      return sample_refractive_pdf(m.ior, m.roughness, normal, outgoing, incoming);
    case mat_gltfpbr:
      return sample_gltfpbr_pdf(m.color, m.ior, m.roughness, m.metallic, normal, outgoing,
          incoming);
    default: return 0;
  }
}
VPT_CALL float sample_delta_pdf(const material_point& m, f3 normal, f3 outgoing, f3 incoming) {
  if (m.roughness != 0) return 0;
  switch (m.type) {
    case mat_reflective: return sample_reflective_delta_pdf(normal, outgoing, incoming);
    case mat_transparent: return sample_transparent_delta_pdf(m.ior, normal, outgoing, incoming);
    case mat_refractive: return sample_refractive_delta_pdf(m.ior, normal, outgoing, incoming);
    case mat_volumetric: return sample_passthrough_pdf(normal, outgoing, incoming);
    default: return 0;
  }
}

// ---- homogeneous volumes (yocto_shading.h:1047-1102, yocto_pathtrace.cpp:238-255) ----
VPT_FN f3 eval_transmittance(f3 density, float distance) { return vexp(-density * distance); }
VPT_FN float sample_transmittance(f3 density, float max_distance, float rl, float rd) {
  auto channel  = vclamp((int)(rl * 3), 0, 2);
  auto sigma    = comp(density, channel);
  auto distance = (sigma == 0) ? flt_max : -vlog(1 - rd) / sigma;
  return vmin(distance, max_distance);
}
VPT_FN float sample_transmittance_pdf(f3 density, float distance, float max_distance) {
  if (distance < max_distance) {
    return sum(density * vexp(-density * distance)) / 3;
  } else {
    return sum(vexp(-density * max_distance)) / 3;
  }
}
VPT_FN float eval_phasefunction(float anisotropy, f3 outgoing, f3 incoming) {
  auto cosine = -dot(outgoing, incoming);
  auto denom  = 1 + anisotropy * anisotropy - 2 * anisotropy * cosine;
  return (1 - anisotropy * anisotropy) / (4 * pif * denom * vsqrt(denom));
}
VPT_FN f3 sample_phasefunction(float anisotropy, f3 outgoing, f2 rn) {
  auto cos_theta = 0.0f;
  if (vabs(anisotropy) < 1e-3f) {
    cos_theta = 1 - 2 * rn.y;
  } else {
    auto square = (1 - anisotropy * anisotropy) / (1 + anisotropy - 2 * anisotropy * rn.y);
    cos_theta   = (1 + anisotropy * anisotropy - square * square) / (2 * anisotropy);
  }
  auto sin_theta = vsqrt(vmax(0.0f, 1 - cos_theta * cos_theta));
  auto phi       = 2 * pif * rn.x;
  auto local     = f3{sin_theta * vcos(phi), sin_theta * vsin(phi), cos_theta};
  return basis_fromz(-outgoing) * local;
}
// eval_scattering: note the reference passes (incoming, outgoing) to the phase
// function (yocto_pathtrace.cpp:241-242); HG is symmetric in the pair.
VPT_FN f3 eval_scattering(const material_point& m, f3 outgoing, f3 incoming) {
  return m.density * m.scattering * eval_phasefunction(m.scanisotropy, incoming, outgoing);
}
VPT_FN float sample_scattering_pdf(const material_point& m, f3 outgoing, f3 incoming) {
  return eval_phasefunction(m.scanisotropy, outgoing, incoming);
}

}  // namespace vpt
