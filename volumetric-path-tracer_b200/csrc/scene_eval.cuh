// scene_eval.cuh — per-hit scene evaluation on the device: camera rays,
// texture fetches, interpolated positions / normals / texcoords, material
// points and environment lookups. Semantics (expression order, wrap rules,
// roughness clamps) follow yocto_scene.cpp:67-651 so the deterministic shaders
// reproduce the reference bit for bit.
#pragma once

#include "dscene.cuh"

namespace vpt {

// ---- camera (yocto_scene.cpp:67-102) ----------------------------------------
VPT_FN ray3 eval_camera(const DScene& sc, int cam, f2 image_uv, f2 lens_uv) {
  auto p  = sc.cameras + 5 * cam;
  auto c0 = ld4(p), c1 = ld4(p + 1), c2 = ld4(p + 2), c3 = ld4(p + 3), c4 = ld4(p + 4);
  auto frame = frame3{xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
  auto ortho = f2i(c0.w) != 0;
  auto lens = c1.w, film_ = c2.w, aspect = c3.w, focus = c4.x, aperture = c4.y;
  auto film = aspect >= 1 ? f2{film_, film_ / aspect} : f2{film_ * aspect, film_};
  if (!ortho) {
    auto q  = f3{film.x * (0.5f - image_uv.x), film.y * (image_uv.y - 0.5f), lens};
    auto dc = -normalize(q);
    auto e  = f3{lens_uv.x * aperture / 2, lens_uv.y * aperture / 2, 0};
    auto pf = dc * focus / vabs(dc.z);
    auto d  = normalize(pf - e);
    return make_ray(transform_point(frame, e), transform_direction(frame, d));
  } else {
    auto scale = 1 / lens;
    auto q     = f3{film.x * (0.5f - image_uv.x) * scale, film.y * (image_uv.y - 0.5f) * scale,
        lens};
    auto e  = f3{-q.x, -q.y, 0} + f3{lens_uv.x * aperture / 2, lens_uv.y * aperture / 2, 0};
    auto pf = f3{-q.x, -q.y, -focus};
    auto d  = normalize(pf - e);
    return make_ray(transform_point(frame, e), transform_direction(frame, d));
  }
}

// ---- textures (yocto_scene.cpp:112-169, yocto_color.h:212-238) ---------------
VPT_FN float srgb_to_rgb1(float srgb) {
  return ((double)srgb <= 0.04045) ? srgb / 12.92f
                                   : vpow((srgb + 0.055f) / (1.0f + 0.055f), 2.4f);
}

VPT_FN f4 lookup_texture(const DScene& sc, i4 tex, int i, int j, bool as_linear) {
  VPT_WORK(texel_reads, 1);
  auto idx    = tex.w + j * tex.x + i;
  auto decode = as_linear && !(tex.z & 1);
  if (tex.z & 2) {
    auto c = ld4(sc.texf + idx);
    if (decode) return {srgb_to_rgb1(c.x), srgb_to_rgb1(c.y), srgb_to_rgb1(c.z), c.w};
    return c;
  } else {
#if defined(__CUDA_ARCH__)
    auto packed = __ldg(reinterpret_cast<const unsigned int*>(sc.texb) + idx);
#else
    auto packed = reinterpret_cast<const unsigned int*>(sc.texb)[idx];
#endif
    int r = packed & 255, g = (packed >> 8) & 255, b = (packed >> 16) & 255, a = packed >> 24;
    if (decode) {
      // byte -> linear through a 256-entry table built on the host with the
      // host libm: srgb_to_rgb(b / 255.0f) per texel before interpolation.
      return {ld1(sc.srgb_lut + r), ld1(sc.srgb_lut + g), ld1(sc.srgb_lut + b), a / 255.0f};
    }
    return {r / 255.0f, g / 255.0f, b / 255.0f, a / 255.0f};
  }
}

VPT_FN f4 eval_texture_rec(const DScene& sc, i4 tex, f2 uv, bool as_linear) {
  if (tex.x == 0 || tex.y == 0) return {0, 0, 0, 0};
  auto s = vfmod(uv.x, 1.0f) * tex.x;
  if (s < 0) s += tex.x;
  auto t = vfmod(uv.y, 1.0f) * tex.y;
  if (t < 0) t += tex.y;
  auto i = vclamp((int)s, 0, tex.x - 1), j = vclamp((int)t, 0, tex.y - 1);
  auto ii = (i + 1) % tex.x, jj = (j + 1) % tex.y;
  auto u = s - i, v = t - j;
  return lookup_texture(sc, tex, i, j, as_linear) * (1 - u) * (1 - v) +
         lookup_texture(sc, tex, i, jj, as_linear) * (1 - u) * v +
         lookup_texture(sc, tex, ii, j, as_linear) * u * (1 - v) +
         lookup_texture(sc, tex, ii, jj, as_linear) * u * v;
}

VPT_CALL f4 eval_texture(const DScene& sc, int texture, f2 uv, bool as_linear) {
  if (texture == invalidid) return {1, 1, 1, 1};
  return eval_texture_rec(sc, ld4(sc.textures + texture), uv, as_linear);
}

// ---- shapes ------------------------------------------------------------------
struct shape_ref {
  int kind, pos, nrm, tex, col, elem;
};
struct inst_ref {
  frame3    frame;
  int       shape, material;
  shape_ref s;
};

VPT_FN inst_ref ld_instance(const DScene& sc, int instance) {
  auto p  = sc.inst_frame + 4 * instance;
  auto c0 = ld4(p), c1 = ld4(p + 1), c2 = ld4(p + 2), c3 = ld4(p + 3);
  auto r  = inst_ref{};
  r.frame = {xyz(c0), xyz(c1), xyz(c2), xyz(c3)};
  r.shape = f2i(c0.w), r.material = f2i(c1.w);
  auto a = ld4(sc.shapes + 2 * r.shape), b = ld4(sc.shapes + 2 * r.shape + 1);
  r.s    = {a.x, a.y, a.z, a.w, b.x, b.y};
  return r;
}

VPT_FN f3 interp_tri(f3 p0, f3 p1, f3 p2, f2 uv) {
  return p0 * (1 - uv.x - uv.y) + p1 * uv.x + p2 * uv.y;
}
VPT_FN f2 interp_tri(f2 p0, f2 p1, f2 p2, f2 uv) {
  return p0 * (1 - uv.x - uv.y) + p1 * uv.x + p2 * uv.y;
}
VPT_FN f4 interp_tri(f4 p0, f4 p1, f4 p2, f2 uv) {
  return p0 * (1 - uv.x - uv.y) + p1 * uv.x + p2 * uv.y;
}
template <typename T>
VPT_FN T interp_quad(T p0, T p1, T p2, T p3, f2 uv) {
  if (uv.x + uv.y <= 1) return interp_tri(p0, p1, p3, uv);
  return interp_tri(p2, p3, p1, f2{1 - uv.x, 1 - uv.y});
}

VPT_FN f3 vertex3(const float* pool, int off, int v) { return ld3(pool + 3 * (size_t)(off + v)); }

VPT_FN f3 eval_local_position(const DScene& sc, const shape_ref& s, int element, f2 uv) {
  auto q = ld4(sc.elems + s.elem + element);
  if (s.kind == kind_triangles) {
    return interp_tri(vertex3(sc.vpos, s.pos, q.x), vertex3(sc.vpos, s.pos, q.y),
        vertex3(sc.vpos, s.pos, q.z), uv);
  } else if (s.kind == kind_quads) {
    return interp_quad(vertex3(sc.vpos, s.pos, q.x), vertex3(sc.vpos, s.pos, q.y),
        vertex3(sc.vpos, s.pos, q.z), vertex3(sc.vpos, s.pos, q.w), uv);
  }
  return {0, 0, 0};
}

// eval_position (yocto_scene.cpp:279-302)
VPT_FN f3 eval_position(const DScene& sc, const inst_ref& in, int element, f2 uv) {
  if (in.s.kind == kind_empty) return {0, 0, 0};
  return transform_point(in.frame, eval_local_position(sc, in.s, element, uv));
}

VPT_FN f3 triangle_normal(f3 p0, f3 p1, f3 p2) { return normalize(cross(p1 - p0, p2 - p0)); }
VPT_FN f3 quad_normal(f3 p0, f3 p1, f3 p2, f3 p3) {
  return normalize(triangle_normal(p0, p1, p3) + triangle_normal(p2, p3, p1));
}
// transform_normal(frame, n) with non_rigid = false (yocto_math.h:3106-3113)
VPT_FN f3 transform_normal(const frame3& a, f3 b) { return normalize(transform_vector(a, b)); }

// eval_element_normal (yocto_scene.cpp:305-327)
VPT_FN f3 eval_element_normal(const DScene& sc, const inst_ref& in, int element) {
  if (in.s.kind == kind_empty) return {0, 0, 0};
  auto q = ld4(sc.elems + in.s.elem + element);
  if (in.s.kind == kind_triangles) {
    return transform_normal(in.frame, triangle_normal(vertex3(sc.vpos, in.s.pos, q.x),
                                          vertex3(sc.vpos, in.s.pos, q.y),
                                          vertex3(sc.vpos, in.s.pos, q.z)));
  }
  return transform_normal(
      in.frame, quad_normal(vertex3(sc.vpos, in.s.pos, q.x), vertex3(sc.vpos, in.s.pos, q.y),
                    vertex3(sc.vpos, in.s.pos, q.z), vertex3(sc.vpos, in.s.pos, q.w)));
}

// eval_normal (yocto_scene.cpp:330-356)
VPT_FN f3 eval_normal(const DScene& sc, const inst_ref& in, int element, f2 uv) {
  if (in.s.nrm < 0) return eval_element_normal(sc, in, element);
  auto q = ld4(sc.elems + in.s.elem + element);
  if (in.s.kind == kind_triangles) {
    return transform_normal(
        in.frame, normalize(interp_tri(vertex3(sc.vnrm, in.s.nrm, q.x),
                      vertex3(sc.vnrm, in.s.nrm, q.y), vertex3(sc.vnrm, in.s.nrm, q.z), uv)));
  } else if (in.s.kind == kind_quads) {
    return transform_normal(in.frame,
        normalize(interp_quad(vertex3(sc.vnrm, in.s.nrm, q.x), vertex3(sc.vnrm, in.s.nrm, q.y),
            vertex3(sc.vnrm, in.s.nrm, q.z), vertex3(sc.vnrm, in.s.nrm, q.w), uv)));
  }
  return {0, 0, 0};
}

VPT_FN f2 vertex2(const float* pool, int off, int v) { return ld2(pool + 2 * (size_t)(off + v)); }

// eval_texcoord (yocto_scene.cpp:359-378)
VPT_FN f2 eval_texcoord(const DScene& sc, const inst_ref& in, int element, f2 uv) {
  if (in.s.tex < 0) return uv;
  auto q = ld4(sc.elems + in.s.elem + element);
  if (in.s.kind == kind_triangles) {
    return interp_tri(vertex2(sc.vtex, in.s.tex, q.x), vertex2(sc.vtex, in.s.tex, q.y),
        vertex2(sc.vtex, in.s.tex, q.z), uv);
  } else if (in.s.kind == kind_quads) {
    return interp_quad(vertex2(sc.vtex, in.s.tex, q.x), vertex2(sc.vtex, in.s.tex, q.y),
        vertex2(sc.vtex, in.s.tex, q.z), vertex2(sc.vtex, in.s.tex, q.w), uv);
  }
  return {0, 0};
}

// eval_color (yocto_scene.cpp:506-526)
VPT_FN f4 eval_color(const DScene& sc, const inst_ref& in, int element, f2 uv) {
  if (in.s.col < 0) return {1, 1, 1, 1};
  auto q = ld4(sc.elems + in.s.elem + element);
  if (in.s.kind == kind_triangles) {
    return interp_tri(ld4(sc.vcol + in.s.col + q.x), ld4(sc.vcol + in.s.col + q.y),
        ld4(sc.vcol + in.s.col + q.z), uv);
  } else if (in.s.kind == kind_quads) {
    return interp_quad(ld4(sc.vcol + in.s.col + q.x), ld4(sc.vcol + in.s.col + q.y),
        ld4(sc.vcol + in.s.col + q.z), ld4(sc.vcol + in.s.col + q.w), uv);
  }
  return {0, 0, 0, 0};
}

// ---- materials ---------------------------------------------------------------
struct material_t {  // material_data (yocto_scene.h:119-139)
  int   type;
  f3    emission, color;
  float roughness, metallic, ior;
  f3    scattering;
  float scanisotropy, trdepth, opacity;
  int   emission_tex, color_tex, roughness_tex, scattering_tex, normal_tex;
};
struct material_point {  // yocto_scene.h:292
  int   type;
  f3    emission, color;
  float opacity, roughness, metallic, ior;
  f3    density, scattering;
  float scanisotropy, trdepth;
};

enum : int {
  mat_matte = 0, mat_glossy, mat_reflective, mat_transparent, mat_refractive, mat_subsurface,
  mat_volumetric, mat_gltfpbr
};
constexpr float min_roughness = 0.03f * 0.03f;  // yocto_scene.h:191

VPT_FN material_t ld_material(const DScene& sc, int id) {
  auto p = sc.materials + 6 * id;
  auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2), d = ld4(p + 3), e = ld4(p + 4),
       f = ld4(p + 5);
  auto m = material_t{};
  m.type = f2i(a.x), m.emission = {a.y, a.z, a.w};
  m.color = {b.x, b.y, b.z}, m.roughness = b.w;
  m.metallic = c.x, m.ior = c.y, m.scattering = {c.z, c.w, d.x};
  m.scanisotropy = d.y, m.trdepth = d.z, m.opacity = d.w;
  m.emission_tex = f2i(e.x), m.color_tex = f2i(e.y), m.roughness_tex = f2i(e.z),
  m.scattering_tex = f2i(e.w), m.normal_tex = f2i(f.x);
  return m;
}

VPT_FN bool is_delta(const material_point& m) {
  return (m.type == mat_reflective && m.roughness == 0) ||
         (m.type == mat_refractive && m.roughness == 0) ||
         (m.type == mat_transparent && m.roughness == 0) || (m.type == mat_volumetric);
}
VPT_FN bool is_volumetric_type(int type) {
  return type == mat_refractive || type == mat_volumetric || type == mat_subsurface;
}

VPT_FN void finish_material_point(material_point& point) {
  if (is_volumetric_type(point.type)) {
    point.density = -vlog(vclamp(point.color, 0.0001f, 1.0f)) / point.trdepth;
  } else {
    point.density = {0, 0, 0};
  }
  if (point.type == mat_matte || point.type == mat_gltfpbr || point.type == mat_glossy) {
    point.roughness = vclamp(point.roughness, min_roughness, 1.0f);
  } else if (point.type == mat_volumetric) {
    point.roughness = 0;
  } else {
    if (point.roughness < min_roughness) point.roughness = 0;
  }
}

// eval_material(scene, instance, element, uv) (yocto_scene.cpp:529-579)
VPT_CALL material_point eval_material(
    const DScene& sc, const inst_ref& in, const material_t& m, int element, f2 uv) {
  auto texcoord       = eval_texcoord(sc, in, element, uv);
  auto emission_tex   = eval_texture(sc, m.emission_tex, texcoord, true);
  auto color_shp      = eval_color(sc, in, element, uv);
  auto color_tex      = eval_texture(sc, m.color_tex, texcoord, true);
  auto roughness_tex  = eval_texture(sc, m.roughness_tex, texcoord, false);
  auto scattering_tex = eval_texture(sc, m.scattering_tex, texcoord, true);
  auto point          = material_point{};
  point.type          = m.type;
  point.emission      = m.emission * xyz(emission_tex);
  point.color         = m.color * xyz(color_tex) * xyz(color_shp);
  point.opacity       = m.opacity * color_tex.w * color_shp.w;
  point.metallic      = m.metallic * roughness_tex.z;
  point.roughness     = m.roughness * roughness_tex.y;
  point.roughness     = point.roughness * point.roughness;
  point.ior           = m.ior;
  point.scattering    = m.scattering * xyz(scattering_tex);
  point.scanisotropy  = m.scanisotropy;
  point.trdepth       = m.trdepth;
  finish_material_point(point);
  return point;
}

// eval_material(scene, int) (yocto_scene.cpp:581-619): no textures
VPT_FN material_point eval_material_plain(const material_t& m) {
  auto point         = material_point{};
  point.type         = m.type;
  point.emission     = m.emission;
  point.color        = m.color;
  point.opacity      = m.opacity;
  point.metallic     = m.metallic;
  point.roughness    = m.roughness;
  point.roughness    = point.roughness * point.roughness;
  point.ior          = m.ior;
  point.scattering   = m.scattering;
  point.scanisotropy = m.scanisotropy;
  point.trdepth      = m.trdepth;
  finish_material_point(point);
  return point;
}

// ---- shading normal (yocto_scene.cpp:414-503) --------------------------------
VPT_FN void triangle_tangents_fromuv(
    f3 p0, f3 p1, f3 p2, f2 uv0, f2 uv1, f2 uv2, f3& tu, f3& tv) {
  auto p   = p1 - p0;
  auto q   = p2 - p0;
  auto s   = f2{uv1.x - uv0.x, uv2.x - uv0.x};
  auto t   = f2{uv1.y - uv0.y, uv2.y - uv0.y};
  auto div = s.x * t.y - s.y * t.x;
  if (div != 0) {
    tu = f3{t.y * p.x - t.x * q.x, t.y * p.y - t.x * q.y, t.y * p.z - t.x * q.z} / div;
    tv = f3{s.x * q.x - s.y * p.x, s.x * q.y - s.y * p.y, s.x * q.z - s.y * p.z} / div;
  } else {
    tu = {1, 0, 0};
    tv = {0, 1, 0};
  }
}

VPT_FN void eval_element_tangents(
    const DScene& sc, const inst_ref& in, int element, f3& tu, f3& tv) {
  tu = {0, 0, 0}, tv = {0, 0, 0};
  if (in.s.tex < 0) return;
  auto q = ld4(sc.elems + in.s.elem + element);
  if (in.s.kind == kind_triangles) {
    triangle_tangents_fromuv(vertex3(sc.vpos, in.s.pos, q.x), vertex3(sc.vpos, in.s.pos, q.y),
        vertex3(sc.vpos, in.s.pos, q.z), vertex2(sc.vtex, in.s.tex, q.x),
        vertex2(sc.vtex, in.s.tex, q.y), vertex2(sc.vtex, in.s.tex, q.z), tu, tv);
  } else if (in.s.kind == kind_quads) {
    // quad_tangents_fromuv with current_uv = {0,0}: always the (p0,p1,p3) half
    triangle_tangents_fromuv(vertex3(sc.vpos, in.s.pos, q.x), vertex3(sc.vpos, in.s.pos, q.y),
        vertex3(sc.vpos, in.s.pos, q.w), vertex2(sc.vtex, in.s.tex, q.x),
        vertex2(sc.vtex, in.s.tex, q.y), vertex2(sc.vtex, in.s.tex, q.w), tu, tv);
  } else {
    return;
  }
  tu = transform_direction(in.frame, tu);
  tv = transform_direction(in.frame, tv);
}

VPT_FN f3 eval_normalmap(
    const DScene& sc, const inst_ref& in, const material_t& m, int element, f2 uv) {
  auto normal   = eval_normal(sc, in, element, uv);
  auto texcoord = eval_texcoord(sc, in, element, uv);
  if (m.normal_tex != invalidid && in.s.kind != kind_empty) {
    auto normalmap = -1 + 2 * xyz(eval_texture_rec(sc, ld4(sc.textures + m.normal_tex), texcoord,
                                  false));
    f3 tu, tv;
    eval_element_tangents(sc, in, element, tu, tv);
    auto fx     = orthonormalize(tu, normal);
    auto fy     = normalize(cross(normal, fx));
    auto flip_v = dot(fy, tv) < 0;
    normalmap.y *= flip_v ? 1 : -1;
    // transform_normal(frame, n) with a non-rigid flag of false: normalize(frame * n)
    normal = normalize(fx * normalmap.x + fy * normalmap.y + normal * normalmap.z);
  }
  return normal;
}

VPT_CALL f3 eval_shading_normal(const DScene& sc, const inst_ref& in, const material_t& m,
    int element, f2 uv, f3 outgoing) {
  if (in.s.kind == kind_empty) return {0, 0, 0};
  auto normal = eval_normal(sc, in, element, uv);
  if (m.normal_tex != invalidid) normal = eval_normalmap(sc, in, m, element, uv);
  if (m.type == mat_refractive) return normal;
  return dot(normal, outgoing) >= 0 ? normal : -normal;
}

// ---- environments (yocto_scene.cpp:634-651) ----------------------------------
VPT_FN f2 environment_texcoord(f3 wl) {
  auto texcoord = f2{vatan2(wl.z, wl.x) / (2 * pif), vacos(vclamp(wl.y, -1.0f, 1.0f)) / pif};
  if (texcoord.x < 0) texcoord.x += 1;
  return texcoord;
}

VPT_CALL f3 eval_environment(const DScene& sc, f3 direction) {
  auto emission = f3{0, 0, 0};
  for (auto e = 0; e < sc.num_environments; e++) {
    auto p   = sc.environments + 9 * e;
    auto inv = ld_frame(p + 4);
    auto em  = ld4(p + 8);
    auto wl  = transform_direction(inv, direction);
    auto tc  = environment_texcoord(wl);
    emission += f3{em.x, em.y, em.z} * xyz(eval_texture(sc, f2i(em.w), tc, false));
  }
  return emission;
}

}  // namespace vpt
