// ref_scene_from_archive.h — TEST INFRASTRUCTURE (oracle side).
// Rebuilds the reference's scene_data (yocto_scene.h:221-249) from a
// ".vptscene" archive written by vpt_export or by the Python fixture
// generators, so the UNMODIFIED reference CPU tracer can render exactly the
// scene the CUDA path renders, on a box where /root/reference is absent.
// Analytic SDFs get their std::function closures back from the typed
// parameters, bound to the reference's own sd_* (yocto_sdfs.h:43-80) exactly
// as the JSON loader binds them (yocto_sceneio.cpp:3685-3728).
#pragma once

#include <yocto/yocto_scene.h>
#include <yocto/yocto_sdfs.h>

#include "vpt_archive.h"

namespace vptref {

using namespace yocto;

inline frame3f get_frame(const float* p) {
  auto f = frame3f{};
  memcpy(&f, p, 48);
  return f;
}

inline bool scene_from_archive(
    const vpt::archive& ar, scene_data& scene, std::string& error) {
  auto need = [&](const char* name) -> const vpt::archive_array* {
    auto a = ar.find(name);
    if (!a) error = std::string("missing array ") + name;
    return a;
  };
  scene = scene_data{};
  if (auto a = need("cameras")) {
    for (size_t i = 0; i < a->dims[0]; i++) {
      auto  p        = a->f32() + i * 18;
      auto& c        = scene.cameras.emplace_back();
      c.frame        = get_frame(p);
      c.orthographic = p[12] != 0;
      c.lens = p[13], c.film = p[14], c.aspect = p[15], c.focus = p[16],
      c.aperture = p[17];
    }
  } else
    return false;
  auto inf = need("instances.frame"), ini = need("instances.ids");
  if (!inf || !ini) return false;
  for (size_t i = 0; i < inf->dims[0]; i++) {
    auto& in    = scene.instances.emplace_back();
    in.frame    = get_frame(inf->f32() + i * 12);
    in.shape    = ini->i32()[i * 2 + 0];
    in.material = ini->i32()[i * 2 + 1];
  }
  auto mf = need("materials.f"), mi = need("materials.i");
  if (!mf || !mi) return false;
  for (size_t i = 0; i < mf->dims[0]; i++) {
    auto  p        = mf->f32() + i * 16;
    auto  q        = mi->i32() + i * 6;
    auto& m        = scene.materials.emplace_back();
    m.type         = (material_type)q[0];
    m.emission     = {p[0], p[1], p[2]};
    m.color        = {p[3], p[4], p[5]};
    m.roughness    = p[6];
    m.metallic     = p[7];
    m.ior          = p[8];
    m.scattering   = {p[9], p[10], p[11]};
    m.scanisotropy = p[12];
    m.trdepth      = p[13];
    m.opacity      = p[14];
    m.emission_tex = q[1], m.color_tex = q[2], m.roughness_tex = q[3],
    m.scattering_tex = q[4], m.normal_tex = q[5];
  }
  auto ef = need("environments.f"), ei = need("environments.i");
  if (!ef || !ei) return false;
  for (size_t i = 0; i < ef->dims[0]; i++) {
    auto  p        = ef->f32() + i * 15;
    auto& e        = scene.environments.emplace_back();
    e.frame        = get_frame(p);
    e.emission     = {p[12], p[13], p[14]};
    e.emission_tex = ei->i32()[i];
  }
  auto sc = need("shapes.count");
  if (!sc) return false;
  for (int i = 0; i < sc->i32()[0]; i++) {
    auto& s   = scene.shapes.emplace_back();
    auto  pre = "shape" + std::to_string(i) + ".";
    if (auto a = ar.find(pre + "positions"))
      s.positions.assign(
          (const vec3f*)a->f32(), (const vec3f*)a->f32() + a->dims[0]);
    if (auto a = ar.find(pre + "normals"))
      s.normals.assign(
          (const vec3f*)a->f32(), (const vec3f*)a->f32() + a->dims[0]);
    if (auto a = ar.find(pre + "texcoords"))
      s.texcoords.assign(
          (const vec2f*)a->f32(), (const vec2f*)a->f32() + a->dims[0]);
    if (auto a = ar.find(pre + "colors"))
      s.colors.assign(
          (const vec4f*)a->f32(), (const vec4f*)a->f32() + a->dims[0]);
    if (auto a = ar.find(pre + "triangles"))
      s.triangles.assign(
          (const vec3i*)a->i32(), (const vec3i*)a->i32() + a->dims[0]);
    if (auto a = ar.find(pre + "quads"))
      s.quads.assign(
          (const vec4i*)a->i32(), (const vec4i*)a->i32() + a->dims[0]);
  }
  auto tc = need("textures.count");
  if (!tc) return false;
  for (int i = 0; i < tc->i32()[0]; i++) {
    auto& t   = scene.textures.emplace_back();
    auto  pre = "tex" + std::to_string(i) + ".";
    auto  m   = ar.find(pre + "meta");
    if (!m) {
      error = "missing " + pre + "meta";
      return false;
    }
    t.width = m->i32()[0], t.height = m->i32()[1], t.linear = m->i32()[2] != 0;
    if (auto a = ar.find(pre + "pixelsf"))
      t.pixelsf.assign(
          (const vec4f*)a->f32(), (const vec4f*)a->f32() + a->count() / 4);
    if (auto a = ar.find(pre + "pixelsb"))
      t.pixelsb.assign(
          (const vec4b*)a->u8(), (const vec4b*)a->u8() + a->count() / 4);
  }
  if (auto vc = ar.find("volumes.count")) {
    for (int i = 0; i < vc->i32()[0]; i++) {
      auto& v   = scene.volumes.emplace_back();
      auto  pre = "vol" + std::to_string(i) + ".";
      auto  whd = ar.find(pre + "whd");
      auto  res = ar.find(pre + "res");
      auto  dat = ar.find(pre + "data");
      if (!whd || !res || !dat) {
        error = "missing " + pre + "*";
        return false;
      }
      v.whd = {whd->i32()[0], whd->i32()[1], whd->i32()[2]};
      v.res = res->f32()[0];
      v.vol.assign(dat->f32(), dat->f32() + dat->count());
    }
  }
  if (auto vf = ar.find("vol_instances.f")) {
    auto vi = ar.find("vol_instances.i");
    for (size_t i = 0; vi && i < vf->dims[0]; i++) {
      auto& in    = scene.vol_instances.emplace_back();
      in.frame    = get_frame(vf->f32() + i * 13);
      in.scalef   = vf->f32()[i * 13 + 12];
      in.volume   = vi->i32()[i * 2 + 0];
      in.material = vi->i32()[i * 2 + 1];
    }
  }
  if (auto sf = ar.find("sdfs.f")) {
    auto si = ar.find("sdfs.i");
    for (size_t i = 0; si && i < sf->dims[0]; i++) {
      auto  p      = sf->f32() + i * 19;
      auto& sdf    = scene.sdfs.emplace_back();
      sdf.frame    = get_frame(p);
      sdf.material = si->i32()[i * 2 + 1];
      sdf.whd      = {p[16], p[17], p[18]};
      switch ((sdf_type)si->i32()[i * 2 + 0]) {
        case sdf_type::bbox: {
          auto th = p[12];
          auto w  = vec3f{p[13], p[14], p[15]};
          sdf.f   = [th, w](const vec3f& q) { return sd_bbox(q, w, th); };
        } break;
        case sdf_type::box: {
          auto copy = sdf;
          sdf.f     = [copy](const vec3f& q) {
            return sd_box(q - (copy.whd * 0.5f), copy.whd * 0.5f);
          };
        } break;
        case sdf_type::capped_cone: {
          auto h = p[12], r1 = p[13], r2 = p[14];
          sdf.f = [h, r1, r2](
                      const vec3f& q) { return sd_capped_cone(q, h, r1, r2); };
        } break;
        case sdf_type::plane: {
          sdf.f = [](const vec3f& q) { return sd_plane(q); };
        } break;
        case sdf_type::sphere: {
          auto r = p[12];
          sdf.f  = [r](const vec3f& q) { return sd_sphere(q, r); };
        } break;
        case sdf_type::torus: {
          auto r1 = p[12], r2 = p[13];
          sdf.f = [r1, r2](const vec3f& q) { return sd_torus(q, r1, r2); };
        } break;
        default: error = "bad sdf type"; return false;
      }
    }
  }
  return true;
}

}  // namespace vptref
