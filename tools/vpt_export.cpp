// vpt_export — the product's scene-input tool: JSON / glTF / OBJ / PLY scene -> .vptscene, the
// archive `ypathtrace` and the bindings read. Scene loading is outside the hot path:
// it is done by the reference's own loaders, so this tool is compiled against the reference
// (oracle/Makefile builds it where /root/reference exists). Loads a scene with the
// reference's own loaders (load_scene, yocto_sceneio.cpp:2676; JSON 3544),
// runs the reference's host-side tesselate_surfaces (yocto_pathtrace.cpp:1275)
// and writes the result as a ".vptscene" named-array archive
// (volumetric-path-tracer_b200/host/vpt_archive.h) that both the product host
// library and the oracle renderer (ref_render) can read. Compiled against the
// reference headers where they lie; exists only where /root/reference does.
//
// usage: vpt_export <scene.json> <out.vptscene> [--keep-subdivs] [--dump-subdivs <file>]
//   --keep-subdivs        leave subdivision surfaces un-tessellated: the archive carries the
//                         control cages (subdiv<i>.*) and ypathtrace tessellates them on the GPU
//   --dump-subdivs FILE   additionally write only the control cages (tests compare the device
//                         tessellation of FILE with the reference-tessellated archive)
#include <yocto/yocto_scene.h>
#include <yocto/yocto_sceneio.h>
#include <yocto/yocto_sdfs.h>
#include <yocto_pathtrace/yocto_pathtrace.h>

#include <fstream>
#include <sstream>

#include "json.hpp"
#include "vpt_archive.h"

using namespace yocto;
using std::string;
using std::vector;

static void put_frame(float* dst, const frame3f& f) { memcpy(dst, &f, 48); }

int main(int argc, const char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: vpt_export <scene.json> <out.vptscene>\n");
    return 2;
  }
  auto filename = string{argv[1]};
  auto error    = string{};
  auto scene    = scene_data{};
  if (!load_scene(filename, scene, error)) {
    fprintf(stderr, "vpt_export: %s\n", error.c_str());
    return 1;
  }
  auto keep_subdivs = false;
  auto dump_subdivs = string{};
  for (int i = 3; i < argc; i++) {
    if (string{argv[i]} == "--keep-subdivs") keep_subdivs = true;
    if (string{argv[i]} == "--dump-subdivs" && i + 1 < argc) dump_subdivs = argv[++i];
  }
  auto put_subdivs = [&](vpt::archive& to) {
    int n = (int)scene.subdivs.size();
    to.add_i32("subdivs.count", {1}, &n);
    for (int i = 0; i < n; i++) {
      auto& s   = scene.subdivs[i];
      auto  pre = "subdiv" + std::to_string(i) + ".";
      to.add_i32(pre + "quadspos", {s.quadspos.size(), 4}, (const int*)s.quadspos.data());
      if (!s.quadsnorm.empty())
        to.add_i32(pre + "quadsnorm", {s.quadsnorm.size(), 4}, (const int*)s.quadsnorm.data());
      if (!s.quadstexcoord.empty())
        to.add_i32(pre + "quadstexcoord", {s.quadstexcoord.size(), 4}, (const int*)s.quadstexcoord.data());
      to.add_f32(pre + "positions", {s.positions.size(), 3}, (const float*)s.positions.data());
      if (!s.normals.empty())
        to.add_f32(pre + "normals", {s.normals.size(), 3}, (const float*)s.normals.data());
      if (!s.texcoords.empty())
        to.add_f32(pre + "texcoords", {s.texcoords.size(), 2}, (const float*)s.texcoords.data());
      int params[5] = {s.subdivisions, s.catmullclark ? 1 : 0, s.smooth ? 1 : 0, s.displacement_tex, s.shape};
      to.add_i32(pre + "params", {5}, params);
      to.add_f32(pre + "displacement", {1}, &s.displacement);
    }
  };
  if (!dump_subdivs.empty()) {
    auto only = vpt::archive{};
    put_subdivs(only);
    if (!only.save(dump_subdivs, error)) {
      fprintf(stderr, "vpt_export: %s\n", error.c_str());
      return 1;
    }
  }
  if (!keep_subdivs) tesselate_surfaces(scene);

  auto ar = vpt::archive{};
  if (keep_subdivs) put_subdivs(ar);
  int  meta[4] = {1, 0, 0, 0};
  ar.add_i32("meta", {4}, meta);

  // cameras (yocto_scene.h:84-92)
  {
    auto v = vector<float>(scene.cameras.size() * 18);
    for (size_t i = 0; i < scene.cameras.size(); i++) {
      auto& c = scene.cameras[i];
      auto  p = v.data() + i * 18;
      put_frame(p, c.frame);
      p[12] = c.orthographic ? 1.0f : 0.0f;
      p[13] = c.lens;
      p[14] = c.film;
      p[15] = c.aspect;
      p[16] = c.focus;
      p[17] = c.aperture;
    }
    ar.add_f32("cameras", {scene.cameras.size(), 18}, v.data());
  }
  // instances (yocto_scene.h:142-148)
  {
    auto f  = vector<float>(scene.instances.size() * 12);
    auto id = vector<int>(scene.instances.size() * 2);
    for (size_t i = 0; i < scene.instances.size(); i++) {
      put_frame(f.data() + i * 12, scene.instances[i].frame);
      id[i * 2 + 0] = scene.instances[i].shape;
      id[i * 2 + 1] = scene.instances[i].material;
    }
    ar.add_f32("instances.frame", {scene.instances.size(), 12}, f.data());
    ar.add_i32("instances.ids", {scene.instances.size(), 2}, id.data());
  }
  // materials (yocto_scene.h:119-139)
  {
    auto f  = vector<float>(scene.materials.size() * 16, 0.0f);
    auto id = vector<int>(scene.materials.size() * 6);
    for (size_t i = 0; i < scene.materials.size(); i++) {
      auto& m = scene.materials[i];
      auto  p = f.data() + i * 16;
      p[0] = m.emission.x, p[1] = m.emission.y, p[2] = m.emission.z;
      p[3] = m.color.x, p[4] = m.color.y, p[5] = m.color.z;
      p[6] = m.roughness, p[7] = m.metallic, p[8] = m.ior;
      p[9] = m.scattering.x, p[10] = m.scattering.y, p[11] = m.scattering.z;
      p[12] = m.scanisotropy, p[13] = m.trdepth, p[14] = m.opacity;
      auto q = id.data() + i * 6;
      q[0] = (int)m.type, q[1] = m.emission_tex, q[2] = m.color_tex;
      q[3] = m.roughness_tex, q[4] = m.scattering_tex, q[5] = m.normal_tex;
    }
    ar.add_f32("materials.f", {scene.materials.size(), 16}, f.data());
    ar.add_i32("materials.i", {scene.materials.size(), 6}, id.data());
  }
  // environments (yocto_scene.h:151-156)
  {
    auto f  = vector<float>(scene.environments.size() * 15);
    auto id = vector<int>(scene.environments.size());
    for (size_t i = 0; i < scene.environments.size(); i++) {
      auto& e = scene.environments[i];
      auto  p = f.data() + i * 15;
      put_frame(p, e.frame);
      p[12] = e.emission.x, p[13] = e.emission.y, p[14] = e.emission.z;
      id[i] = e.emission_tex;
    }
    ar.add_f32("environments.f", {scene.environments.size(), 15}, f.data());
    ar.add_i32("environments.i", {scene.environments.size(), 1}, id.data());
  }
  // shapes (yocto_shape.h:74)
  {
    int n = (int)scene.shapes.size();
    ar.add_i32("shapes.count", {1}, &n);
    for (int i = 0; i < n; i++) {
      auto& s   = scene.shapes[i];
      auto  pre = "shape" + std::to_string(i) + ".";
      ar.add_f32(pre + "positions", {s.positions.size(), 3},
          (const float*)s.positions.data());
      if (!s.normals.empty())
        ar.add_f32(pre + "normals", {s.normals.size(), 3},
            (const float*)s.normals.data());
      if (!s.texcoords.empty())
        ar.add_f32(pre + "texcoords", {s.texcoords.size(), 2},
            (const float*)s.texcoords.data());
      if (!s.colors.empty())
        ar.add_f32(pre + "colors", {s.colors.size(), 4},
            (const float*)s.colors.data());
      if (!s.triangles.empty())
        ar.add_i32(pre + "triangles", {s.triangles.size(), 3},
            (const int*)s.triangles.data());
      if (!s.quads.empty())
        ar.add_i32(
            pre + "quads", {s.quads.size(), 4}, (const int*)s.quads.data());
      int unsupported[2] = {(int)s.points.size(), (int)s.lines.size()};
      ar.add_i32(pre + "unsupported", {2}, unsupported);
    }
  }
  // textures (yocto_scene.h:96-102)
  {
    int n = (int)scene.textures.size();
    ar.add_i32("textures.count", {1}, &n);
    for (int i = 0; i < n; i++) {
      auto& t   = scene.textures[i];
      auto  pre = "tex" + std::to_string(i) + ".";
      int   m[3] = {t.width, t.height, t.linear ? 1 : 0};
      ar.add_i32(pre + "meta", {3}, m);
      if (!t.pixelsf.empty())
        ar.add_f32(pre + "pixelsf", {(uint64_t)t.height, (uint64_t)t.width, 4},
            (const float*)t.pixelsf.data());
      else if (!t.pixelsb.empty())
        ar.add(pre + "pixelsb", vpt::dt_u8,
            {(uint64_t)t.height, (uint64_t)t.width, 4}, t.pixelsb.data());
    }
  }
  // volumes / vol_instances (yocto_scene.h:203-219)
  {
    int n = (int)scene.volumes.size();
    ar.add_i32("volumes.count", {1}, &n);
    for (int i = 0; i < n; i++) {
      auto& v   = scene.volumes[i];
      auto  pre = "vol" + std::to_string(i) + ".";
      int   whd[3] = {v.whd.x, v.whd.y, v.whd.z};
      ar.add_i32(pre + "whd", {3}, whd);
      ar.add_f32(pre + "res", {1}, &v.res);
      ar.add_f32(pre + "data",
          {(uint64_t)v.whd.z, (uint64_t)v.whd.y, (uint64_t)v.whd.x},
          v.vol.data());
    }
    auto f  = vector<float>(scene.vol_instances.size() * 13);
    auto id = vector<int>(scene.vol_instances.size() * 2);
    for (size_t i = 0; i < scene.vol_instances.size(); i++) {
      auto& vi = scene.vol_instances[i];
      put_frame(f.data() + i * 13, vi.frame);
      f[i * 13 + 12] = vi.scalef;
      id[i * 2 + 0]  = vi.volume;
      id[i * 2 + 1]  = vi.material;
    }
    ar.add_f32("vol_instances.f", {scene.vol_instances.size(), 13}, f.data());
    ar.add_i32("vol_instances.i", {scene.vol_instances.size(), 2}, id.data());
  }
  // analytic SDFs: sdf_data keeps only a std::function closure
  // (yocto_sceneio.cpp:3685-3728), so the typed parameters are re-read from
  // the JSON and cross-checked against the closure at probe points.
  {
    auto f  = vector<float>(scene.sdfs.size() * 19, 0.0f);
    auto id = vector<int>(scene.sdfs.size() * 2);
    if (!scene.sdfs.empty()) {
      auto fs = std::ifstream(filename);
      auto ss = std::stringstream{};
      ss << fs.rdbuf();
      auto js    = nlohmann::ordered_json::parse(ss.str());
      auto group = js.at("sdfunctions");
      auto names = vector<string>{
          "bbox", "box", "capped_cone", "plane", "sphere", "torus"};
      size_t i = 0;
      for (auto& el : group) {
        auto& sdf  = scene.sdfs[i];
        auto  p    = f.data() + i * 19;
        auto  type = -1;
        auto  tname = el.value("type", string{});
        for (int k = 0; k < 6; k++)
          if (names[k] == tname) type = k;
        if (type < 0) {
          fprintf(stderr, "vpt_export: unknown sdf type %s\n", tname.c_str());
          return 1;
        }
        put_frame(p, sdf.frame);
        auto getf = [&](const char* key) { return el.value(key, 0.0f); };
        auto getv = [&](const char* key) {
          auto v = vec3f{0, 0, 0};
          if (el.contains(key)) {
            auto a = el.at(key).get<std::array<float, 3>>();
            v      = {a[0], a[1], a[2]};
          }
          return v;
        };
        auto typed = std::function<float(const vec3f&)>{};
        switch ((sdf_type)type) {
          case sdf_type::bbox: {
            auto th = getf("thickness");
            auto w  = getv("whd");
            p[12] = th, p[13] = w.x, p[14] = w.y, p[15] = w.z;
            typed = [=](const vec3f& q) { return sd_bbox(q, w, th); };
          } break;
          case sdf_type::box: {
            auto w = getv("whd");
            p[16] = w.x, p[17] = w.y, p[18] = w.z;
            typed = [=](const vec3f& q) {
              return sd_box(q - (w * 0.5f), w * 0.5f);
            };
          } break;
          case sdf_type::capped_cone: {
            auto h = getf("height"), r1 = getf("r1"), r2 = getf("r2");
            p[12] = h, p[13] = r1, p[14] = r2;
            typed = [=](const vec3f& q) { return sd_capped_cone(q, h, r1, r2); };
          } break;
          case sdf_type::plane: {
            typed = [](const vec3f& q) { return sd_plane(q); };
          } break;
          case sdf_type::sphere: {
            auto r = getf("radius");
            p[12]  = r;
            typed  = [=](const vec3f& q) { return sd_sphere(q, r); };
          } break;
          case sdf_type::torus: {
            auto r1 = getf("r1"), r2 = getf("r2");
            p[12] = r1, p[13] = r2;
            typed = [=](const vec3f& q) { return sd_torus(q, r1, r2); };
          } break;
        }
        // probe the closure: typed parameters must reproduce it bit for bit
        for (int k = 0; k < 64; k++) {
          auto q = vec3f{0.37f * (k % 4) - 0.5f, 0.21f * ((k / 4) % 4) - 0.3f,
              0.45f * (k / 16) - 0.6f};
          auto a = sdf.f(q), b = typed(q);
          if (memcmp(&a, &b, 4) != 0) {
            fprintf(stderr, "vpt_export: sdf %zu typed/closure mismatch\n", i);
            return 1;
          }
        }
        id[i * 2 + 0] = type;
        id[i * 2 + 1] = sdf.material;
        i++;
      }
    }
    ar.add_f32("sdfs.f", {scene.sdfs.size(), 19}, f.data());
    ar.add_i32("sdfs.i", {scene.sdfs.size(), 2}, id.data());
  }

  if (!ar.save(argv[2], error)) {
    fprintf(stderr, "vpt_export: %s\n", error.c_str());
    return 1;
  }
  printf("vpt_export: %s -> %s (%zu arrays; %zu shapes, %zu instances, "
         "%zu textures, %zu volumes, %zu sdfs)\n",
      argv[1], argv[2], ar.arrays.size(), scene.shapes.size(),
      scene.instances.size(), scene.textures.size(), scene.volumes.size(),
      scene.sdfs.size());
  return 0;
}
