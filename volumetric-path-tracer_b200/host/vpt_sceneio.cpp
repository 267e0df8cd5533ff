// vpt_sceneio.cpp — ".vptscene" archive -> vpt::scene_data.
// Array names and layouts are the ones tools/vpt_export.cpp writes from
// the reference's scene_data and archive.py writes from numpy fixtures.
#include "vpt_archive.h"
#include "vpt_scene.h"

namespace vpt {

static frame3f get_frame(const float* p) {
  return {{p[0], p[1], p[2]}, {p[3], p[4], p[5]}, {p[6], p[7], p[8]}, {p[9], p[10], p[11]}};
}

template <typename T>
static void assign(std::vector<T>& dst, const archive_array* a, size_t per_item_words) {
  if (!a) return;
  auto n = a->count() / per_item_words;
  dst.resize(n);
  if (n) memcpy(dst.data(), a->data.data(), n * sizeof(T));
}

// subdivs.count + subdiv<i>.{quadspos, quadsnorm, quadstexcoord, positions, normals, texcoords,
// params = {subdivisions, catmullclark, smooth, displacement_tex, shape}, displacement}
static bool read_subdivs(const archive& ar, scene_data& scene) {
  scene.subdivs.clear();
  auto count = ar.find("subdivs.count");
  if (!count || count->dtype != dt_i32) return true;
  for (int i = 0; i < count->i32()[0]; i++) {
    auto  pre = "subdiv" + std::to_string(i) + ".";
    auto& s   = scene.subdivs.emplace_back();
    assign(s.quadspos, ar.find(pre + "quadspos"), 4);
    assign(s.quadsnorm, ar.find(pre + "quadsnorm"), 4);
    assign(s.quadstexcoord, ar.find(pre + "quadstexcoord"), 4);
    assign(s.positions, ar.find(pre + "positions"), 3);
    assign(s.normals, ar.find(pre + "normals"), 3);
    assign(s.texcoords, ar.find(pre + "texcoords"), 2);
    auto params = ar.find(pre + "params"), disp = ar.find(pre + "displacement");
    if (!params || params->dtype != dt_i32 || params->count() < 5 || !disp) return false;
    auto q = params->i32();
    s.subdivisions = q[0], s.catmullclark = q[1] != 0, s.smooth = q[2] != 0;
    s.displacement_tex = q[3], s.shape = q[4];
    s.displacement = disp->f32()[0];
  }
  return true;
}

bool load_subdivs(const std::string& filename, scene_data& scene, std::string& error) {
  auto ar = archive{};
  if (!ar.load(filename, error)) return false;
  if (!read_subdivs(ar, scene)) return error = "malformed subdiv arrays in " + filename, false;
  return true;
}

bool load_scene(const std::string& filename, scene_data& scene, std::string& error) {
  auto ar = archive{};
  if (!ar.load(filename, error)) return false;
  scene = scene_data{};
  if (!read_subdivs(ar, scene)) return error = "malformed subdiv arrays in " + filename, false;
  auto missing = [&](const char* name) {
    error = std::string("missing array ") + name + " in " + filename;
    return false;
  };
  auto expect = [&](const archive_array* a, uint32_t dtype, size_t inner) {
    return a && a->dtype == dtype && a->dims.size() == 2 && a->dims[1] == inner;
  };

  auto cams = ar.find("cameras");
  if (!expect(cams, dt_f32, 18)) return missing("cameras");
  for (size_t i = 0; i < cams->dims[0]; i++) {
    auto  p = cams->f32() + i * 18;
    auto& c = scene.cameras.emplace_back();
    c.frame = get_frame(p), c.orthographic = p[12] != 0;
    c.lens = p[13], c.film = p[14], c.aspect = p[15], c.focus = p[16], c.aperture = p[17];
  }
  auto inf = ar.find("instances.frame"), ini = ar.find("instances.ids");
  if (!expect(inf, dt_f32, 12) || !expect(ini, dt_i32, 2)) return missing("instances.*");
  for (size_t i = 0; i < inf->dims[0]; i++) {
    auto& in    = scene.instances.emplace_back();
    in.frame    = get_frame(inf->f32() + i * 12);
    in.shape    = ini->i32()[i * 2 + 0];
    in.material = ini->i32()[i * 2 + 1];
  }
  auto mf = ar.find("materials.f"), mi = ar.find("materials.i");
  if (!expect(mf, dt_f32, 16) || !expect(mi, dt_i32, 6)) return missing("materials.*");
  for (size_t i = 0; i < mf->dims[0]; i++) {
    auto  p = mf->f32() + i * 16;
    auto  q = mi->i32() + i * 6;
    auto& m = scene.materials.emplace_back();
    m.type  = (material_type)q[0];
    m.emission = {p[0], p[1], p[2]}, m.color = {p[3], p[4], p[5]};
    m.roughness = p[6], m.metallic = p[7], m.ior = p[8];
    m.scattering = {p[9], p[10], p[11]};
    m.scanisotropy = p[12], m.trdepth = p[13], m.opacity = p[14];
    m.emission_tex = q[1], m.color_tex = q[2], m.roughness_tex = q[3], m.scattering_tex = q[4],
    m.normal_tex = q[5];
  }
  auto ef = ar.find("environments.f"), ei = ar.find("environments.i");
  if (!expect(ef, dt_f32, 15) || !expect(ei, dt_i32, 1)) return missing("environments.*");
  for (size_t i = 0; i < ef->dims[0]; i++) {
    auto  p = ef->f32() + i * 15;
    auto& e = scene.environments.emplace_back();
    e.frame = get_frame(p), e.emission = {p[12], p[13], p[14]}, e.emission_tex = ei->i32()[i];
  }
  auto sc = ar.find("shapes.count");
  if (!sc || sc->dtype != dt_i32) return missing("shapes.count");
  for (int i = 0; i < sc->i32()[0]; i++) {
    auto& s   = scene.shapes.emplace_back();
    auto  pre = "shape" + std::to_string(i) + ".";
    assign(s.positions, ar.find(pre + "positions"), 3);
    assign(s.normals, ar.find(pre + "normals"), 3);
    assign(s.texcoords, ar.find(pre + "texcoords"), 2);
    assign(s.colors, ar.find(pre + "colors"), 4);
    assign(s.triangles, ar.find(pre + "triangles"), 3);
    assign(s.quads, ar.find(pre + "quads"), 4);
    if (auto u = ar.find(pre + "unsupported")) {
      s.points.resize(u->i32()[0]);
      s.lines.resize(2 * (size_t)u->i32()[1]);
    }
  }
  auto tc = ar.find("textures.count");
  if (!tc || tc->dtype != dt_i32) return missing("textures.count");
  for (int i = 0; i < tc->i32()[0]; i++) {
    auto& t   = scene.textures.emplace_back();
    auto  pre = "tex" + std::to_string(i) + ".";
    auto  m   = ar.find(pre + "meta");
    if (!m) return missing((pre + "meta").c_str());
    t.width = m->i32()[0], t.height = m->i32()[1], t.linear = m->i32()[2] != 0;
    assign(t.pixelsf, ar.find(pre + "pixelsf"), 4);
    if (auto a = ar.find(pre + "pixelsb")) {
      t.pixelsb.resize(a->count() / 4);
      if (!t.pixelsb.empty()) memcpy(t.pixelsb.data(), a->data.data(), a->data.size());
    }
  }
  if (auto vc = ar.find("volumes.count")) {
    for (int i = 0; i < vc->i32()[0]; i++) {
      auto& v   = scene.volumes.emplace_back();
      auto  pre = "vol" + std::to_string(i) + ".";
      auto  whd = ar.find(pre + "whd"), res = ar.find(pre + "res"), dat = ar.find(pre + "data");
      if (!whd || !res || !dat) return missing((pre + "*").c_str());
      v.whd = {whd->i32()[0], whd->i32()[1], whd->i32()[2]};
      v.res = res->f32()[0];
      v.vol.assign(dat->f32(), dat->f32() + dat->count());
      if (v.vol.size() != (size_t)v.whd.x * v.whd.y * v.whd.z) {
        error = "volume " + std::to_string(i) + " size mismatch in " + filename;
        return false;
      }
    }
  }
  if (auto vf = ar.find("vol_instances.f")) {
    auto vi = ar.find("vol_instances.i");
    if (!expect(vf, dt_f32, 13) || !expect(vi, dt_i32, 2)) return missing("vol_instances.*");
    for (size_t i = 0; i < vf->dims[0]; i++) {
      auto& in    = scene.vol_instances.emplace_back();
      in.frame    = get_frame(vf->f32() + i * 13);
      in.scalef   = vf->f32()[i * 13 + 12];
      in.volume   = vi->i32()[i * 2 + 0];
      in.material = vi->i32()[i * 2 + 1];
    }
  }
  if (auto sf = ar.find("sdfs.f")) {
    auto si = ar.find("sdfs.i");
    if (!expect(sf, dt_f32, 19) || !expect(si, dt_i32, 2)) return missing("sdfs.*");
    for (size_t i = 0; i < sf->dims[0]; i++) {
      auto  p      = sf->f32() + i * 19;
      auto& sdf    = scene.sdfs.emplace_back();
      sdf.frame    = get_frame(p);
      sdf.type     = (sdf_type)si->i32()[i * 2 + 0];
      sdf.material = si->i32()[i * 2 + 1];
      for (int k = 0; k < 4; k++) sdf.params[k] = p[12 + k];
      sdf.whd = {p[16], p[17], p[18]};
    }
  }
  return true;
}

}  // namespace vpt
