// yocto_pathtrace_cuda.cpp — the reference-side adapter of INTEGRATION.md, written out in full.
//
// Compiled AGAINST THE REFERENCE'S OWN HEADERS (yocto_pathtrace.h:119-139, scene_data,
// bvh_scene, pathtrace_lights) and linked with libvpt_cuda.so. It defines
//     yocto::pathtrace_samples   (replaces yocto_pathtrace.cpp:1052-1092)
//     yocto::get_render x2       (replaces yocto_pathtrace.cpp:1105-1116)
// with the reference's signatures; make_state / make_bvh / make_lights / tesselate_surfaces stay
// the reference's. What a maintainer would add to libs/yocto_pathtrace/ and build into
// `yocto_pathtrace` instead of the two CPU bodies; here it is built as an integration test
// target (oracle/Makefile: the reference's yocto_pathtrace.cpp is compiled with the two CPU
// bodies renamed, never edited on disk) together with the reference's loaders and a headless
// run_offline (integration/ypathtrace_headless.cpp).
//
// TEST / INTEGRATION TARGET: like oracle/, nothing under volumetric-path-tracer_b200/ builds or
// links this file. It exists only where /root/reference does (its headers are included).
#include <yocto/yocto_scene.h>
#include <yocto/yocto_sdfs.h>
#include <yocto_pathtrace/yocto_pathtrace.h>

#include <cstring>
#include <fstream>
#include <memory>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <unordered_map>

#include "json.hpp"  // the reference's own libs/yocto/ext/json.hpp
#include "vpt_cuda.h"
#include "yocto_pathtrace_cuda.h"

namespace yocto {

static_assert(sizeof(bvh_node) == sizeof(vpt_bvh_node), "bvh_node must stay 32 bytes");
static_assert(sizeof(rng_state) == 16, "rng_state = {uint64 state, uint64 inc}");
static_assert(sizeof(vec4f) == 16 && sizeof(frame3f) == 48, "plain float aggregates");

namespace {

void throw_on(int status) {  // the reference's error convention: std::runtime_error
  if (status != VPT_OK) throw std::runtime_error(vpt_cuda_last_error());
}

// Typed parameters of the analytic SDFs. sdf_data keeps only a std::function closure
// (yocto_sceneio.cpp:3685-3728); {type, params} are re-read from the scene JSON and every entry
// is cross-checked against the closure at probe points, bit for bit.
struct sdf_typed {
  int   type = -1;
  float params[4] = {0, 0, 0, 0};
  float whd[3]    = {0, 0, 0};
};

std::mutex                    g_mutex;
std::string                   g_scene_file;
std::vector<sdf_typed>        g_sdf_cache;
const scene_data*             g_sdf_cache_scene = nullptr;

std::vector<sdf_typed> read_sdf_params(const std::string& filename, const scene_data& scene) {
  auto result = std::vector<sdf_typed>{};
  if (scene.sdfs.empty()) return result;
  if (filename.empty())
    throw std::runtime_error(
        "analytic SDF parameters unknown: call pathtrace_cuda_scene_file(path) after load_scene");
  auto fs = std::ifstream(filename);
  if (!fs) throw std::runtime_error("cannot re-read " + filename);
  auto ss = std::stringstream{};
  ss << fs.rdbuf();
  auto js = nlohmann::ordered_json::parse(ss.str());
  if (!js.contains("sdfunctions") || js.at("sdfunctions").size() != scene.sdfs.size())
    throw std::runtime_error("sdfunctions of " + filename + " do not match the loaded scene");
  const char* names[6] = {"bbox", "box", "capped_cone", "plane", "sphere", "torus"};
  size_t      i        = 0;
  for (auto& el : js.at("sdfunctions")) {
    auto& sdf   = scene.sdfs[i];
    auto  t     = sdf_typed{};
    auto  tname = el.value("type", std::string{});
    for (int k = 0; k < 6; k++)
      if (tname == names[k]) t.type = k;
    if (t.type < 0) throw std::runtime_error("unknown sdf type " + tname);
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
    switch ((sdf_type)t.type) {
      case sdf_type::bbox: {
        auto th = getf("thickness");
        auto w  = getv("whd");
        t.params[0] = th, t.params[1] = w.x, t.params[2] = w.y, t.params[3] = w.z;
        typed = [=](const vec3f& q) { return sd_bbox(q, w, th); };
      } break;
      case sdf_type::box: {
        auto w = getv("whd");
        t.whd[0] = w.x, t.whd[1] = w.y, t.whd[2] = w.z;
        typed = [=](const vec3f& q) { return sd_box(q - (w * 0.5f), w * 0.5f); };
      } break;
      case sdf_type::capped_cone: {
        auto h = getf("height"), r1 = getf("r1"), r2 = getf("r2");
        t.params[0] = h, t.params[1] = r1, t.params[2] = r2;
        typed = [=](const vec3f& q) { return sd_capped_cone(q, h, r1, r2); };
      } break;
      case sdf_type::plane: typed = [](const vec3f& q) { return sd_plane(q); }; break;
      case sdf_type::sphere: {
        auto r      = getf("radius");
        t.params[0] = r;
        typed       = [=](const vec3f& q) { return sd_sphere(q, r); };
      } break;
      case sdf_type::torus: {
        auto r1 = getf("r1"), r2 = getf("r2");
        t.params[0] = r1, t.params[1] = r2;
        typed = [=](const vec3f& q) { return sd_torus(q, r1, r2); };
      } break;
    }
    for (int k = 0; k < 64; k++) {
      auto q = vec3f{0.37f * (k % 4) - 0.5f, 0.21f * ((k / 4) % 4) - 0.3f, 0.45f * (k / 16) - 0.6f};
      auto a = sdf.f(q), b = typed(q);
      if (memcmp(&a, &b, 4) != 0)
        throw std::runtime_error("sdf " + std::to_string(i) + ": JSON parameters do not reproduce sdf.f");
    }
    result.push_back(t);
    i++;
  }
  return result;
}

// Everything the C ABI takes, as pointers into the caller's own vectors: vec3f / vec2f / vec4f /
// vec3i / vec4i / frame3f are plain float / int arrays and bvh_node is byte-compatible with
// vpt_bvh_node, so nothing is copied except the small per-object structs.
struct desc_storage {
  std::vector<vpt_camera_desc>          cameras;
  std::vector<vpt_instance_desc>        instances;
  std::vector<vpt_shape_desc>           shapes;
  std::vector<vpt_material_desc>        materials;
  std::vector<vpt_texture_desc>         textures;
  std::vector<vpt_environment_desc>     environments;
  std::vector<vpt_volume_desc>          volumes;
  std::vector<vpt_volume_instance_desc> vol_instances;
  std::vector<vpt_sdf_desc>             sdfs;
  std::vector<vpt_light_desc>           lights;
  std::vector<vpt_bvh_desc>             shape_bvhs;
  vpt_scene_desc                        desc{};
};

vpt_bvh_desc describe_bvh(const bvh_data& bvh) {
  return {(const vpt_bvh_node*)bvh.nodes.data(), (int32_t)bvh.nodes.size(), bvh.primitives.data(),
      (int32_t)bvh.primitives.size()};
}

void describe(desc_storage& s, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, const std::vector<sdf_typed>& typed) {
  if (bvh.embree_bvh) throw std::runtime_error("Embree BVHs have no node array to upload");
  if (bvh.shapes.size() != scene.shapes.size())
    throw std::runtime_error("bvh does not match the scene");
  for (auto& c : scene.cameras) {
    auto d = vpt_camera_desc{};
    memcpy(d.frame, &c.frame, 48);
    d.orthographic = c.orthographic ? 1 : 0;
    d.lens = c.lens, d.film = c.film, d.aspect = c.aspect, d.focus = c.focus, d.aperture = c.aperture;
    s.cameras.push_back(d);
  }
  for (auto& i : scene.instances) {
    auto d = vpt_instance_desc{};
    memcpy(d.frame, &i.frame, 48);
    d.shape = i.shape, d.material = i.material;
    s.instances.push_back(d);
  }
  for (auto& sh : scene.shapes) {
    auto d = vpt_shape_desc{};
    d.positions = (const float*)sh.positions.data(), d.num_positions = (int)sh.positions.size();
    d.normals = sh.normals.empty() ? nullptr : (const float*)sh.normals.data();
    d.num_normals = (int)sh.normals.size();
    d.texcoords = sh.texcoords.empty() ? nullptr : (const float*)sh.texcoords.data();
    d.num_texcoords = (int)sh.texcoords.size();
    d.colors = sh.colors.empty() ? nullptr : (const float*)sh.colors.data();
    d.num_colors = (int)sh.colors.size();
    d.triangles = (const int32_t*)sh.triangles.data(), d.num_triangles = (int)sh.triangles.size();
    d.quads = (const int32_t*)sh.quads.data(), d.num_quads = (int)sh.quads.size();
    d.num_points = (int)sh.points.size(), d.num_lines = (int)sh.lines.size();
    s.shapes.push_back(d);
  }
  for (auto& m : scene.materials) {
    auto d = vpt_material_desc{};
    d.type = (int)m.type;
    memcpy(d.emission, &m.emission, 12), memcpy(d.color, &m.color, 12);
    d.roughness = m.roughness, d.metallic = m.metallic, d.ior = m.ior;
    memcpy(d.scattering, &m.scattering, 12);
    d.scanisotropy = m.scanisotropy, d.trdepth = m.trdepth, d.opacity = m.opacity;
    d.emission_tex = m.emission_tex, d.color_tex = m.color_tex, d.roughness_tex = m.roughness_tex;
    d.scattering_tex = m.scattering_tex, d.normal_tex = m.normal_tex;
    s.materials.push_back(d);
  }
  for (auto& t : scene.textures)
    s.textures.push_back({t.width, t.height, t.linear ? 1 : 0,
        t.pixelsf.empty() ? nullptr : (const float*)t.pixelsf.data(),
        t.pixelsb.empty() ? nullptr : (const uint8_t*)t.pixelsb.data()});
  for (auto& e : scene.environments) {
    auto d = vpt_environment_desc{};
    memcpy(d.frame, &e.frame, 48), memcpy(d.emission, &e.emission, 12);
    d.emission_tex = e.emission_tex;
    s.environments.push_back(d);
  }
  for (auto& v : scene.volumes)
    s.volumes.push_back({{v.whd.x, v.whd.y, v.whd.z}, v.res, v.vol.data()});
  for (auto& v : scene.vol_instances) {
    auto d = vpt_volume_instance_desc{};
    memcpy(d.frame, &v.frame, 48);
    d.scalef = v.scalef, d.volume = v.volume, d.material = v.material;
    s.vol_instances.push_back(d);
  }
  for (size_t i = 0; i < scene.sdfs.size(); i++) {
    auto d = vpt_sdf_desc{};
    memcpy(d.frame, &scene.sdfs[i].frame, 48);
    d.type = typed[i].type, d.material = scene.sdfs[i].material;
    memcpy(d.params, typed[i].params, 16), memcpy(d.whd, typed[i].whd, 12);
    s.sdfs.push_back(d);
  }
  for (auto& l : lights.lights)
    s.lights.push_back({l.instance, l.environment, l.sdf, l.elements_cdf.data(),
        (int32_t)l.elements_cdf.size()});
  for (auto& b : bvh.shapes) s.shape_bvhs.push_back(describe_bvh(b));
  auto& d = s.desc;
  d.num_cameras = (int)s.cameras.size(), d.cameras = s.cameras.data();
  d.num_instances = (int)s.instances.size(), d.instances = s.instances.data();
  d.num_shapes = (int)s.shapes.size(), d.shapes = s.shapes.data();
  d.num_materials = (int)s.materials.size(), d.materials = s.materials.data();
  d.num_textures = (int)s.textures.size(), d.textures = s.textures.data();
  d.num_environments = (int)s.environments.size(), d.environments = s.environments.data();
  d.num_volumes = (int)s.volumes.size(), d.volumes = s.volumes.data();
  d.num_vol_instances = (int)s.vol_instances.size(), d.vol_instances = s.vol_instances.data();
  d.num_sdfs = (int)s.sdfs.size(), d.sdfs = s.sdfs.data();
  d.num_lights = (int)s.lights.size(), d.lights = s.lights.data();
  d.scene_bvh = describe_bvh(bvh), d.shape_bvhs = s.shape_bvhs.data();
}

// Device mirrors, keyed by the caller's objects (the reference's structs are not ours to extend):
// one uploaded scene and one device state per pathtrace_state.
struct device_binding {
  vpt_scene*        scene = nullptr;
  vpt_state*        state = nullptr;
  const scene_data* scene_key = nullptr;
  const bvh_scene*  bvh_key = nullptr;
  ~device_binding() {
    if (state) vpt_cuda_free_state(state);
    if (scene) vpt_cuda_free_scene(scene);
  }
};
std::unordered_map<const pathtrace_state*, std::unique_ptr<device_binding>> g_bindings;

device_binding& binding_of(const pathtrace_state& state) {
  auto& slot = g_bindings[&state];
  if (!slot) slot = std::make_unique<device_binding>();
  return *slot;
}

}  // namespace

void pathtrace_cuda_scene_file(const std::string& filename) {
  auto lock         = std::lock_guard{g_mutex};
  g_scene_file      = filename;
  g_sdf_cache_scene = nullptr;
}

void pathtrace_cuda_release(const pathtrace_state& state) {
  auto lock = std::lock_guard{g_mutex};
  g_bindings.erase(&state);
}

// Progressively computes an image: one sample per call, as yocto_pathtrace.cpp:1052-1092.
// The accumulation image and the per-pixel generators live on the device between calls;
// state.samples is kept, state.image / state.hits / state.rngs are refreshed by get_render
// (image) or pathtrace_cuda_sync (all three), not per sample.
void pathtrace_samples(pathtrace_state& state, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, const pathtrace_params& params) {
  if (state.samples >= params.samples) return;  // :1055
  auto  lock = std::lock_guard{g_mutex};
  auto& dev  = binding_of(state);
  // the state is being reused with another scene: drop the mirrors
  if (dev.scene && (dev.scene_key != &scene || dev.bvh_key != &bvh)) g_bindings.erase(&state);
  pathtrace_samples_unlocked(state, scene, bvh, lights, params);
}

void pathtrace_samples_unlocked(pathtrace_state& state, const scene_data& scene,
    const bvh_scene& bvh, const pathtrace_lights& lights, const pathtrace_params& params) {
  auto& dev = binding_of(state);
  if (!dev.scene) {
    if (g_sdf_cache_scene != &scene) {
      g_sdf_cache       = read_sdf_params(g_scene_file, scene);
      g_sdf_cache_scene = &scene;
    }
    auto storage = desc_storage{};
    describe(storage, scene, bvh, lights, g_sdf_cache);
    throw_on(vpt_cuda_upload_scene(&storage.desc, &dev.scene));
    dev.scene_key = &scene, dev.bvh_key = &bvh;
  }
  if (!dev.state || state.samples == 0) {
    // a fresh (or reset) pathtrace_state: its generators are make_state's (:975-978)
    if ((int)state.rngs.size() != state.width * state.height)
      throw std::invalid_argument{"state should come from make_state"};
    if (!dev.state)
      throw_on(vpt_cuda_create_state(state.width, state.height,
          (const uint64_t*)state.rngs.data(), &dev.state));
    else
      throw_on(vpt_cuda_reset_state(dev.state, (const uint64_t*)state.rngs.data()));
  }
  auto p                = vpt_params{};
  p.camera              = params.camera;
  p.resolution          = params.resolution;
  p.shader              = (int)params.shader;
  p.samples             = params.samples;
  p.bounces             = params.bounces;
  p.noparallel          = params.noparallel ? 1 : 0;
  p.pratio              = params.pratio;
  p.exposure            = params.exposure;
  p.filmic              = params.filmic ? 1 : 0;
  p.noimplicit_mis      = params.noimplicit_mis ? 1 : 0;
  p.spheretrace_maxiter = params.spheretrace_maxiter;
  p.pipeline            = VPT_PIPELINE_AUTO;
  throw_on(vpt_cuda_trace_samples(dev.scene, dev.state, &p, 1));  // get_shader's throw: :948
  state.samples += 1;
}

// Copies the device sums, hit counts and generators back into the caller's pathtrace_state
// (what the reference's loop leaves there after every sample).
void pathtrace_cuda_sync(pathtrace_state& state) {
  auto lock = std::lock_guard{g_mutex};
  auto it   = g_bindings.find(&state);
  if (it == g_bindings.end() || !it->second->state) return;
  auto n = (size_t)state.width * state.height;
  state.image.resize(n), state.hits.resize(n), state.rngs.resize(n);
  throw_on(vpt_cuda_download_image(it->second->state, &state.image[0].x, n * 4));
  throw_on(vpt_cuda_download_rngs(it->second->state, (uint64_t*)state.rngs.data(), n * 2));
  for (auto& h : state.hits) h = state.samples;  // hits[idx] += 1 per sample (:1089), never read
}

static void check_image(const color_image& image, int width, int height, bool linear) {
  if (image.width != width || image.height != height)
    throw std::invalid_argument{"image should have the same size"};
  if (image.linear != linear)
    throw std::invalid_argument{linear ? "expected linear image" : "expected srgb image"};
}

// Get resulting render (yocto_pathtrace.cpp:1105-1116): sums * (1 / samples)
color_image get_render(const pathtrace_state& state) {
  auto image = make_image(state.width, state.height, true);
  get_render(image, state);
  return image;
}
void get_render(color_image& image, const pathtrace_state& state) {
  check_image(image, state.width, state.height, true);
  auto lock  = std::lock_guard{g_mutex};
  auto it    = g_bindings.find(&state);
  auto scale = 1.0f / (float)state.samples;
  if (it == g_bindings.end() || !it->second->state) {
    // nothing traced on the device yet: the host sums are the state
    for (auto idx = 0; idx < state.width * state.height; idx++)
      image.pixels[idx] = state.image[idx] * scale;
    return;
  }
  auto sums = std::vector<vec4f>((size_t)state.width * state.height);
  throw_on(vpt_cuda_download_image(it->second->state, &sums[0].x, sums.size() * 4));
  for (auto idx = 0; idx < state.width * state.height; idx++)
    image.pixels[idx] = sums[idx] * scale;
}

}  // namespace yocto
