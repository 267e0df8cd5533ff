// vpt_pathtrace.cpp — host half of the yocto_pathtrace API (see
// vpt_pathtrace.h): make_state, make_lights, get_render on the CPU as in the
// reference, and pathtrace_samples as a thin driver of the extern "C" CUDA
// layer (include/vpt_cuda.h). No rendering happens on the host.
#include "vpt_pathtrace.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>

#include <cmath>
#include <cstring>

namespace vpt {

// ---------------------------------------------------------------------------
// PCG32 as in yocto_sampling.h:184-205 (host copy, used for seeding only)
// ---------------------------------------------------------------------------
static uint32_t advance_rng(rng_state& rng) {
  uint64_t oldstate = rng.state;
  rng.state         = oldstate * 6364136223846793005ULL + rng.inc;
  auto xorshifted   = (uint32_t)(((oldstate >> 18u) ^ oldstate) >> 27u);
  auto rot          = (uint32_t)(oldstate >> 59u);
  return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
}
static rng_state make_rng(uint64_t seed, uint64_t seq = 1) {
  auto rng  = rng_state{};
  rng.state = 0U;
  rng.inc   = (seq << 1u) | 1u;
  advance_rng(rng);
  rng.state += seed;
  advance_rng(rng);
  return rng;
}

pathtrace_state make_state(const scene_data& scene, const pathtrace_params& params) {
  auto& camera = scene.cameras.at(params.camera);
  auto  state  = pathtrace_state{};
  if (camera.aspect >= 1) {
    state.width  = params.resolution;
    state.height = (int)std::round(params.resolution / camera.aspect);
  } else {
    state.height = params.resolution;
    state.width  = (int)std::round(params.resolution * camera.aspect);
  }
  state.samples = 0;
  state.image.assign((size_t)state.width * state.height, {0, 0, 0, 0});
  state.hits.assign((size_t)state.width * state.height, 0);
  state.rngs.assign((size_t)state.width * state.height, {});
  // one generator feeds every pixel's stream id, in pixel order
  // (yocto_pathtrace.cpp:975-978; rand1i(rng, 1 << 31) is a modulus of 2^31)
  auto rng_ = make_rng(1301081);
  for (auto& rng : state.rngs) {
    auto r = (int)(advance_rng(rng_) % 2147483648u);
    rng    = make_rng(961748941ull, (uint64_t)(r / 2 + 1));
  }
  return state;
}

// ---------------------------------------------------------------------------
// lights (yocto_pathtrace.cpp:983-1049)
// ---------------------------------------------------------------------------
static float triangle_area(const vec3f& p0, const vec3f& p1, const vec3f& p2) {
  return length(cross(p1 - p0, p2 - p0)) / 2;
}
static float quad_area(const vec3f& p0, const vec3f& p1, const vec3f& p2, const vec3f& p3) {
  return triangle_area(p0, p1, p3) + triangle_area(p2, p3, p1);
}

pathtrace_lights make_lights(const scene_data& scene, const pathtrace_params& params) {
  auto lights = pathtrace_lights{};
  for (auto handle = 0; handle < (int)scene.instances.size(); handle++) {
    auto& instance = scene.instances[handle];
    auto& material = scene.materials.at(instance.material);
    if (is_zero(material.emission)) continue;
    auto& shape = scene.shapes.at(instance.shape);
    if (shape.triangles.empty() && shape.quads.empty()) continue;
    auto& light    = lights.lights.emplace_back();
    light.instance = handle;
    if (!shape.triangles.empty()) {
      light.elements_cdf = std::vector<float>(shape.triangles.size());
      for (size_t idx = 0; idx < light.elements_cdf.size(); idx++) {
        auto& t                 = shape.triangles[idx];
        light.elements_cdf[idx] = triangle_area(
            shape.positions[t.x], shape.positions[t.y], shape.positions[t.z]);
        if (idx != 0) light.elements_cdf[idx] += light.elements_cdf[idx - 1];
      }
    }
    if (!shape.quads.empty()) {
      light.elements_cdf = std::vector<float>(shape.quads.size());
      for (size_t idx = 0; idx < light.elements_cdf.size(); idx++) {
        auto& t                 = shape.quads[idx];
        light.elements_cdf[idx] = quad_area(shape.positions[t.x], shape.positions[t.y],
            shape.positions[t.z], shape.positions[t.w]);
        if (idx != 0) light.elements_cdf[idx] += light.elements_cdf[idx - 1];
      }
    }
  }
  for (auto handle = 0; handle < (int)scene.environments.size(); handle++) {
    auto& environment = scene.environments[handle];
    if (is_zero(environment.emission)) continue;
    auto& light       = lights.lights.emplace_back();
    light.environment = handle;
    if (environment.emission_tex != invalidid) {
      auto& texture      = scene.textures.at(environment.emission_tex);
      light.elements_cdf = std::vector<float>((size_t)texture.width * texture.height);
      for (size_t idx = 0; idx < light.elements_cdf.size(); idx++) {
        auto i = (int)(idx % texture.width), j = (int)(idx / texture.width);
        auto th    = (j + 0.5f) * pif / texture.height;
        auto value = vec4f{0, 0, 0, 0};
        if (!texture.pixelsf.empty()) {
          value = texture.pixelsf[(size_t)j * texture.width + i];
        } else {
          auto b = texture.pixelsb[(size_t)j * texture.width + i];
          value  = {b.x / 255.0f, b.y / 255.0f, b.z / 255.0f, b.w / 255.0f};
        }
        auto vmax4 = vmax(vmax(vmax(value.x, value.y), value.z), value.w);
        light.elements_cdf[idx] = vmax4 * std::sin(th);
        if (idx != 0) light.elements_cdf[idx] += light.elements_cdf[idx - 1];
      }
    }
  }
  for (auto handle = 0; handle < (int)scene.sdfs.size(); ++handle) {
    auto& sdf      = scene.sdfs[handle];
    auto& material = scene.materials.at(sdf.material);
    if (is_zero(material.emission)) continue;
    auto& light = lights.lights.emplace_back();
    light.sdf   = handle;
    // flat xy box assumed: area = w * h
    light.elements_cdf = std::vector<float>{sdf.whd.x * sdf.whd.y};
  }
  return lights;
}

// ---------------------------------------------------------------------------
// device binding
// ---------------------------------------------------------------------------
struct device_scene {
  vpt_scene* handle = nullptr;
  ~device_scene() {
    if (handle) vpt_cuda_free_scene(handle);
  }
};

// scene replica + accumulation state of one more GPU (multi-GPU rendering, params.gpus > 1)
struct device_replica {
  int        device = 0;
  vpt_scene* scene  = nullptr;
  vpt_state* state  = nullptr;
  ~device_replica() {
    // free on the device that owns them
    int current = 0;
    if (scene || state) {
      current = vpt_cuda_current_device();
      vpt_cuda_set_device(device);
    }
    if (state) vpt_cuda_free_state(state);
    if (scene) vpt_cuda_free_scene(scene);
    if (scene || state) vpt_cuda_set_device(current);
  }
};

struct device_binding {
  std::shared_ptr<device_scene> scene = {};
  std::vector<std::unique_ptr<device_replica>> replicas = {};
  int         root_device = 0;
  // what every GPU's generators derive from: base_rngs advanced by rng_offset (+ rank << 40)
  std::shared_ptr<const std::vector<rng_state>> base_rngs;
  uint64_t    rng_offset  = 0;
  vpt_state*  state     = nullptr;
  const void* scene_key = nullptr;
  const void* bvh_key   = nullptr;
  const void* light_key = nullptr;
  bool        dirty     = false;  // device image is ahead of state.image
  bool        dirty_rngs = false; // device generators are ahead of state.rngs
  ~device_binding() {
    if (state) vpt_cuda_state_remove_peers(state);
    replicas.clear();
    if (state) vpt_cuda_free_state(state);
  }
};

static void check(int status, const char* what) {
  if (status != VPT_OK)
    throw std::runtime_error(std::string(what) + ": " + vpt_cuda_last_error());
}

void make_scene_desc(scene_desc_storage& st, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights) {
  if (bvh.shapes.size() != scene.shapes.size())
    throw std::runtime_error("bvh does not match the scene (shape count differs)");
  auto put_frame = [](float* dst, const frame3f& f) { memcpy(dst, &f, 48); };
  auto& cameras = st.cameras; auto& instances = st.instances; auto& shapes = st.shapes;
  auto& shape_bvhs = st.shape_bvhs; auto& materials = st.materials; auto& textures = st.textures;
  auto& environments = st.environments; auto& volumes = st.volumes;
  auto& vol_instances = st.vol_instances; auto& sdfs = st.sdfs; auto& ldesc = st.lights;
  auto& desc = st.desc;
  cameras = std::vector<vpt_camera_desc>(scene.cameras.size());
  for (size_t i = 0; i < cameras.size(); i++) {
    auto& c = scene.cameras[i];
    put_frame(cameras[i].frame, c.frame);
    cameras[i].orthographic = c.orthographic ? 1 : 0;
    cameras[i].lens = c.lens, cameras[i].film = c.film, cameras[i].aspect = c.aspect;
    cameras[i].focus = c.focus, cameras[i].aperture = c.aperture;
  }
  instances = std::vector<vpt_instance_desc>(scene.instances.size());
  for (size_t i = 0; i < instances.size(); i++) {
    put_frame(instances[i].frame, scene.instances[i].frame);
    instances[i].shape    = scene.instances[i].shape;
    instances[i].material = scene.instances[i].material;
  }
  shapes = std::vector<vpt_shape_desc>(scene.shapes.size());
  shape_bvhs = std::vector<vpt_bvh_desc>(scene.shapes.size());
  for (size_t i = 0; i < shapes.size(); i++) {
    auto& s = scene.shapes[i];
    auto& d = shapes[i];
    d.positions = (const float*)s.positions.data(), d.num_positions = (int)s.positions.size();
    d.normals = (const float*)s.normals.data(), d.num_normals = (int)s.normals.size();
    d.texcoords = (const float*)s.texcoords.data(), d.num_texcoords = (int)s.texcoords.size();
    d.colors = (const float*)s.colors.data(), d.num_colors = (int)s.colors.size();
    d.triangles = (const int*)s.triangles.data(), d.num_triangles = (int)s.triangles.size();
    d.quads = (const int*)s.quads.data(), d.num_quads = (int)s.quads.size();
    d.num_points = (int)s.points.size(), d.num_lines = (int)s.lines.size() / 2;
    auto& b       = bvh.shapes[i];
    shape_bvhs[i] = {b.nodes.data(), (int)b.nodes.size(), b.primitives.data(),
        (int)b.primitives.size()};
  }
  materials = std::vector<vpt_material_desc>(scene.materials.size());
  for (size_t i = 0; i < materials.size(); i++) {
    auto& m = scene.materials[i];
    auto& d = materials[i];
    d.type  = (int)m.type;
    memcpy(d.emission, &m.emission, 12), memcpy(d.color, &m.color, 12);
    d.roughness = m.roughness, d.metallic = m.metallic, d.ior = m.ior;
    memcpy(d.scattering, &m.scattering, 12);
    d.scanisotropy = m.scanisotropy, d.trdepth = m.trdepth, d.opacity = m.opacity;
    d.emission_tex = m.emission_tex, d.color_tex = m.color_tex, d.roughness_tex = m.roughness_tex;
    d.scattering_tex = m.scattering_tex, d.normal_tex = m.normal_tex;
  }
  textures = std::vector<vpt_texture_desc>(scene.textures.size());
  for (size_t i = 0; i < textures.size(); i++) {
    auto& t     = scene.textures[i];
    textures[i] = {t.width, t.height, t.linear ? 1 : 0,
        t.pixelsf.empty() ? nullptr : (const float*)t.pixelsf.data(),
        t.pixelsb.empty() ? nullptr : (const uint8_t*)t.pixelsb.data()};
  }
  environments = std::vector<vpt_environment_desc>(scene.environments.size());
  for (size_t i = 0; i < environments.size(); i++) {
    auto& e = scene.environments[i];
    put_frame(environments[i].frame, e.frame);
    memcpy(environments[i].emission, &e.emission, 12);
    environments[i].emission_tex = e.emission_tex;
  }
  volumes = std::vector<vpt_volume_desc>(scene.volumes.size());
  for (size_t i = 0; i < volumes.size(); i++) {
    auto& v    = scene.volumes[i];
    volumes[i] = {{v.whd.x, v.whd.y, v.whd.z}, v.res, v.vol.data()};
  }
  vol_instances = std::vector<vpt_volume_instance_desc>(scene.vol_instances.size());
  for (size_t i = 0; i < vol_instances.size(); i++) {
    auto& v = scene.vol_instances[i];
    put_frame(vol_instances[i].frame, v.frame);
    vol_instances[i].scalef = v.scalef, vol_instances[i].volume = v.volume;
    vol_instances[i].material = v.material;
  }
  sdfs = std::vector<vpt_sdf_desc>(scene.sdfs.size());
  for (size_t i = 0; i < sdfs.size(); i++) {
    auto& s = scene.sdfs[i];
    put_frame(sdfs[i].frame, s.frame);
    sdfs[i].type = (int)s.type, sdfs[i].material = s.material;
    memcpy(sdfs[i].params, s.params, 16), memcpy(sdfs[i].whd, &s.whd, 12);
  }
  ldesc = std::vector<vpt_light_desc>(lights.lights.size());
  for (size_t i = 0; i < ldesc.size(); i++) {
    auto& l  = lights.lights[i];
    ldesc[i] = {l.instance, l.environment, l.sdf, l.elements_cdf.data(),
        (int)l.elements_cdf.size()};
  }
  desc = vpt_scene_desc{};
  desc.num_cameras = (int)cameras.size(), desc.cameras = cameras.data();
  desc.num_instances = (int)instances.size(), desc.instances = instances.data();
  desc.num_shapes = (int)shapes.size(), desc.shapes = shapes.data();
  desc.num_materials = (int)materials.size(), desc.materials = materials.data();
  desc.num_textures = (int)textures.size(), desc.textures = textures.data();
  desc.num_environments = (int)environments.size(), desc.environments = environments.data();
  desc.num_volumes = (int)volumes.size(), desc.volumes = volumes.data();
  desc.num_vol_instances = (int)vol_instances.size(), desc.vol_instances = vol_instances.data();
  desc.num_sdfs = (int)sdfs.size(), desc.sdfs = sdfs.data();
  desc.num_lights = (int)ldesc.size(), desc.lights = ldesc.data();
  desc.scene_bvh  = {bvh.nodes.data(), (int)bvh.nodes.size(), bvh.primitives.data(),
      (int)bvh.primitives.size()};
  desc.shape_bvhs = shape_bvhs.data();
}

static vpt_scene* upload_scene_raw(
    const scene_data& scene, const bvh_scene& bvh, const pathtrace_lights& lights) {
  auto storage = scene_desc_storage{};
  make_scene_desc(storage, scene, bvh, lights);
  vpt_scene* out = nullptr;
  check(vpt_cuda_upload_scene(&storage.desc, &out), "vpt_cuda_upload_scene");
  return out;
}

// make_lights on the device: the light list and CDFs are built in HBM by
// vpt_cuda_upload_scene_device_lights (bit-identical to make_lights above)
std::shared_ptr<device_scene> upload_scene_device_lights(
    const scene_data& scene, const bvh_scene& bvh, vpt_lights_stats* stats) {
  auto storage = scene_desc_storage{};
  make_scene_desc(storage, scene, bvh, pathtrace_lights{});
  auto dev = std::make_shared<device_scene>();
  check(vpt_cuda_upload_scene_device_lights(&storage.desc, &dev->handle, stats),
      "vpt_cuda_upload_scene_device_lights");
  return dev;
}

std::shared_ptr<device_scene> upload_scene(
    const scene_data& scene, const bvh_scene& bvh, const pathtrace_lights& lights) {
  auto dev    = std::make_shared<device_scene>();
  dev->handle = upload_scene_raw(scene, bvh, lights);
  return dev;
}
vpt_scene* device_handle(const std::shared_ptr<device_scene>& scene) {
  return scene ? scene->handle : nullptr;
}
void bind_device_scene(pathtrace_state& state, const std::shared_ptr<device_scene>& dscene,
    const scene_data& scene, const bvh_scene& bvh, const pathtrace_lights& lights) {
  if (!state.device) state.device = std::make_shared<device_binding>();
  state.device->scene     = dscene;
  state.device->scene_key = &scene, state.device->bvh_key = &bvh,
  state.device->light_key = &lights;
}
vpt_state* device_handle(pathtrace_state& state) {
  return state.device ? state.device->state : nullptr;
}

void mark_device_dirty(pathtrace_state& state, int samples) {
  if (!state.device) return;
  state.device->dirty = true;
  state.samples       = samples;
}

static vpt_params to_c_params(const pathtrace_params& p) {
  auto c = vpt_params{};
  c.camera = p.camera, c.resolution = p.resolution, c.shader = (int)p.shader;
  c.samples = p.samples, c.bounces = p.bounces, c.noparallel = p.noparallel ? 1 : 0;
  c.pratio = p.pratio, c.exposure = p.exposure, c.filmic = p.filmic ? 1 : 0;
  c.noimplicit_mis = p.noimplicit_mis ? 1 : 0, c.spheretrace_maxiter = p.spheretrace_maxiter;
  c.pipeline = (int)p.pipeline;
  return c;
}

static device_binding& bind(pathtrace_state& state, const scene_data& scene,
    const bvh_scene& bvh, const pathtrace_lights& lights) {
  if (!state.device) state.device = std::make_shared<device_binding>();
  auto& dev = *state.device;
  if (!dev.scene || dev.scene_key != &scene || dev.bvh_key != &bvh || dev.light_key != &lights) {
    dev.scene     = upload_scene(scene, bvh, lights);
    dev.scene_key = &scene, dev.bvh_key = &bvh, dev.light_key = &lights;
  }
  if (!dev.state) {
    if (state.rngs.size() != (size_t)state.width * state.height)
      throw std::invalid_argument("state.rngs does not match the image size");
    check(vpt_cuda_create_state(state.width, state.height, (const uint64_t*)state.rngs.data(),
              &dev.state),
        "vpt_cuda_create_state");
    if (!dev.base_rngs)
      dev.base_rngs = std::make_shared<const std::vector<rng_state>>(state.rngs), dev.rng_offset = 0;
  }
  return dev;
}

// GPUs 1 .. gpus-1: upload a replica of the scene, create a state whose generators run
// rank << 40 draws ahead, and join it to the root state (the reduce happens in the downloads)
static void bind_replicas(device_binding& dev, pathtrace_state& state, const scene_data& scene,
    const bvh_scene& bvh, const pathtrace_lights& lights, int gpus) {
  if ((int)dev.replicas.size() == gpus - 1) return;
  int visible = vpt_cuda_device_count();
  // VPT_GPUS_OVERSUBSCRIBE=1 (tests on a one-GPU box): ranks wrap around the visible devices
  auto wrap_env = getenv("VPT_GPUS_OVERSUBSCRIBE");
  bool wrap     = wrap_env && atoi(wrap_env) != 0;
  if (gpus > visible && !wrap)
    throw std::runtime_error(std::to_string(gpus) + " GPUs requested, " + std::to_string(visible) +
                             " visible");
  check(vpt_cuda_state_remove_peers(dev.state), "vpt_cuda_state_remove_peers");
  dev.replicas.clear();
  dev.root_device = vpt_cuda_current_device();
  auto storage = scene_desc_storage{};
  make_scene_desc(storage, scene, bvh, lights);
  auto errors  = std::vector<std::string>(gpus);
  auto workers = std::vector<std::thread>{};
  for (int r = 1; r < gpus; r++) dev.replicas.push_back(std::make_unique<device_replica>());
  for (int r = 1; r < gpus; r++) {
    workers.emplace_back([&, r] {
      auto& rep  = *dev.replicas[r - 1];
      rep.device = (dev.root_device + r) % visible;
      auto ok = [&](int status) {
        if (status != VPT_OK && errors[r].empty()) errors[r] = vpt_cuda_last_error();
        return status == VPT_OK;
      };
      if (!ok(vpt_cuda_set_device(rep.device))) return;
      if (!ok(vpt_cuda_upload_scene(&storage.desc, &rep.scene))) return;
      if (!ok(vpt_cuda_create_state(state.width, state.height,
              (const uint64_t*)dev.base_rngs->data(), &rep.state)))
        return;
      ok(vpt_cuda_advance_rngs(rep.state, dev.rng_offset + rank_rng_offset(r)));
    });
  }
  for (auto& w : workers) w.join();
  for (auto& e : errors)
    if (!e.empty()) {
      dev.replicas.clear();
      throw std::runtime_error("multi-GPU set-up: " + e);
    }
  for (auto& rep : dev.replicas)
    check(vpt_cuda_state_add_peer(dev.state, rep->state), "vpt_cuda_state_add_peer");
}

void bind_device_replicas(pathtrace_state& state, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, int gpus) {
  if (gpus <= 1) return;
  bind_replicas(bind(state, scene, bvh, lights), state, scene, bvh, lights, gpus);
}

void pathtrace_samples(pathtrace_state& state, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, const pathtrace_params& params) {
  if (state.samples >= params.samples) return;
  if ((int)params.shader < 0 || (int)params.shader >= (int)pathtrace_shader_names.size())
    throw std::runtime_error("sampler unknown");
  auto& dev     = bind(state, scene, bvh, lights);
  auto  cparams = to_c_params(params);
  int   batch   = params.batch > 0 ? params.batch : 1;
  int   gpus    = params.gpus > 1 ? params.gpus : 1;
  if (gpus == 1) {
    check(vpt_cuda_trace_samples(dev.scene->handle, dev.state, &cparams, batch),
        "vpt_cuda_trace_samples");
    state.samples = vpt_cuda_state_samples(dev.state);
    dev.dirty = dev.dirty_rngs = true;
    return;
  }
  // sample-range sharding: this call's `batch` samples (capped by the total target) are split
  // over the GPUs; one host thread per GPU drives its wavefront, the caller's thread the root's
  bind_replicas(dev, state, scene, bvh, lights, gpus);
  batch = std::min(batch, params.samples - state.samples);
  // a rank's own target is open-ended (the split decides how much it runs); samples == 1 keeps
  // selecting the preview branch (yocto_pathtrace.cpp:1059-1068), which only rank 0 ever runs
  if (cparams.samples != 1) cparams.samples = 1 << 30;
  auto errors  = std::vector<std::string>(gpus);
  auto workers = std::vector<std::thread>{};
  // VPT_GPUS_TRACE=1: per-rank wall-clock interval of every call (host-thread overlap check)
  static const bool timeline = getenv("VPT_GPUS_TRACE") != nullptr;
  auto t0    = std::chrono::steady_clock::now();
  auto since = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
  auto spans = std::vector<std::array<double, 2>>(gpus, {0, 0});
  for (int r = 1; r < gpus; r++) {
    int share = sample_share(batch, gpus, r);
    if (share <= 0) continue;
    workers.emplace_back([&, r, share] {
      auto& rep   = *dev.replicas[r - 1];
      spans[r][0] = since();
      if (vpt_cuda_set_device(rep.device) != VPT_OK ||
          vpt_cuda_trace_samples(rep.scene, rep.state, &cparams, share) != VPT_OK)
        errors[r] = vpt_cuda_last_error();
      spans[r][1] = since();
    });
  }
  spans[0][0] = since();
  int status  = vpt_cuda_trace_samples(dev.scene->handle, dev.state, &cparams,
      sample_share(batch, gpus, 0));
  spans[0][1] = since();
  if (status != VPT_OK) errors[0] = vpt_cuda_last_error();
  for (auto& w : workers) w.join();
  if (timeline) {
    for (int r = 0; r < gpus; r++)
      fprintf(stderr, "[vpt gpus] rank %d: %.1f .. %.1f ms (%d samples)\n", r, spans[r][0], spans[r][1],
          sample_share(batch, gpus, r));
  }
  for (auto& e : errors)
    if (!e.empty()) throw std::runtime_error("vpt_cuda_trace_samples: " + e);
  state.samples = vpt_cuda_state_total_samples(dev.state);
  dev.dirty = dev.dirty_rngs = true;
}

// the accumulation image only (what get_render needs): half the bytes of sync_state
void sync_image(pathtrace_state& state) {
  if (!state.device || !state.device->state || !state.device->dirty) return;
  auto& dev = *state.device;
  check(vpt_cuda_download_image(dev.state, (float*)state.image.data(), state.image.size() * 4),
      "vpt_cuda_download_image");
  state.hits.assign(state.hits.size(), state.samples);
  dev.dirty = false;
}

void sync_state(pathtrace_state& state) {
  sync_image(state);
  if (!state.device || !state.device->state || !state.device->dirty_rngs) return;
  auto& dev = *state.device;
  check(vpt_cuda_download_rngs(dev.state, (uint64_t*)state.rngs.data(), state.rngs.size() * 2),
      "vpt_cuda_download_rngs");
  dev.dirty_rngs = false;
}

void reset_device_state(pathtrace_state& state, const std::vector<rng_state>& rngs,
    uint64_t rng_offset) {
  reset_device_state(state, std::make_shared<const std::vector<rng_state>>(rngs), rng_offset);
}

// the caller shares ownership of the generators: no 16-bytes-per-pixel copy per restart
void reset_device_state(pathtrace_state& state,
    const std::shared_ptr<const std::vector<rng_state>>& shared_rngs, uint64_t rng_offset) {
  if (!shared_rngs) throw std::invalid_argument("reset_device_state: null generators");
  if (!state.device) state.device = std::make_shared<device_binding>();
  auto& dev     = *state.device;
  auto& rngs    = *shared_rngs;
  dev.base_rngs = shared_rngs;
  if (!dev.state) {
    // no device state yet: restart the host copies (same O(log n) PCG32 jump as the device's)
    state.rngs = rngs;
    state.image.assign(state.image.size(), {0, 0, 0, 0});
    for (auto& g : state.rngs) {
      uint64_t cur_mult = 6364136223846793005ULL, cur_plus = g.inc, acc_mult = 1, acc_plus = 0;
      for (auto delta = rng_offset; delta > 0; delta >>= 1) {
        if (delta & 1) acc_mult *= cur_mult, acc_plus = acc_plus * cur_mult + cur_plus;
        cur_plus = (cur_mult + 1) * cur_plus, cur_mult *= cur_mult;
      }
      g.state = acc_mult * g.state + acc_plus;
    }
    dev.rng_offset = rng_offset;
    state.samples  = 0;
    return;
  }
  check(vpt_cuda_reset_state(dev.state, (const uint64_t*)rngs.data()), "vpt_cuda_reset_state");
  if (rng_offset) check(vpt_cuda_advance_rngs(dev.state, rng_offset), "vpt_cuda_advance_rngs");
  dev.rng_offset = rng_offset;
  int current    = vpt_cuda_current_device();
  for (size_t r = 0; r < dev.replicas.size(); r++) {
    auto& rep = *dev.replicas[r];
    check(vpt_cuda_set_device(rep.device), "vpt_cuda_set_device");
    auto a = vpt_cuda_reset_state(rep.state, (const uint64_t*)rngs.data());
    auto b = a == VPT_OK ? vpt_cuda_advance_rngs(rep.state, rng_offset + rank_rng_offset((int)r + 1)) : a;
    vpt_cuda_set_device(current);
    check(b, "resetting a GPU replica");
  }
  state.samples = 0;
  dev.dirty = dev.dirty_rngs = true;
}

void mark_host_stale(pathtrace_state& state) {
  if (state.device) state.device->dirty = state.device->dirty_rngs = true;
}

std::vector<uint8_t> get_render_srgb8(pathtrace_state& state) {
  return get_render_tonemapped8(state, 0, false);
}

std::vector<uint8_t> get_render_tonemapped8(pathtrace_state& state, float exposure, bool filmic) {
  auto dstate = device_handle(state);
  if (!dstate) throw std::invalid_argument("nothing rendered yet");
  auto bytes = std::vector<uint8_t>((size_t)state.width * state.height * 4);
  if (vpt_cuda_download_tonemapped8(dstate, state.samples, exposure, filmic ? 1 : 0, bytes.data(),
          bytes.size()) != VPT_OK)
    throw std::runtime_error(vpt_cuda_last_error());
  return bytes;
}

color_image get_render(const pathtrace_state& state) {
  auto image   = color_image{};
  image.width  = state.width, image.height = state.height, image.linear = true;
  image.pixels.assign((size_t)state.width * state.height, {0, 0, 0, 0});
  get_render(image, state);
  return image;
}

void get_render(color_image& image, const pathtrace_state& state) {
  if (image.width != state.width || image.height != state.height)
    throw std::invalid_argument{"image should have the same size"};
  if (!image.linear) throw std::invalid_argument{"expected linear image"};
  // refresh the host mirror; the reference API takes the state by const ref
  sync_image(const_cast<pathtrace_state&>(state));
  auto scale = 1.0f / (float)state.samples;
  for (size_t idx = 0; idx < (size_t)state.width * state.height; idx++)
    image.pixels[idx] = state.image[idx] * scale;
}

vpt_trace_stats get_trace_stats(const pathtrace_state& state) {
  auto stats = vpt_trace_stats{};
  if (state.device && state.device->state)
    check(vpt_cuda_get_trace_stats(state.device->state, &stats), "vpt_cuda_get_trace_stats");
  return stats;
}

std::vector<bvh_intersection> intersect_bvh(const bvh_scene& bvh, const scene_data& scene,
    const pathtrace_lights& lights, const std::vector<float>& rays, int instance) {
  auto n      = (int)(rays.size() / 8);
  auto dscene = upload_scene(scene, bvh, lights);
  auto ids    = std::vector<int32_t>((size_t)n * 3);
  auto uvd    = std::vector<float>((size_t)n * 3);
  check(vpt_cuda_intersect(dscene->handle, rays.data(), n, instance, ids.data(), uvd.data()),
      "vpt_cuda_intersect");
  auto out = std::vector<bvh_intersection>(n);
  for (auto i = 0; i < n; i++) {
    out[i].hit      = ids[3 * i] != 0;
    out[i].instance = ids[3 * i + 1], out[i].element = ids[3 * i + 2];
    out[i].uv       = {uvd[3 * i], uvd[3 * i + 1]};
    out[i].distance = uvd[3 * i + 2];
  }
  return out;
}

}  // namespace vpt
