// vpt_host_c.cpp — extern "C" wrappers of the C++ facade (include/vpt_host.h),
// used by the Python tests / bench through ctypes. Exceptions never cross the
// C boundary: they become a negative status + vpth_last_error().
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../include/vpt_host.h"
#include "vpt_pathtrace.h"

using namespace vpt;

struct vpth_scene {
  scene_data                    scene;
  bvh_scene                     bvh;
  pathtrace_lights              lights;
  std::shared_ptr<device_scene> device;
};

struct vpth_render {
  vpth_scene*            owner = nullptr;
  pathtrace_params       params;
  pathtrace_state        state;
  std::shared_ptr<const std::vector<rng_state>> initial_rngs;
  std::vector<void*>     pinned;
  int                    transfers = 0;
  // Page-locking the three host mirrors costs 60 - 230 ms (24 MB each at 1920x800; the calls of
  // several processes serialise in the kernel: 8 ranks of one box paid 150 - 230 ms each) and
  // saves about a millisecond per transfer, so it happens on request (vpth_render_pin) or once a
  // render has moved its mirrors three times (a preview / end-to-end loop), not for a render
  // that downloads one image at the end.
  void note_transfer() {
    if (++transfers >= 3) pin();
  }
  // page-lock the host mirrors once a device exists (best effort)
  void pin() {
    if (!pinned.empty() || vpt_cuda_device_count() < 1) return;
    auto reg = [&](void* p, size_t n) {
      if (n && vpt_cuda_host_register(p, n) == VPT_OK) pinned.push_back(p);
    };
    if (initial_rngs) reg((void*)initial_rngs->data(), initial_rngs->size() * sizeof(rng_state));
    reg(state.rngs.data(), state.rngs.size() * sizeof(rng_state));
    reg(state.image.data(), state.image.size() * sizeof(vec4f));
  }
  ~vpth_render() {
    for (auto p : pinned) vpt_cuda_host_unregister(p);
  }
};

static thread_local std::string g_error;

template <typename F>
static int guarded(F&& f) {
  try {
    f();
    return VPT_OK;
  } catch (std::invalid_argument& e) {
    g_error = e.what();
    return VPT_ERR_INVALID;
  } catch (std::exception& e) {
    g_error = e.what();
    auto msg = std::string(e.what());
    if (msg.find("no CUDA device") != std::string::npos) return VPT_ERR_NO_DEVICE;
    if (msg.find("unsupported") != std::string::npos ||
        msg.find("not supported") != std::string::npos)
      return VPT_ERR_UNSUPPORTED;
    return VPT_ERR_CUDA;
  }
}

static vpt_scene* device_scene_of(vpth_scene* s) {
  if (!s->device) s->device = upload_scene(s->scene, s->bvh, s->lights);
  return device_handle(s->device);
}

static pathtrace_params from_c_params(const vpt_params& c) {
  auto p = pathtrace_params{};
  p.camera = c.camera, p.resolution = c.resolution, p.shader = (pathtrace_shader_type)c.shader;
  p.samples = c.samples, p.bounces = c.bounces, p.noparallel = c.noparallel != 0;
  p.pratio = c.pratio, p.exposure = c.exposure, p.filmic = c.filmic != 0;
  p.noimplicit_mis = c.noimplicit_mis != 0, p.spheretrace_maxiter = c.spheretrace_maxiter;
  p.pipeline = (pathtrace_pipeline)c.pipeline;
  return p;
}

extern "C" {

const char* vpth_last_error(void) { return g_error.c_str(); }

int vpth_load_scene(const char* path, vpth_scene** out) {
  if (!path || !out) {
    g_error = "null argument";
    return VPT_ERR_INVALID;
  }
  *out   = nullptr;
  auto s = new vpth_scene();
  auto r = guarded([&] {
    auto error = std::string{};
    if (!load_scene(path, s->scene, error)) throw std::invalid_argument(error);
    auto params = pathtrace_params{};
    s->bvh      = make_bvh(s->scene, params);
    s->lights   = make_lights(s->scene, params);
  });
  if (r != VPT_OK) {
    delete s;
    return r;
  }
  *out = s;
  return VPT_OK;
}

void vpth_free_scene(vpth_scene* scene) { delete scene; }

int vpth_scene_counts(const vpth_scene* s, int32_t counts[12]) {
  if (!s || !counts) return VPT_ERR_INVALID;
  size_t shape_nodes = 0;
  for (auto& b : s->bvh.shapes) shape_nodes += b.nodes.size();
  size_t v[12] = {s->scene.cameras.size(), s->scene.instances.size(), s->scene.shapes.size(),
      s->scene.materials.size(), s->scene.textures.size(), s->scene.environments.size(),
      s->scene.volumes.size(), s->scene.vol_instances.size(), s->scene.sdfs.size(),
      s->lights.lights.size(), s->bvh.nodes.size(), shape_nodes};
  for (int i = 0; i < 12; i++) counts[i] = (int32_t)v[i];
  return VPT_OK;
}

static const bvh_data* pick_bvh(const vpth_scene* s, int shape) {
  if (shape < 0) return &s->bvh;
  if (shape >= (int)s->bvh.shapes.size()) return nullptr;
  return &s->bvh.shapes[shape];
}

int vpth_bvh_size(const vpth_scene* s, int shape, int32_t* num_nodes, int32_t* num_prims) {
  auto b = s ? pick_bvh(s, shape) : nullptr;
  if (!b || !num_nodes || !num_prims) return VPT_ERR_INVALID;
  *num_nodes = (int32_t)b->nodes.size(), *num_prims = (int32_t)b->primitives.size();
  return VPT_OK;
}

int vpth_bvh_copy(const vpth_scene* s, int shape, vpt_bvh_node* nodes, int32_t* prims) {
  auto b = s ? pick_bvh(s, shape) : nullptr;
  if (!b || !nodes || !prims) return VPT_ERR_INVALID;
  memcpy(nodes, b->nodes.data(), b->nodes.size() * sizeof(vpt_bvh_node));
  memcpy(prims, b->primitives.data(), b->primitives.size() * sizeof(int32_t));
  return VPT_OK;
}

int vpth_light_info(const vpth_scene* s, int light, int32_t ids[3], int32_t* cdf_size) {
  if (!s || light < 0 || light >= (int)s->lights.lights.size() || !ids || !cdf_size)
    return VPT_ERR_INVALID;
  auto& l = s->lights.lights[light];
  ids[0] = l.instance, ids[1] = l.environment, ids[2] = l.sdf;
  *cdf_size = (int32_t)l.elements_cdf.size();
  return VPT_OK;
}

int vpth_light_cdf(const vpth_scene* s, int light, float* cdf) {
  if (!s || light < 0 || light >= (int)s->lights.lights.size() || !cdf) return VPT_ERR_INVALID;
  auto& l = s->lights.lights[light];
  memcpy(cdf, l.elements_cdf.data(), l.elements_cdf.size() * sizeof(float));
  return VPT_OK;
}

int vpth_scene_load_subdivs(vpth_scene* s, const char* path) {
  if (!s || !path) return VPT_ERR_INVALID;
  return guarded([&] {
    auto error = std::string{};
    if (!load_subdivs(path, s->scene, error)) throw std::runtime_error(error);
  });
}

// tesselate_surfaces on the device; the BVHs, lights and the device scene are rebuilt
int vpth_scene_tesselate(vpth_scene* s) {
  if (!s) return VPT_ERR_INVALID;
  return guarded([&] {
    for (auto& sd : s->scene.subdivs)  // nothing of an earlier tessellation may survive
      if (sd.shape >= 0 && sd.shape < (int)s->scene.shapes.size()) s->scene.shapes[sd.shape] = shape_data{};
    tesselate_surfaces(s->scene);
    auto params = pathtrace_params{};
    s->bvh      = make_bvh(s->scene, params);
    s->lights   = make_lights(s->scene, params);
    s->device   = nullptr;
  });
}

int vpth_shape_sizes(const vpth_scene* s, int shape, int32_t sizes[6]) {
  if (!s || !sizes || shape < 0 || shape >= (int)s->scene.shapes.size()) return VPT_ERR_INVALID;
  auto& sh = s->scene.shapes[shape];
  sizes[0] = (int)sh.positions.size(), sizes[1] = (int)sh.normals.size(), sizes[2] = (int)sh.texcoords.size();
  sizes[3] = (int)sh.triangles.size(), sizes[4] = (int)sh.quads.size(), sizes[5] = (int)sh.colors.size();
  return VPT_OK;
}

int vpth_shape_copy(const vpth_scene* s, int shape, float* positions, float* normals, float* texcoords,
    int32_t* triangles, int32_t* quads) {
  if (!s || shape < 0 || shape >= (int)s->scene.shapes.size()) return VPT_ERR_INVALID;
  auto& sh = s->scene.shapes[shape];
  if (positions) memcpy(positions, sh.positions.data(), sizeof(vec3f) * sh.positions.size());
  if (normals) memcpy(normals, sh.normals.data(), sizeof(vec3f) * sh.normals.size());
  if (texcoords) memcpy(texcoords, sh.texcoords.data(), sizeof(vec2f) * sh.texcoords.size());
  if (triangles) memcpy(triangles, sh.triangles.data(), sizeof(vec3i) * sh.triangles.size());
  if (quads) memcpy(quads, sh.quads.data(), sizeof(vec4i) * sh.quads.size());
  return VPT_OK;
}

int vpth_scene_use_device_bvh(vpth_scene* s, vpt_bvh_build_stats* total) {
  if (!s) return VPT_ERR_INVALID;
  return guarded([&] {
    auto params       = pathtrace_params{};
    params.device_bvh = true;
    auto sum          = vpt_bvh_build_stats{};
    // shape by shape for the statistics, then the (host) scene level over the new root boxes
    for (auto& shape : s->scene.shapes) {
      auto st = vpt_bvh_build_stats{};
      make_bvh_device(shape, &st);
      sum.primitives += st.primitives, sum.nodes += st.nodes, sum.keys_ms += st.keys_ms;
      sum.sort_ms += st.sort_ms, sum.hierarchy_ms += st.hierarchy_ms;
    }
    s->bvh    = make_bvh(s->scene, params);
    s->device = nullptr;  // uploaded again on next use
    if (total) *total = sum;
  });
}

int vpth_scene_use_device_lights(vpth_scene* s, vpt_lights_stats* stats) {
  if (!s) return VPT_ERR_INVALID;
  return guarded([&] { s->device = upload_scene_device_lights(s->scene, s->bvh, stats); });
}

int vpth_device_light(vpth_scene* s, int light, int32_t ids[3], float* cdf, size_t capacity,
    int32_t* num_elements) {
  if (!s) return VPT_ERR_INVALID;
  vpt_scene* dev = nullptr;
  auto r = guarded([&] { dev = device_scene_of(s); });
  if (r != VPT_OK) return r;
  if (light < 0) {  // query: number of lights on the device
    if (!num_elements) return VPT_ERR_INVALID;
    *num_elements = vpt_cuda_scene_num_lights(dev);
    return VPT_OK;
  }
  r = vpt_cuda_download_light(dev, light, ids, cdf, capacity, num_elements);
  if (r != VPT_OK) g_error = vpt_cuda_last_error();
  return r;
}

int vpth_render_create(vpth_scene* scene, const vpt_params* params, vpth_render** out) {
  if (!scene || !params || !out) {
    g_error = "null argument";
    return VPT_ERR_INVALID;
  }
  *out   = nullptr;
  auto r = new vpth_render();
  auto e = guarded([&] {
    if (params->camera < 0 || params->camera >= (int)scene->scene.cameras.size())
      throw std::invalid_argument("camera index out of range");
    r->owner        = scene;
    r->params       = from_c_params(*params);
    r->state        = make_state(scene->scene, r->params);
    r->initial_rngs = std::make_shared<const std::vector<rng_state>>(r->state.rngs);
  });
  if (e != VPT_OK) {
    delete r;
    return e;
  }
  *out = r;
  return VPT_OK;
}

void vpth_render_free(vpth_render* render) { delete render; }

int vpth_render_size(const vpth_render* r, int32_t* width, int32_t* height) {
  if (!r || !width || !height) return VPT_ERR_INVALID;
  *width = r->state.width, *height = r->state.height;
  return VPT_OK;
}

int vpth_render_get_rngs(vpth_render* r, uint64_t* rngs) {
  if (!r || !rngs) return VPT_ERR_INVALID;
  return guarded([&] {
    sync_state(r->state);
    memcpy(rngs, r->state.rngs.data(), r->state.rngs.size() * sizeof(rng_state));
  });
}

int vpth_render_reset(vpth_render* r, uint64_t rng_offset) {
  if (!r) return VPT_ERR_INVALID;
  return guarded([&] {
    r->note_transfer();
    reset_device_state(r->state, r->initial_rngs, rng_offset);
  });
}

int vpth_render_pin(vpth_render* r) {
  if (!r) return VPT_ERR_INVALID;
  return guarded([&] { r->pin(); });
}

int vpth_render_samples(vpth_render* r, int nsamples) {
  if (!r || nsamples <= 0) return VPT_ERR_INVALID;
  return guarded([&] {
    auto s = r->owner;
    // VPT_HOST_TRACE=1: wall-clock split of the call (scene upload / binding / trace)
    static const bool trace = getenv("VPT_HOST_TRACE") != nullptr;
    auto t0 = std::chrono::steady_clock::now();
    auto lap = [&] {
      auto t1 = std::chrono::steady_clock::now();
      auto ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
      t0 = t1;
      return ms;
    };
    double ms_upload = 0, ms_bind = 0, ms_pin = 0;
    if (!s->device) s->device = upload_scene(s->scene, s->bvh, s->lights);
    ms_upload = lap();
    bind_device_scene(r->state, s->device, s->scene, s->bvh, s->lights);
    ms_bind = lap();
    ms_pin = lap();
    auto params  = r->params;
    params.batch = nsamples;
    pathtrace_samples(r->state, s->scene, s->bvh, s->lights, params);
    if (trace)
      fprintf(stderr, "[vpt host] upload %.1f ms, bind %.1f ms, trace %.1f ms (%d samples)\n", ms_upload,
          ms_bind + ms_pin, lap(), nsamples);
  });
}

int vpth_render_set_gpus(vpth_render* r, int gpus) {
  if (!r || gpus < 1) return VPT_ERR_INVALID;
  r->params.gpus = gpus;
  return VPT_OK;
}

int vpth_render_done_samples(const vpth_render* r) { return r ? r->state.samples : 0; }

int vpth_render_get_image(vpth_render* r, float* rgba, int normalized) {
  if (!r || !rgba) return VPT_ERR_INVALID;
  return guarded([&] {
    r->note_transfer();
    auto n = (size_t)r->state.width * r->state.height;
    if (normalized) {
      auto img = get_render(r->state);
      memcpy(rgba, img.pixels.data(), n * sizeof(vec4f));
    } else {
      sync_image(r->state);
      memcpy(rgba, r->state.image.data(), n * sizeof(vec4f));
    }
  });
}

int vpth_render_image_view(vpth_render* r, const float** rgba) {
  if (!r || !rgba) return VPT_ERR_INVALID;
  return guarded([&] {
    r->note_transfer();
    sync_image(r->state);
    *rgba = (const float*)r->state.image.data();
  });
}

int vpth_render_get_srgb8(vpth_render* r, uint8_t* rgba) {
  if (!r || !rgba) return VPT_ERR_INVALID;
  return guarded([&] {
    auto bytes = get_render_srgb8(r->state);
    memcpy(rgba, bytes.data(), bytes.size());
  });
}

int vpth_render_get_tonemapped8(vpth_render* r, float exposure, int filmic, uint8_t* rgba) {
  if (!r || !rgba) return VPT_ERR_INVALID;
  return guarded([&] {
    auto bytes = get_render_tonemapped8(r->state, exposure, filmic != 0);
    memcpy(rgba, bytes.data(), bytes.size());
  });
}

int vpth_render_stats(const vpth_render* r, vpt_trace_stats* stats) {
  if (!r || !stats) return VPT_ERR_INVALID;
  return guarded([&] { *stats = get_trace_stats(r->state); });
}

void* vpth_render_device_image(vpth_render* r) {
  if (!r) return nullptr;
  auto dstate = device_handle(r->state);
  return dstate ? vpt_cuda_image_device_ptr(dstate) : nullptr;
}

int vpth_render_export_image(vpth_render* r, void* device_dst) {
  if (!r) return VPT_ERR_INVALID;
  return guarded([&] {
    auto dstate = device_handle(r->state);
    if (!dstate) throw std::invalid_argument("nothing rendered yet");
    auto n = (size_t)r->state.width * r->state.height * 4;
    if (vpt_cuda_export_image(dstate, device_dst, n) != VPT_OK)
      throw std::runtime_error(vpt_cuda_last_error());
  });
}

// after a cross-GPU reduce: the summed image and the total sample count come back
int vpth_render_import_image(vpth_render* r, const void* device_src, int total_samples) {
  if (!r) return VPT_ERR_INVALID;
  return guarded([&] {
    auto dstate = device_handle(r->state);
    if (!dstate) throw std::invalid_argument("nothing rendered yet");
    auto n = (size_t)r->state.width * r->state.height * 4;
    if (vpt_cuda_import_image(dstate, device_src, n) != VPT_OK)
      throw std::runtime_error(vpt_cuda_last_error());
    mark_device_dirty(r->state, total_samples);
  });
}

vpt_scene* vpth_scene_device_handle(vpth_scene* scene) {
  if (!scene) return nullptr;
  try {
    return device_scene_of(scene);
  } catch (const std::exception& e) {
    g_error = e.what();
    return nullptr;
  }
}

int vpth_intersect(vpth_scene* s, const float* rays, int n, int instance, int32_t* ids,
    float* uvd) {
  if (!s) return VPT_ERR_INVALID;
  int status = VPT_OK;
  auto r = guarded([&] {
    status = vpt_cuda_intersect(device_scene_of(s), rays, n, instance, ids, uvd);
    if (status != VPT_OK) g_error = vpt_cuda_last_error();
  });
  return r != VPT_OK ? r : status;
}

int vpth_spheretrace(vpth_scene* s, const float* rays, int n, int maxiter, int32_t* ids,
    float* dn) {
  if (!s) return VPT_ERR_INVALID;
  int status = VPT_OK;
  auto r = guarded([&] {
    status = vpt_cuda_spheretrace(device_scene_of(s), rays, n, maxiter, ids, dn);
    if (status != VPT_OK) g_error = vpt_cuda_last_error();
  });
  return r != VPT_OK ? r : status;
}

}  // extern "C"
