// vpt_pathtrace.h — the yocto_pathtrace API, served by the B200.
//
// Same names, argument meaning, defaults and error behaviour as the
// reference's public header (libs/yocto_pathtrace/yocto_pathtrace.h:57-139):
// pathtrace_params, pathtrace_state, pathtrace_lights, the shader enum and its
// names, make_state / make_bvh / make_lights / tesselate_surfaces /
// pathtrace_samples / get_render. trace_samples is an alias of
// pathtrace_samples (the name BASELINE.json uses). The host builds (BVH, light
// CDFs, per-pixel generators) run on the CPU as in the reference; everything
// executed per path runs on the GPU through include/vpt_cuda.h. There is no
// CPU rendering path: without the CUDA library / a device the calls throw.
#pragma once

#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "vpt_scene.h"

namespace vpt {

struct rng_state {
  uint64_t state = 0x853c49e6748fea9bULL;
  uint64_t inc   = 0xda3e39cb94b95bdbULL;
};

struct device_binding;  // device-resident mirrors owned by the library

// Rendering state (yocto_pathtrace.h:57-64). image / hits mirror the device
// accumulation buffers and are refreshed by sync_state() / get_render().
struct pathtrace_state {
  int                    width   = 0;
  int                    height  = 0;
  int                    samples = 0;
  std::vector<vec4f>     image   = {};
  std::vector<int>       hits    = {};
  std::vector<rng_state> rngs    = {};
  std::shared_ptr<device_binding> device = {};
};

// Type of tracing algorithm (order is ABI, yocto_pathtrace.h:74-84)
enum struct pathtrace_shader_type {
  volpathtrace, pathtrace, naive, eyelight, normal, texcoord, color, implicit, implicit_normal,
};

enum struct pathtrace_pipeline { automatic = 0, megakernel = 1, wavefront = 2 };

// Options for trace functions (yocto_pathtrace.h:87-99)
struct pathtrace_params {
  int                   camera              = 0;
  int                   resolution          = 720;
  pathtrace_shader_type shader              = pathtrace_shader_type::pathtrace;
  int                   samples             = 512;
  int                   bounces             = 4;
  bool                  noparallel          = false;  // accepted, ignored on the GPU
  int                   pratio              = 8;
  float                 exposure            = 0;
  bool                  filmic              = false;
  bool                  noimplicit_mis      = false;
  int                   spheretrace_maxiter = 450;
  // device options (not in the reference)
  pathtrace_pipeline    pipeline            = pathtrace_pipeline::automatic;
  int                   batch               = 1;  // samples per pathtrace_samples call
  // GPUs of this box to render on: every GPU holds the scene and a full
  // accumulation image and takes a disjoint range of each call's samples; GPU r > 0 runs every
  // pixel's PCG32 stream from r << 40 draws ahead; get_render sums the images on the first GPU
  // (P2P over NVLink, vpt_cuda_state_add_peer). 1 GPU = the reference's streams exactly.
  int                   gpus                = 1;
  // make_bvh builds the shape BVHs on the device (LBVH) instead of the
  // reference's middle-split trees: same closest hits, exact-distance ties may pick another
  // primitive, so the bit-exact image gates need the default (false)
  bool                  device_bvh          = false;
};

// Samples GPU `rank` of `gpus` takes out of a call of `batch` samples: contiguous ranges, the
// first batch % gpus ranks one more (the C++ twin of sharding.sample_ranges).
inline int sample_share(int batch, int gpus, int rank) {
  return batch / gpus + (rank < batch % gpus ? 1 : 0);
}
// draws every pixel's generator of GPU `rank` is advanced by (PCG32 jump-ahead)
inline uint64_t rank_rng_offset(int rank) { return (uint64_t)rank << 40; }

inline const auto pathtrace_shader_names = std::vector<std::string>{"volpathtrace", "pathtrace",
    "naive", "eyelight", "normal", "texcoord", "color", "implicit", "implicit_normal"};

// Scene lights (yocto_pathtrace.h:106-116)
struct pathtrace_light {
  int                instance     = invalidid;
  int                environment  = invalidid;
  int                sdf          = invalidid;
  std::vector<float> elements_cdf = {};
};
struct pathtrace_lights {
  std::vector<pathtrace_light> lights = {};
};

// Linear float image returned by get_render (image_data, yocto_image.h:68)
struct color_image {
  int                width  = 0;
  int                height = 0;
  bool               linear = true;
  std::vector<vec4f> pixels = {};
};

// Initialize state: image size from the camera aspect and one PCG32 stream per
// pixel, seeded sequentially from one generator (yocto_pathtrace.cpp:960-980).
pathtrace_state make_state(const scene_data& scene, const pathtrace_params& params);

// Build the two-level BVH on the host (middle split, <= 4 primitives per leaf;
// yocto_pathtrace.cpp:955, yocto_bvh.cpp:411-611).
bvh_scene make_bvh(const scene_data& scene, const pathtrace_params& params);
bvh_shape make_bvh(const shape_data& shape);
bvh_shape make_bvh_device(const shape_data& shape, vpt_bvh_build_stats* stats = nullptr);

// Light list and sampling CDFs (yocto_pathtrace.cpp:983-1049).
pathtrace_lights make_lights(const scene_data& scene, const pathtrace_params& params);

// Tesselate subdivs (yocto_pathtrace.h:128, yocto_pathtrace.cpp:1230-1280): Catmull-Clark levels,
// smooth normals and texture displacement run on the device (vpt_cuda_catmullclark /
// _vertex_normals / _displace), the face-varying split and triangulation on the host; shapes
// equal the reference's bit for bit. Archives written by tools/vpt_export without
// --keep-subdivs arrive tessellated (by the reference's own code) and have no subdivs left.
void tesselate_surfaces(scene_data& scene);
void tesselate_surface(shape_data& shape, const subdiv_data& subdiv, const scene_data& scene);

// Progressively computes an image: params.batch samples per call (default 1,
// as the reference). Blocking. Throws std::runtime_error("sampler unknown") for
// a bad shader and on any CUDA failure.
void pathtrace_samples(pathtrace_state& state, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, const pathtrace_params& params);
inline void trace_samples(pathtrace_state& state, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, const pathtrace_params& params) {
  pathtrace_samples(state, scene, bvh, lights, params);
}

// Copies the device accumulation image and generators back into `state`.
void sync_state(pathtrace_state& state);
// the same for state.image / state.hits only (get_render uses it)
void sync_image(pathtrace_state& state);

// Get resulting render (divides by state.samples, yocto_pathtrace.cpp:1105-1116);
// the second form throws std::invalid_argument on a size / linearity mismatch.
color_image get_render(const pathtrace_state& state);
// Output stage on the device: get_render + rgb_to_srgb + float_to_byte (what save_image encodes
// for 8-bit formats, yocto_sceneio.cpp:523-529); width*height*4 bytes, RGBA.
std::vector<uint8_t> get_render_srgb8(pathtrace_state& state);
// the same behind the display tonemap (tonemap(), yocto_color.h:306-316: exposure, filmic) that
// the reference's viewer applies per frame (ypathtrace.cpp:159-189); bytes equal the reference's
std::vector<uint8_t> get_render_tonemapped8(pathtrace_state& state, float exposure, bool filmic);
void        get_render(color_image& render, const pathtrace_state& state);

// The C-ABI scene description (include/vpt_cuda.h) of a host scene; the
// pointers inside `desc` refer to `scene`, `bvh`, `lights` and to this storage.
struct scene_desc_storage {
  std::vector<vpt_camera_desc>          cameras;
  std::vector<vpt_instance_desc>        instances;
  std::vector<vpt_shape_desc>           shapes;
  std::vector<vpt_bvh_desc>             shape_bvhs;
  std::vector<vpt_material_desc>        materials;
  std::vector<vpt_texture_desc>         textures;
  std::vector<vpt_environment_desc>     environments;
  std::vector<vpt_volume_desc>          volumes;
  std::vector<vpt_volume_instance_desc> vol_instances;
  std::vector<vpt_sdf_desc>             sdfs;
  std::vector<vpt_light_desc>           lights;
  vpt_scene_desc                        desc{};
};
void make_scene_desc(scene_desc_storage& storage, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights);

// Explicit device-scene management, for callers that render the same scene
// from several states (the first pathtrace_samples call otherwise uploads).
struct device_scene;
// the same with make_lights run on the device; `lights` passed to later
// pathtrace_samples calls is then only the cache key
std::shared_ptr<device_scene> upload_scene_device_lights(
    const scene_data& scene, const bvh_scene& bvh, vpt_lights_stats* stats = nullptr);
std::shared_ptr<device_scene> upload_scene(
    const scene_data& scene, const bvh_scene& bvh, const pathtrace_lights& lights);
void bind_device_scene(pathtrace_state& state, const std::shared_ptr<device_scene>& dscene,
    const scene_data& scene, const bvh_scene& bvh, const pathtrace_lights& lights);
// Contexts, scene replicas and states of GPUs 1 .. gpus-1 ahead of time (the first
// pathtrace_samples call with params.gpus > 1 otherwise sets them up: about a second per process
// for the extra contexts and the module load)
void bind_device_replicas(pathtrace_state& state, const scene_data& scene, const bvh_scene& bvh,
    const pathtrace_lights& lights, int gpus);
vpt_scene* device_handle(const std::shared_ptr<device_scene>& scene);
vpt_state* device_handle(pathtrace_state& state);
// the device image was replaced (e.g. by a cross-GPU reduce): state.samples becomes
// `samples` and the host mirror is refreshed on the next sync_state / get_render
void mark_device_dirty(pathtrace_state& state, int samples);
// the device state was re-initialised behind the host mirrors' back (reset): image and
// generators are downloaded again on the next sync
void mark_host_stale(pathtrace_state& state);

// Restart a render: the device image(s) are cleared and every generator is set to `rngs`
// advanced by `rng_offset` draws (+ rank << 40 on the GPUs of a multi-GPU render); the host
// mirrors are refreshed on the next sync. Before the first sample the host copies are restarted.
void reset_device_state(pathtrace_state& state, const std::vector<rng_state>& rngs,
    uint64_t rng_offset);
void reset_device_state(pathtrace_state& state,
    const std::shared_ptr<const std::vector<rng_state>>& rngs, uint64_t rng_offset = 0);

// Device statistics of the last pathtrace_samples call on this state.
vpt_trace_stats get_trace_stats(const pathtrace_state& state);

// Batched first-hit queries through the device traversal (intersect_bvh,
// yocto_bvh.cpp:1097-1113). rays: n x {o.xyz, d.xyz, tmin, tmax}.
struct bvh_intersection {
  int   instance = -1;
  int   element  = -1;
  vec2f uv       = {0, 0};
  float distance = 0;
  bool  hit      = false;
};
std::vector<bvh_intersection> intersect_bvh(const bvh_scene& bvh, const scene_data& scene,
    const pathtrace_lights& lights, const std::vector<float>& rays, int instance = -1);

}  // namespace vpt
