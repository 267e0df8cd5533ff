// ref_render — TEST INFRASTRUCTURE / TIMED CPU BASELINE (oracle side).
// Drives the UNMODIFIED reference CPU path (libyocto_ref.a, built by
// oracle/Makefile from /root/reference) on a ".vptscene" archive:
//   make_bvh / make_lights / make_state / pathtrace_samples / get_render
//   (yocto_pathtrace.cpp:955, 983, 960, 1052, 1105) — the same sequence as
//   run_offline (apps/ypathtrace/ypathtrace.cpp:41-87), headless.
// Also dumps first hits for a given ray batch through the reference's
// intersect_bvh (yocto_bvh.cpp:1097, :1105) and spheretrace
// (yocto_pathtrace.cpp:267, :289) so the CUDA traversal can be checked
// bit-for-bit. Output is a named-array archive read by tests/ and bench.py.
//
// usage: ref_render <scene.vptscene> <out.vptarch> [--shader NAME]
//        [--samples N] [--resolution R] [--bounces B] [--noimplicitmis]
//        [--stmaxiter M] [--noparallel] [--rays rays.vptarch]
//        [--instance K] [--dump-state] [--dump-host] [--camera N]
#include <yocto/yocto_bvh.h>
#include <yocto/yocto_image.h>
#include <yocto/yocto_scene.h>
#include <yocto/yocto_sdfs.h>
#include <yocto_pathtrace/yocto_pathtrace.h>

#include <chrono>
#include <thread>

#include "ref_scene_from_archive.h"

using namespace yocto;
using std::string;
using std::vector;

// Declared (not static) in yocto_pathtrace.cpp:259-307; re-declared here with
// the same layout so the dump can call the reference's own sphere tracer.
namespace yocto {
struct spheretrace_result {
  bool  hit      = false;
  float dist     = flt_max;
  int   instance = invalidid;
  int   sdf      = invalidid;
};
spheretrace_result spheretrace(
    const scene_data& scene, const ray3f& ray, int sdf_handle, int maxiter);
spheretrace_result spheretrace(
    const scene_data& scene, const ray3f& ray, int maxiter);
}  // namespace yocto

int main(int argc, const char** argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: ref_render scene.vptscene out.vptarch [opts]\n");
    return 2;
  }
  auto params    = pathtrace_params{};
  auto rays_file = string{};
  auto instance  = -1;
  auto dump_state = false;
  auto dump_host  = false;
  for (int i = 3; i < argc; i++) {
    auto a    = string{argv[i]};
    auto next = [&]() -> string {
      if (i + 1 >= argc) {
        fprintf(stderr, "ref_render: missing value for %s\n", a.c_str());
        exit(2);
      }
      return argv[++i];
    };
    if (a == "--shader") {
      auto name  = next();
      auto found = false;
      for (size_t k = 0; k < pathtrace_shader_names.size(); k++)
        if (pathtrace_shader_names[k] == name) {
          params.shader = (pathtrace_shader_type)k;
          found         = true;
        }
      if (!found) {
        fprintf(stderr, "ref_render: unknown shader %s\n", name.c_str());
        return 2;
      }
    } else if (a == "--samples")
      params.samples = std::stoi(next());
    else if (a == "--resolution")
      params.resolution = std::stoi(next());
    else if (a == "--bounces")
      params.bounces = std::stoi(next());
    else if (a == "--camera")
      params.camera = std::stoi(next());
    else if (a == "--stmaxiter")
      params.spheretrace_maxiter = std::stoi(next());
    else if (a == "--noimplicitmis")
      params.noimplicit_mis = true;
    else if (a == "--noparallel")
      params.noparallel = true;
    else if (a == "--rays")
      rays_file = next();
    else if (a == "--instance")
      instance = std::stoi(next());
    else if (a == "--dump-state")
      dump_state = true;
    else if (a == "--dump-host")
      dump_host = true;
    else {
      fprintf(stderr, "ref_render: unknown option %s\n", a.c_str());
      return 2;
    }
  }

  auto error = string{};
  auto in    = vpt::archive{};
  if (!in.load(argv[1], error)) {
    fprintf(stderr, "ref_render: %s\n", error.c_str());
    return 1;
  }
  auto scene = scene_data{};
  if (!vptref::scene_from_archive(in, scene, error)) {
    fprintf(stderr, "ref_render: %s\n", error.c_str());
    return 1;
  }
  auto out = vpt::archive{};
  auto bvh = make_bvh(scene, params);

  if (dump_host) {
    // host-built structures of the reference: bvh nodes / primitives (make_bvh,
    // yocto_bvh.cpp:574), light list + cdfs (make_lights, yocto_pathtrace.cpp:983)
    // and the initial per-pixel generators (make_state, :960)
    static_assert(sizeof(bvh_node) == 32, "bvh_node layout");
    auto put_bvh = [&](const string& pre, const bvh_data& b) {
      out.add(pre + "nodes", vpt::dt_u8, {(uint64_t)b.nodes.size(), 32}, b.nodes.data());
      out.add_i32(pre + "prims", {(uint64_t)b.primitives.size()}, b.primitives.data());
    };
    put_bvh("bvh.scene.", bvh);
    for (size_t i = 0; i < bvh.shapes.size(); i++)
      put_bvh("bvh.shape" + std::to_string(i) + ".", bvh.shapes[i]);
    auto lights = make_lights(scene, params);
    auto ids    = vector<int>(lights.lights.size() * 3);
    for (size_t i = 0; i < lights.lights.size(); i++) {
      auto& l        = lights.lights[i];
      ids[i * 3 + 0] = l.instance, ids[i * 3 + 1] = l.environment, ids[i * 3 + 2] = l.sdf;
      out.add_f32("lights.cdf" + std::to_string(i), {(uint64_t)l.elements_cdf.size()},
          l.elements_cdf.data());
    }
    out.add_i32("lights.ids", {(uint64_t)lights.lights.size(), 3}, ids.data());
    auto state = make_state(scene, params);
    int  wh[2] = {state.width, state.height};
    out.add_i32("state.size", {2}, wh);
    out.add("state.rngs", vpt::dt_u64, {(uint64_t)state.rngs.size(), 2}, state.rngs.data());
  } else if (!rays_file.empty()) {
    auto ra = vpt::archive{};
    if (!ra.load(rays_file, error)) {
      fprintf(stderr, "ref_render: %s\n", error.c_str());
      return 1;
    }
    auto rays = ra.find("rays");
    if (!rays || rays->dims.size() != 2 || rays->dims[1] != 8) {
      fprintf(stderr, "ref_render: rays must be f32[n,8]\n");
      return 1;
    }
    auto n   = (size_t)rays->dims[0];
    auto ids = vector<int>(n * 3);
    auto fl  = vector<float>(n * 3);
    auto sid = vector<int>(n * 3);
    auto sfl = vector<float>(n * 4);
    auto has_sdf = !scene.sdfs.empty() || !scene.vol_instances.empty();
    for (size_t i = 0; i < n; i++) {
      auto p   = rays->f32() + i * 8;
      auto ray = ray3f{{p[0], p[1], p[2]}, {p[3], p[4], p[5]}, p[6], p[7]};
      auto isec = instance >= 0 ? intersect_bvh(bvh, scene, instance, ray)
                                : intersect_bvh(bvh, scene, ray);
      ids[i * 3 + 0] = isec.hit ? 1 : 0;
      ids[i * 3 + 1] = isec.hit ? isec.instance : -1;
      ids[i * 3 + 2] = isec.hit ? isec.element : -1;
      fl[i * 3 + 0]  = isec.hit ? isec.uv.x : 0;
      fl[i * 3 + 1]  = isec.hit ? isec.uv.y : 0;
      fl[i * 3 + 2]  = isec.hit ? isec.distance : 0;
      if (has_sdf) {
        auto st = spheretrace(scene, ray, params.spheretrace_maxiter);
        sid[i * 3 + 0] = st.hit ? 1 : 0;
        sid[i * 3 + 1] = st.hit ? st.instance : -1;
        sid[i * 3 + 2] = st.hit ? st.sdf : -1;
        auto nrm       = vec3f{0, 0, 0};
        if (st.hit) {
          auto pos = ray_point(ray, st.dist);
          if (st.instance != invalidid) {
            auto& vi = scene.vol_instances[st.instance];
            nrm = eval_sdf_normal(scene.volumes[vi.volume], vi, pos, st.dist);
          } else {
            nrm = eval_sdf_normal(scene.sdfs[st.sdf], pos, st.dist);
          }
        }
        sfl[i * 4 + 0] = st.hit ? st.dist : 0;
        sfl[i * 4 + 1] = nrm.x, sfl[i * 4 + 2] = nrm.y, sfl[i * 4 + 3] = nrm.z;
      }
    }
    out.add_i32("hits.ids", {n, 3}, ids.data());
    out.add_f32("hits.f", {n, 3}, fl.data());
    if (has_sdf) {
      out.add_i32("sdfhits.ids", {n, 3}, sid.data());
      out.add_f32("sdfhits.f", {n, 4}, sfl.data());
    }
  } else {
    auto lights = make_lights(scene, params);
    auto state  = make_state(scene, params);
    if (dump_state) {
      out.add("rngs", vpt::dt_u64, {(uint64_t)state.rngs.size(), 2},
          state.rngs.data());
    }
    auto t0 = std::chrono::steady_clock::now();
    for (auto s = 0; s < params.samples; s++)
      pathtrace_samples(state, scene, bvh, lights, params);
    auto t1 = std::chrono::steady_clock::now();
    auto seconds = (float)std::chrono::duration<double>(t1 - t0).count();
    auto image   = get_render(state);
    out.add_f32("image",
        {(uint64_t)state.height, (uint64_t)state.width, 4},
        (const float*)image.pixels.data());
    out.add_f32("image_sum",
        {(uint64_t)state.height, (uint64_t)state.width, 4},
        (const float*)state.image.data());
    if (dump_state) {
      out.add("rngs_after", vpt::dt_u64, {(uint64_t)state.rngs.size(), 2},
          state.rngs.data());
      // what save_image encodes for an 8-bit format (yocto_sceneio.cpp:523-529):
      // rgb_to_srgb + float_to_byte of the get_render image
      auto ldr = std::vector<vec4b>{};
      rgb_to_srgb(ldr, image.pixels);
      out.add("image_srgb8", vpt::dt_u8, {(uint64_t)state.height, (uint64_t)state.width, 4},
          ldr.data());
      // what the viewer displays (ypathtrace.cpp:170, :186): tonemap(hdr, exposure, filmic)
      // (yocto_color.h:306-316) and the 8-bit conversion, for two settings
      auto tm = [&](const char* name, float exposure, bool filmic) {
        auto bytes = std::vector<vec4b>(image.pixels.size());
        for (size_t i = 0; i < bytes.size(); i++)
          bytes[i] = float_to_byte(tonemap(image.pixels[i], exposure, filmic, true));
        out.add(name, vpt::dt_u8, {(uint64_t)state.height, (uint64_t)state.width, 4}, bytes.data());
      };
      tm("image_tonemapped8_filmic_e1.5", 1.5f, true);
      tm("image_tonemapped8_e-0.75", -0.75f, false);
    }
    int meta[4] = {state.width, state.height, state.samples,
        (int)std::thread::hardware_concurrency()};
    out.add_i32("meta", {4}, meta);
    out.add_f32("seconds", {1}, &seconds);
    auto nlights = (int)lights.lights.size();
    out.add_i32("nlights", {1}, &nlights);
    printf("ref_render: %dx%d %d spp shader=%s %.3f s %.4f Mpaths/s (%u threads)\n",
        state.width, state.height, state.samples,
        pathtrace_shader_names[(int)params.shader].c_str(), seconds,
        (double)state.width * state.height * state.samples / seconds / 1e6,
        params.noparallel ? 1u : std::thread::hardware_concurrency());
  }
  if (!out.save(argv[2], error)) {
    fprintf(stderr, "ref_render: %s\n", error.c_str());
    return 1;
  }
  return 0;
}
