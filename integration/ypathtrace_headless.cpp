// ypathtrace_headless — INTEGRATION TEST DRIVER (not product code): the offline half of the
// reference CLI (apps/ypathtrace/ypathtrace.cpp:41-87, :307-337: same flags, same call sequence)
// on top of the reference's own loaders and host builders, with pathtrace_samples / get_render
// coming from whichever yocto_pathtrace it is linked with:
//   * integration/yocto_pathtrace_cuda.cpp + libvpt_cuda.so  -> ypathtrace_headless_cuda
//   * the reference's CPU yocto_pathtrace.cpp                  -> ypathtrace_headless_cpu
// Besides --output (the reference's save_image) it can --dump the linear float render, the
// un-normalised sums and the per-pixel generators as a named-array archive, which is what
// tests/test_gpu_integration.py compares bit for bit between the two builds.
#include <yocto/yocto_cli.h>
#include <yocto/yocto_image.h>
#include <yocto/yocto_scene.h>
#include <yocto/yocto_sceneio.h>
#include <yocto_pathtrace/yocto_pathtrace.h>

#include "vpt_archive.h"
#if defined(VPT_ADAPTER)
#include "yocto_pathtrace_cuda.h"
#endif

using namespace yocto;
using namespace std::string_literals;

static void run(const vector<string>& args) {
  auto params      = pathtrace_params{};
  auto filename    = "scene.json"s;
  auto output      = ""s;
  auto dump        = ""s;
  auto interactive = false;
  auto cli         = make_cli("ypathtrace", "Raytrace scenes.");
  add_option(cli, "scene", filename, "Scene filename.");
  add_option(cli, "output", output, "Output filename.");
  add_option(cli, "dump", dump, "Archive with the float render, sums and generators.");
  add_option(cli, "interactive", interactive, "Run interactively.");
  add_option(cli, "resolution", params.resolution, "Image resolution.", {1, 4096});
  add_option(cli, "shader", params.shader, "Shader type.", pathtrace_shader_names);
  add_option(cli, "samples", params.samples, "Number of samples.", {1, 4096});
  add_option(cli, "bounces", params.bounces, "Number of bounces.", {1, 128});
  add_option(cli, "noparallel", params.noparallel, "Disable threading.");
  add_option(cli, "noimplicitmis", params.noimplicit_mis, "Disable MIS on implicit shader");
  add_option(cli, "stmaxiter", params.spheretrace_maxiter,
      "Number of maximum iteration while spheretracing", {1, 512});
  parse_cli(cli, args);
  if (interactive) print_fatal("--interactive needs OpenGL: unavailable on this box");

  auto error = string{};
  auto scene = scene_data{};
  if (!load_scene(filename, scene, error)) print_fatal(error);
#if defined(VPT_ADAPTER)
  pathtrace_cuda_scene_file(filename);  // the one line the adapter adds to run_offline
#endif
  tesselate_surfaces(scene);
  auto bvh    = make_bvh(scene, params);
  auto lights = make_lights(scene, params);
  auto state  = make_state(scene, params);
  for (auto sample = 0; sample < params.samples; sample++)
    pathtrace_samples(state, scene, bvh, lights, params);
  auto render = get_render(state);
  if (!output.empty() && !save_image(output, render, error)) print_fatal(error);
  if (!dump.empty()) {
#if defined(VPT_ADAPTER)
    pathtrace_cuda_sync(state);
#endif
    auto ar = vpt::archive{};
    ar.add_f32("render", {(uint64_t)state.height, (uint64_t)state.width, 4},
        (const float*)render.pixels.data());
    ar.add_f32("image_sum", {(uint64_t)state.height, (uint64_t)state.width, 4},
        (const float*)state.image.data());
    ar.add("rngs_after", vpt::dt_u64, {(uint64_t)state.rngs.size(), 2}, state.rngs.data());
    int meta[3] = {state.width, state.height, state.samples};
    ar.add_i32("meta", {3}, meta);
    if (!ar.save(dump, error)) print_fatal(error);
  }
#if defined(VPT_ADAPTER)
  pathtrace_cuda_release(state);
#endif
}

int main(int argc, const char* argv[]) { handle_errors(run, make_cli_args(argc, argv)); }
