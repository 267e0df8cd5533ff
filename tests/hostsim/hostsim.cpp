// hostsim — TEST-ONLY debugging aid. Compiles the device headers
// (volumetric-path-tracer_b200/csrc/*.cuh) for the host CPU and runs them one
// pixel at a time, so that the CUDA code's logic can be single-stepped against
// the reference (oracle/_ref/ref_render) on a box without a GPU. With the host
// libm the stochastic shaders replay the reference's random streams exactly,
// which localises any divergence to a line of code.
//
// It is NOT part of the product: nothing under volumetric-path-tracer_b200/
// builds, links or calls this, and the shipped libraries have no CPU path.
//
// usage: hostsim <scene.vptscene> <out.vptarch> [--shader NAME] [--samples N]
//        [--resolution R] [--bounces B] [--noimplicitmis] [--stmaxiter M]
//        [--rays rays.vptarch] [--instance K] [--threads T] [--cull-check] [--sdf-cull-check]
#define VPT_HOST_POOL 1
#define VPT_STACK_CHECK 1
#define VPT_WORK_COUNTERS 1
#define VPT_CULL_CHECK 1
#define VPT_SDF_CULL_CHECK 1
#include <chrono>
#include <cmath>
#include <cstring>
#include <thread>

#include "../../volumetric-path-tracer_b200/csrc/shaders.cuh"
#include "../../volumetric-path-tracer_b200/csrc/upload.cuh"
#include "../../volumetric-path-tracer_b200/host/vpt_archive.h"
#include "../../volumetric-path-tracer_b200/host/vpt_pathtrace.h"

using namespace vpt;

template <int S>
static void render_rows(const DScene& sc, const DParams& prm, pathtrace_state& state, int nsamples,
    int tid, int nthreads) {
  alignas(8) int stack_mem[vpt_stack_depth * vpt_stack_words];
  auto st = stack_t{stack_mem, 1};
  for (int idx = tid; idx < state.width * state.height; idx += nthreads) {
    auto rng = rng_t{state.rngs[idx].state, state.rngs[idx].inc};
    auto acc = state.image[idx];
    for (int s = 0; s < nsamples; s++) {
      auto ray = sample_camera(sc, prm, idx, rng);
      auto rad = guard_radiance(shade_path<S>(sc, prm, ray, rng, st));
      acc      = acc + rad;
    }
    state.image[idx] = acc;
    state.rngs[idx]  = {rng.state, rng.inc};
  }
}

template <int S>
static void render(const DScene& sc, const DParams& prm, pathtrace_state& state, int nsamples,
    int nthreads) {
  auto threads = std::vector<std::thread>{};
  for (int t = 0; t < nthreads; t++)
    threads.emplace_back([&, t] { render_rows<S>(sc, prm, state, nsamples, t, nthreads); });
  for (auto& t : threads) t.join();
}

// `hostsim --selftest-math N [stride]`: the restatements the DEVICE runs instead of libm
// (vmath.cuh: atan / atan2 / acos from fdlibm; glibc_libm.cuh: sin / cos / exp / log / pow from
// glibc's flt-32 algorithms) against this box's libm, bit for bit. Unary functions sweep every
// `stride`-th float bit pattern (stride 1 = all 2^32), pow sweeps x over the same patterns for
// the exponents the shaders use (5, 2.4) plus N random (x, y) pairs; atan2 takes N random pairs.
// Also checks sincosf == (sinf, cosf). Prints the mismatch count; exit status 0 iff it is 0.
static int selftest_math(long n, unsigned stride) {
  auto bits = [](float x) { uint32_t u; memcpy(&u, &x, 4); return u; };
  auto same = [&](float a, float b) { return bits(a) == bits(b) || (a != a && b != b); };
  int  nthreads = (int)std::max(1u, std::thread::hardware_concurrency());
  std::vector<long> bad(nthreads * 16, 0);
  const char* names[9] = {"atanf", "acosf", "sinf", "cosf", "expf", "logf", "powf(x,5)", "powf(x,2.4)",
      "sincosf"};
  auto worker = [&](int t) {
    long* b = &bad[t * 16];
    uint64_t per = ((1ull << 32) / stride + nthreads - 1) / nthreads;
    for (uint64_t k = per * t; k < per * (t + 1) && k * stride < (1ull << 32); k++) {
      uint32_t u = (uint32_t)(k * stride);
      float    x; memcpy(&x, &u, 4);
      b[0] += !same(vatan(x), atanf(x));
      b[1] += !same(vacos(x), acosf(x));
      b[2] += !same(g_sinf(x), sinf(x));
      b[3] += !same(g_cosf(x), cosf(x));
      b[4] += !same(g_expf(x), expf(x));
      b[5] += !same(g_logf(x), logf(x));
      b[6] += !same(g_powf(x, 5.0f), powf(x, 5.0f));
      b[7] += !same(g_powf(x, 2.4f), powf(x, 2.4f));
      float sn, cs; sincosf(x, &sn, &cs);
      b[8] += !same(sn, sinf(x)) + !same(cs, cosf(x));
    }
  };
  std::vector<std::thread> threads;
  for (int t = 0; t < nthreads; t++) threads.emplace_back(worker, t);
  for (auto& t : threads) t.join();
  long total = 0;
  for (int f = 0; f < 9; f++) {
    long c = 0;
    for (int t = 0; t < nthreads; t++) c += bad[t * 16 + f];
    printf("  %-12s %ld mismatches\n", names[f], c);
    total += c;
  }
  uint64_t s = 0x9e3779b97f4a7c15ull;
  auto next = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (uint32_t)(s >> 16); };
  auto anyf = [&]() { uint32_t u = next(); float f; memcpy(&f, &u, 4); return f; };
  auto unit = [&]() { return ((int32_t)next()) / 2147483648.0f; };
  long pairs = 0;
  float edges[] = {0.0f, -0.0f, 1.0f, -1.0f, 0.5f, -0.5f, 2.0f, 3.0f, -3.0f, 1e-30f, 1e30f, 3.4e38f,
      1e-45f, -1e-45f, 33554432.0f, INFINITY, -INFINITY, NAN};
  for (auto a : edges)
    for (auto b : edges) pairs += !same(vatan2(a, b), atan2f(a, b)) + !same(g_powf(a, b), powf(a, b));
  for (long i = 0; i < n; i++) {
    auto a = anyf(), b = anyf(), u = unit(), v = unit();
    pairs += !same(vatan2(a, b), atan2f(a, b)) + !same(vatan2(u, v), atan2f(u, v));
    pairs += !same(g_powf(a, b), powf(a, b)) + !same(g_powf(vabs(u), v * 8), powf(vabs(u), v * 8));
    pairs += !same(g_powf(vabs(a), u * 40), powf(vabs(a), u * 40));
  }
  printf("  %-12s %ld mismatches\n", "atan2f/powf pairs", pairs);
  total += pairs;
  printf("math selftest: %ld mismatches (stride %u, %ld random pairs)\n", total, stride, n);
  return total == 0 ? 0 : 1;
}

int main(int argc, const char** argv) {
  if (argc >= 3 && std::string(argv[1]) == "--selftest-math")
    return selftest_math(atol(argv[2]), argc > 3 ? (unsigned)atol(argv[3]) : 257u);
  if (argc < 3) {
    fprintf(stderr, "usage: hostsim scene.vptscene out.vptarch [opts]\n");
    return 2;
  }
  auto params    = pathtrace_params{};
  auto rays_file = std::string{};
  auto instance  = -1;
  auto nthreads  = (int)std::max(1u, std::thread::hardware_concurrency());
  for (int i = 3; i < argc; i++) {
    auto a    = std::string{argv[i]};
    auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "--shader") {
      auto name = next();
      for (size_t k = 0; k < pathtrace_shader_names.size(); k++)
        if (pathtrace_shader_names[k] == name) params.shader = (pathtrace_shader_type)k;
    } else if (a == "--samples") params.samples = std::stoi(next());
    else if (a == "--resolution") params.resolution = std::stoi(next());
    else if (a == "--bounces") params.bounces = std::stoi(next());
    else if (a == "--camera") params.camera = std::stoi(next());
    else if (a == "--stmaxiter") params.spheretrace_maxiter = std::stoi(next());
    else if (a == "--noimplicitmis") params.noimplicit_mis = true;
    else if (a == "--rays") rays_file = next();
    else if (a == "--instance") instance = std::stoi(next());
    else if (a == "--threads") nthreads = std::stoi(next());
    else if (a == "--cull-check") vpt_cull_check_on = true;
    else if (a == "--sdf-cull-check") vpt_sdf_cull_check_on = true;
    else if (a == "--sdf-cull") vpt_sdf_cull_mode = std::stoi(next());
  }
  auto error = std::string{};
  auto scene = scene_data{};
  if (!load_scene(argv[1], scene, error)) {
    fprintf(stderr, "hostsim: %s\n", error.c_str());
    return 1;
  }
  auto bvh     = make_bvh(scene, params);
  auto lights  = make_lights(scene, params);
  auto storage = scene_desc_storage{};
  make_scene_desc(storage, scene, bvh, lights);
  auto flat = flat_scene{};
  if (flatten_scene(storage.desc, flat, error) != VPT_OK) {
    fprintf(stderr, "hostsim: %s\n", error.c_str());
    return 1;
  }
  auto& sc  = flat.d;
  auto  out = archive{};
  if (!rays_file.empty()) {
    auto ra = archive{};
    if (!ra.load(rays_file, error)) return 1;
    auto rays = ra.find("rays");
    auto n    = (size_t)rays->dims[0];
    auto ids  = std::vector<int>(n * 3);
    auto fl   = std::vector<float>(n * 3);
    auto sid  = std::vector<int>(n * 3);
    auto sfl  = std::vector<float>(n * 4);
    alignas(8) int stack_mem[vpt_stack_depth * vpt_stack_words];
    auto st      = stack_t{stack_mem, 1};
    auto has_sdf = sc.num_sdfs + sc.num_vol_instances > 0;
    for (size_t i = 0; i < n; i++) {
      auto p   = rays->f32() + i * 8;
      auto ray = ray3{{p[0], p[1], p[2]}, {p[3], p[4], p[5]}, p[6], p[7]};
      auto h   = instance >= 0 ? intersect_instance(sc, instance, ray, st)
                               : intersect_scene(sc, ray, st);
      ids[i * 3 + 0] = h.hit, ids[i * 3 + 1] = h.hit ? h.instance : -1;
      ids[i * 3 + 2] = h.hit ? h.element : -1;
      fl[i * 3 + 0] = h.hit ? h.uv.x : 0, fl[i * 3 + 1] = h.hit ? h.uv.y : 0;
      fl[i * 3 + 2] = h.hit ? h.distance : 0;
      if (has_sdf) {
        auto r   = spheretrace(sc, ray, params.spheretrace_maxiter);
        auto nrm = f3{0, 0, 0};
        if (r.hit) {
          auto pos = ray_point(ray, r.dist);
          nrm = r.instance != invalidid
                    ? eval_sdf_normal_grid(sc, ld_vol_instance(sc, r.instance), pos, r.dist)
                    : eval_sdf_normal_fn(ld_sdf(sc, r.sdf), pos, r.dist);
        }
        sid[i * 3 + 0] = r.hit, sid[i * 3 + 1] = r.hit ? r.instance : -1;
        sid[i * 3 + 2] = r.hit ? r.sdf : -1;
        sfl[i * 4 + 0] = r.hit ? r.dist : 0;
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
    auto state = make_state(scene, params);
    auto prm   = DParams{params.camera, (int)params.shader, params.samples, params.bounces,
        params.noimplicit_mis ? 1 : 0, params.spheretrace_maxiter, state.width, state.height};
    auto t0 = std::chrono::steady_clock::now();
    switch ((int)params.shader) {
      case 0: render<0>(sc, prm, state, params.samples, nthreads); break;
      case 1: render<1>(sc, prm, state, params.samples, nthreads); break;
      case 2: render<2>(sc, prm, state, params.samples, nthreads); break;
      case 3: render<3>(sc, prm, state, params.samples, nthreads); break;
      case 4: render<4>(sc, prm, state, params.samples, nthreads); break;
      case 5: render<5>(sc, prm, state, params.samples, nthreads); break;
      case 6: render<6>(sc, prm, state, params.samples, nthreads); break;
      case 7: render<7>(sc, prm, state, params.samples, nthreads); break;
      default: render<8>(sc, prm, state, params.samples, nthreads); break;
    }
    auto seconds = (float)std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    out.add_f32("image_sum", {(uint64_t)state.height, (uint64_t)state.width, 4},
        (const float*)state.image.data());
    out.add("rngs_after", dt_u64, {(uint64_t)state.rngs.size(), 2}, state.rngs.data());
    int meta[4] = {state.width, state.height, params.samples, nthreads};
    out.add_i32("meta", {4}, meta);
    out.add_f32("seconds", {1}, &seconds);
    {
      // algorithmic work per camera path (feeds bench.py's roofline)
      double paths = (double)state.width * state.height * params.samples;
      long long v[12] = {g_work.box_tests, g_work.tri_tests, g_work.inst_transforms,
          g_work.scene_queries, g_work.instance_queries, g_work.sdf_scene_evals,
          g_work.sdf_fn_evals, g_work.sdf_grid_evals, g_work.trilinear, g_work.bounces,
          g_work.lightpdf_calls, g_work.texel_reads};
      const char* names[12] = {"box_tests", "tri_tests", "inst_transforms", "scene_queries",
          "instance_queries", "sdf_scene_evals", "sdf_fn_evals", "sdf_grid_evals", "trilinear",
          "bounces", "lightpdf_calls", "texel_reads"};
      float per_path[12];
      printf("hostsim work per path:");
      for (int i = 0; i < 12; i++) {
        per_path[i] = (float)(v[i] / paths);
        printf(" %s=%.3f", names[i], per_path[i]);
      }
      printf(" iq_box_tests=%.3f iq_tri_tests=%.3f node_visits=%.3f cull_tests=%.3f\n",
          g_work.iq_box_tests / paths, g_work.iq_tri_tests / paths, g_work.node_visits / paths,
          g_work.cull_tests / paths);
      printf("hostsim traversal per path: rounds=%.2f pops=%.2f leaf_pops=%.2f stale_leaf_pops=%.2f "
             "wide_visits=%.2f wide_slot_tests=%.2f stale_node_pops=%.2f\n",
          g_work.rounds / paths, g_work.pops / paths, g_work.leaf_pops / paths, g_work.stale_pops / paths,
          g_work.wide_visits / paths, g_work.wide_slot_tests / paths, g_work.stale_node_pops / paths);
      if (g_work.sdf_scene_evals > 0)
        printf("hostsim sdf culling: %.2f fields skipped per scene evaluation (%lld evaluations)\n",
            (double)g_work.sdf_skips / (double)g_work.sdf_scene_evals, (long long)g_work.sdf_scene_evals);
      if (vpt_sdf_cull_check_on)
        printf("hostsim: sdf cull check: %lld fields skipped, %lld of them at or below the running "
               "minimum\n", (long long)g_work.sdf_skips, (long long)g_work.sdf_skips_wrong);
      if (vpt_cull_check_on)
        printf("hostsim: cull check: %lld instances culled, %lld of them hold a hit the reference "
               "would find\n", (long long)g_work.culled, (long long)g_work.cull_wrong);
      out.add_f32("work_per_path", {12}, per_path);
    }
    printf("hostsim: %dx%d %d spp shader=%s %.3f s\n", state.width, state.height, params.samples,
        pathtrace_shader_names[(int)params.shader].c_str(), seconds);
  }
  printf("hostsim: traversal stack high water %d entries, uploader bound %d\n",
      vpt_stack_high_water + 1, flat.stack_depth);
  if (!out.save(argv[2], error)) return 1;
  return 0;
}
