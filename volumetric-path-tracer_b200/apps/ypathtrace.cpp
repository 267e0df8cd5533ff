// ypathtrace — offline command line of the B200 path tracer.
//
// Mirrors run() / run_offline() of the reference CLI
// (apps/ypathtrace/ypathtrace.cpp:41-87, :307-341): same flags, defaults and
// call sequence (load_scene -> tesselate_surfaces -> make_bvh -> make_lights ->
// make_state -> pathtrace_samples x N -> get_render -> save_image). Differences:
// the scene is a ".vptscene" archive (asset loading stays with the reference's
// loaders, see tools/vpt_export.cpp); --interactive is unavailable on
// this box (no OpenGL); the image writers are .hdr / .pfm / .ppm / .png.
// Added flags: --pipeline {auto,mega,wavefront}, --batch N, --device K, --device-lights
// (make_lights on the GPU), --gpus N (sample ranges over N GPUs of the box, summed on the first),
// --exposure / --filmic / --pratio (pathtrace_params fields the reference CLI leaves at their
// defaults) with --display IMG / --display-every N: the viewer's display loop
// (ypathtrace.cpp:156-189) without a window.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "vpt_pathtrace.h"

using namespace vpt;

static void print_fatal(const std::string& msg) {
  printf("\n%s\n", msg.c_str());
  exit(1);
}

static float rgb_to_srgb(float rgb) {
  return (rgb <= 0.0031308f) ? 12.92f * rgb : (1 + 0.055f) * std::pow(rgb, 1 / 2.4f) - 0.055f;
}
static uint8_t float_to_byte(float a) {
  auto v = (int)(a * 256);
  return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

static uint32_t crc32_of(const uint8_t* data, size_t n, uint32_t crc = 0) {
  static uint32_t table[256];
  if (!table[1])
    for (uint32_t i = 0; i < 256; i++) {
      auto c = i;
      for (int k = 0; k < 8; k++) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
      table[i] = c;
    }
  crc = ~crc;
  for (size_t i = 0; i < n; i++) crc = table[(crc ^ data[i]) & 255] ^ (crc >> 8);
  return ~crc;
}

// PNG with stored (uncompressed) deflate blocks: no zlib dependency
static bool save_png(const std::string& path, int w, int h, const std::vector<uint8_t>& rgba) {
  auto raw = std::vector<uint8_t>{};
  for (int j = 0; j < h; j++) {
    raw.push_back(0);
    raw.insert(raw.end(), rgba.begin() + (size_t)j * w * 4, rgba.begin() + (size_t)(j + 1) * w * 4);
  }
  auto z = std::vector<uint8_t>{0x78, 0x01};
  uint32_t a = 1, b = 0;
  for (auto v : raw) a = (a + v) % 65521, b = (b + a) % 65521;
  for (size_t pos = 0; pos < raw.size();) {
    auto n = std::min<size_t>(65535, raw.size() - pos);
    z.push_back(pos + n == raw.size() ? 1 : 0);
    z.push_back(n & 255), z.push_back(n >> 8), z.push_back(~n & 255), z.push_back((~n >> 8) & 255);
    z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
    pos += n;
  }
  uint32_t adler = (b << 16) | a;
  for (int k = 3; k >= 0; k--) z.push_back((adler >> (8 * k)) & 255);
  auto fs = fopen(path.c_str(), "wb");
  if (!fs) return false;
  auto chunk = [&](const char* tag, const std::vector<uint8_t>& data) {
    auto buf = std::vector<uint8_t>(tag, tag + 4);
    buf.insert(buf.end(), data.begin(), data.end());
    uint32_t len = (uint32_t)data.size(), crc = crc32_of(buf.data(), buf.size());
    uint8_t  be[4] = {(uint8_t)(len >> 24), (uint8_t)(len >> 16), (uint8_t)(len >> 8), (uint8_t)len};
    fwrite(be, 1, 4, fs);
    fwrite(buf.data(), 1, buf.size(), fs);
    uint8_t ce[4] = {(uint8_t)(crc >> 24), (uint8_t)(crc >> 16), (uint8_t)(crc >> 8), (uint8_t)crc};
    fwrite(ce, 1, 4, fs);
  };
  fwrite("\x89PNG\r\n\x1a\n", 1, 8, fs);
  auto ihdr = std::vector<uint8_t>{(uint8_t)(w >> 24), (uint8_t)(w >> 16), (uint8_t)(w >> 8),
      (uint8_t)w, (uint8_t)(h >> 24), (uint8_t)(h >> 16), (uint8_t)(h >> 8), (uint8_t)h, 8, 6, 0,
      0, 0};
  chunk("IHDR", ihdr), chunk("IDAT", z), chunk("IEND", {});
  fclose(fs);
  return true;
}

// save_image by extension (yocto_sceneio.cpp:509-591): ldr formats get rgb_to_srgb + 8 bit
// `srgb8`: the 8-bit image already converted on the device (get_render_srgb8), or empty
static bool save_image(const std::string& path, const color_image& image, std::string& error,
    const std::vector<uint8_t>& srgb8 = {}) {
  auto ext = path.substr(path.find_last_of('.') == std::string::npos ? path.size()
                                                                      : path.find_last_of('.'));
  auto w = image.width, h = image.height;
  if (ext == ".hdr" || ext == ".pfm") {
    auto fs = fopen(path.c_str(), "wb");
    if (!fs) return error = "cannot create " + path, false;
    if (ext == ".pfm") {
      fprintf(fs, "PF\n%d %d\n-1.0\n", w, h);
      for (int j = h - 1; j >= 0; j--)
        for (int i = 0; i < w; i++) fwrite(&image.pixels[(size_t)j * w + i], 4, 3, fs);
    } else {
      fprintf(fs, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
      for (auto& p : image.pixels) {
        auto    m = std::fmax(p.x, std::fmax(p.y, p.z));
        uint8_t rgbe[4] = {0, 0, 0, 0};
        if (m > 1e-32f) {
          int  e;
          auto s  = std::frexp(m, &e) * 256.0f / m;
          rgbe[0] = (uint8_t)(p.x * s), rgbe[1] = (uint8_t)(p.y * s), rgbe[2] = (uint8_t)(p.z * s);
          rgbe[3] = (uint8_t)(e + 128);
        }
        fwrite(rgbe, 1, 4, fs);
      }
    }
    fclose(fs);
    return true;
  }
  auto bytes = srgb8;
  if (bytes.size() != (size_t)w * h * 4) {
    bytes.resize((size_t)w * h * 4);
    for (size_t i = 0; i < (size_t)w * h; i++) {
      auto& p          = image.pixels[i];
      bytes[4 * i + 0] = float_to_byte(rgb_to_srgb(p.x));
      bytes[4 * i + 1] = float_to_byte(rgb_to_srgb(p.y));
      bytes[4 * i + 2] = float_to_byte(rgb_to_srgb(p.z));
      bytes[4 * i + 3] = float_to_byte(p.w);
    }
  }
  if (ext == ".png") {
    if (!save_png(path, w, h, bytes)) return error = "cannot create " + path, false;
    return true;
  }
  if (ext == ".ppm") {
    auto fs = fopen(path.c_str(), "wb");
    if (!fs) return error = "cannot create " + path, false;
    fprintf(fs, "P6\n%d %d\n255\n", w, h);
    for (size_t i = 0; i < (size_t)w * h; i++) fwrite(&bytes[4 * i], 1, 3, fs);
    fclose(fs);
    return true;
  }
  error = "unsupported image format " + ext + " (use .png, .ppm, .hdr or .pfm)";
  return false;
}

int main(int argc, const char** argv) {
  auto scenename   = std::string{"scene.vptscene"};
  auto outname     = std::string{"out.png"};
  auto interactive = false;
  auto device      = 0;
  auto device_lights = false;
  auto display       = std::string{};  // headless stand-in for the viewer's display loop
  auto display_every = 0;
  auto params      = pathtrace_params{};
  params.batch     = 16;
  auto usage       = [&]() {
    printf("usage: ypathtrace [--scene S.vptscene] [--output IMG] [--interactive]\n"
           "  [--resolution 1..4096] [--shader NAME] [--samples 1..4096] [--bounces 1..128]\n"
           "  [--noparallel] [--noimplicitmis] [--stmaxiter 1..512]\n"
           "  [--pipeline auto|mega|wavefront] [--batch N] [--device K] [--device-lights]\n"
           "  [--device-bvh] [--gpus N] [--exposure E] [--filmic] [--pratio P] [--display IMG.ppm|png]\n"
           "  [--display-every N]\n  shaders:");
    for (auto& n : pathtrace_shader_names) printf(" %s", n.c_str());
    printf("\n");
  };
  auto ranged = [&](const char* name, const std::string& v, int lo, int hi) {
    auto x = std::stoi(v);
    if (x < lo || x > hi) print_fatal(std::string("bad value for ") + name);
    return x;
  };
  for (int i = 1; i < argc; i++) {
    auto a    = std::string{argv[i]};
    auto next = [&]() -> std::string {
      if (i + 1 >= argc) print_fatal("missing value for " + a);
      return argv[++i];
    };
    if (a == "--scene") scenename = next();
    else if (a == "--output") outname = next();
    else if (a == "--interactive") interactive = true;
    else if (a == "--resolution") params.resolution = ranged("resolution", next(), 1, 4096);
    else if (a == "--samples") params.samples = ranged("samples", next(), 1, 4096);
    else if (a == "--bounces") params.bounces = ranged("bounces", next(), 1, 128);
    else if (a == "--stmaxiter") params.spheretrace_maxiter = ranged("stmaxiter", next(), 1, 512);
    else if (a == "--noparallel") params.noparallel = true;
    else if (a == "--noimplicitmis") params.noimplicit_mis = true;
    else if (a == "--batch") params.batch = ranged("batch", next(), 1, 4096);
    else if (a == "--device") device = std::stoi(next());
    else if (a == "--device-lights") device_lights = true;
    else if (a == "--device-bvh") params.device_bvh = true;
    else if (a == "--gpus") params.gpus = ranged("gpus", next(), 1, 64);
    else if (a == "--exposure") params.exposure = std::stof(next());
    else if (a == "--filmic") params.filmic = true;
    else if (a == "--pratio") params.pratio = ranged("pratio", next(), 1, 64);
    else if (a == "--display") display = next();
    else if (a == "--display-every") display_every = ranged("display-every", next(), 1, 4096);
    else if (a == "--shader") {
      auto name = next(), found = std::string{};
      for (size_t k = 0; k < pathtrace_shader_names.size(); k++)
        if (pathtrace_shader_names[k] == name) params.shader = (pathtrace_shader_type)k, found = name;
      if (found.empty()) print_fatal("unknown shader " + name);
    } else if (a == "--pipeline") {
      auto name = next();
      if (name == "auto") params.pipeline = pathtrace_pipeline::automatic;
      else if (name == "mega") params.pipeline = pathtrace_pipeline::megakernel;
      else if (name == "wavefront") params.pipeline = pathtrace_pipeline::wavefront;
      else print_fatal("unknown pipeline " + name);
    } else if (a == "--help" || a == "-h") {
      usage();
      return 0;
    } else {
      usage();
      print_fatal("unknown option " + a);
    }
  }
  if (interactive) print_fatal("interactive view unavailable on this box (no OpenGL)");

  try {
    auto clock = [] { return std::chrono::steady_clock::now(); };
    auto since = [&](auto t0) { return std::chrono::duration<double>(clock() - t0).count(); };
    if (vpt_cuda_set_device(device) != VPT_OK) print_fatal(vpt_cuda_last_error());
    auto t0    = clock();
    auto error = std::string{};
    auto scene = scene_data{};
    if (!load_scene(scenename, scene, error)) print_fatal(error);
    printf("load scene: %.3f s\n", since(t0));
    if (!scene.subdivs.empty()) {  // archives exported with --keep-subdivs
      t0 = clock();
      auto count = scene.subdivs.size();
      tesselate_surfaces(scene);
      printf("tesselate surfaces (device): %.3f s (%zu subdivs)\n", since(t0), count);
    }
    t0       = clock();
    auto bvh = make_bvh(scene, params);
    printf("build bvh: %.3f s\n", since(t0));
    t0          = clock();
    auto lights = pathtrace_lights{};
    auto state  = pathtrace_state{};
    auto dscene = std::shared_ptr<device_scene>{};
    if (device_lights) {
      // make_lights on the device: the host object stays empty and only keys the cache
      auto stats  = vpt_lights_stats{};
      dscene      = upload_scene_device_lights(scene, bvh, &stats);
      printf("build lights (device): %.3f s (%d lights, %llu cdf entries: weights %.2f ms, "
             "running sums %.2f ms)\n", since(t0), stats.num_lights,
          (unsigned long long)stats.cdf_elements, stats.weights_ms, stats.scan_ms);
      state = make_state(scene, params);
      bind_device_scene(state, dscene, scene, bvh, lights);
    } else {
      lights = make_lights(scene, params);
      printf("build lights: %.3f s\n", since(t0));
      t0     = clock();
      dscene = upload_scene(scene, bvh, lights);
      printf("upload scene: %.3f s\n", since(t0));
      state = make_state(scene, params);
      bind_device_scene(state, dscene, scene, bvh, lights);
    }
    if (params.gpus > 1) {
      t0 = clock();
      bind_device_replicas(state, scene, bvh, lights, params.gpus);
      printf("set up %d GPUs: %.3f s\n", params.gpus, since(t0));
    }
    // The viewer's display loop without a window (apps/ypathtrace/ypathtrace.cpp:156-189): a
    // one-sample preview at resolution / pratio blown up to the frame, then the tonemapped
    // frame every `display_every` samples - tonemap + 8-bit conversion run on the device, the
    // bytes arrive through the state's pinned buffer and go to the --display file
    auto show = [&](const std::vector<uint8_t>& rgba, int w, int h) {
      auto none = color_image{};
      none.width = w, none.height = h;
      if (!save_image(display, none, error, rgba)) print_fatal(error);
    };
    if (!display.empty()) {
      auto pparams = params;
      pparams.resolution = std::max(1, params.resolution / params.pratio);
      pparams.samples = 1, pparams.batch = 1, pparams.gpus = 1;
      auto pstate = make_state(scene, pparams);
      bind_device_scene(pstate, dscene, scene, bvh, lights);
      pathtrace_samples(pstate, scene, bvh, lights, pparams);
      auto small = get_render_tonemapped8(pstate, params.exposure, params.filmic);
      auto frame = std::vector<uint8_t>((size_t)state.width * state.height * 4);
      for (auto idx = 0; idx < state.width * state.height; idx++) {
        auto i  = idx % state.width, j = idx / state.width;
        auto pi = std::min(i / params.pratio, pstate.width - 1),
             pj = std::min(j / params.pratio, pstate.height - 1);
        memcpy(&frame[(size_t)idx * 4], &small[((size_t)pj * pstate.width + pi) * 4], 4);
      }
      show(frame, state.width, state.height);
      printf("preview: %dx%d, 1 sample -> %s\n", pstate.width, pstate.height, display.c_str());
    }
    t0         = clock();
    int shown  = 0;
    while (state.samples < params.samples) {
      pathtrace_samples(state, scene, bvh, lights, params);
      printf("\rrender samples: %d/%d", state.samples, params.samples);
      fflush(stdout);
      if (!display.empty() && display_every > 0 && state.samples - shown >= display_every) {
        show(get_render_tonemapped8(state, params.exposure, params.filmic), state.width, state.height);
        shown = state.samples;
      }
    }
    if (!display.empty())
      show(get_render_tonemapped8(state, params.exposure, params.filmic), state.width, state.height);
    auto render  = get_render(state);
    auto seconds = since(t0);
    printf("\nrender: %.3f s, %.2f Mpaths/s (%dx%d, %d spp, %s, %d GPU%s)\n", seconds,
        (double)state.width * state.height * state.samples / seconds / 1e6, state.width,
        state.height, state.samples, pathtrace_shader_names[(int)params.shader].c_str(),
        params.gpus, params.gpus > 1 ? "s" : "");
    auto ldr = outname.size() > 4 && (outname.substr(outname.size() - 4) == ".png" ||
                                         outname.substr(outname.size() - 4) == ".ppm");
    if (!save_image(outname, render, error, ldr ? get_render_srgb8(state) : std::vector<uint8_t>{}))
      print_fatal(error);
    printf("saved %s\n", outname.c_str());
  } catch (std::exception& e) {
    print_fatal(e.what());
  }
  return 0;
}
