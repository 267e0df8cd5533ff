// vpt_cuda.cu — kernels and the extern "C" layer of include/vpt_cuda.h.
//
// Built for sm_100a only, with -fmad=false so that the traversal and the
// deterministic shaders reproduce the reference's IEEE arithmetic bit for bit.
// There is no CPU fallback in this library: every compute entry point needs a
// CUDA device and fails with VPT_ERR_NO_DEVICE otherwise.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/vpt_cuda.h"
#include "shaders.cuh"
#include "tiling.cuh"
#include "upload.cuh"
#include "wavefront.cuh"
#include "bvh_build.cuh"
#include "tesselate.cuh"

using namespace vpt;

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
static thread_local std::string g_last_error;

static int set_error(int code, const std::string& msg) {
  g_last_error = msg;
  return code;
}
static int cuda_fail(cudaError_t e, const char* what) {
  return set_error(e == cudaErrorMemoryAllocation ? VPT_ERR_OUT_OF_MEMORY : VPT_ERR_CUDA,
      std::string(what) + ": " + cudaGetErrorString(e));
}
#define VPT_CUDA_TRY(call)                               \
  do {                                                   \
    cudaError_t e_ = (call);                             \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call);  \
  } while (0)

static int require_device() {
  int  n = 0;
  auto e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return set_error(VPT_ERR_NO_DEVICE,
        "no CUDA device available: the B200 path has no CPU fallback");
  }
  return VPT_OK;
}

// ---------------------------------------------------------------------------
// opaque handles
// ---------------------------------------------------------------------------
struct vpt_scene {
  flat_scene flat;
  int        device = 0;
};

struct vpt_state {
  int             width = 0, height = 0, samples = 0;
  int             device = 0;
  f4*             image  = nullptr;
  rng_t*          rngs   = nullptr;
  wavefront_buffers wf;
  cudaEvent_t     ev0 = nullptr, ev1 = nullptr;
  vpt_trace_stats stats{};
  // multi-GPU accumulation: states whose images are summed into this one's downloads
  std::vector<vpt_state*> peers;
  f4*             reduced = nullptr;  // root image + peers' images (root's device)
  f4*             staging = nullptr;  // a peer image copied over when P2P loads are unavailable
  uchar4*         out8    = nullptr;  // device side of download_srgb8
  uint8_t*        pinned8 = nullptr;  // pinned host side of download_srgb8
};

// ---------------------------------------------------------------------------
// megakernel: one thread owns one pixel and runs whole paths for `nsamples`
// samples; the pixel's generator never leaves registers.
// ---------------------------------------------------------------------------
constexpr int mega_block = 128;

template <int S>
__global__ void __launch_bounds__(mega_block)
    k_render_mega(const DScene sc, const DParams prm, f4* __restrict__ image,
        rng_t* __restrict__ rngs, int npixels, int nsamples) {
  extern __shared__ int stack_mem[];
  // 8x4 pixel tiles per warp keep neighbouring lanes on neighbouring rays
  int tid = blockIdx.x * blockDim.x + threadIdx.x;
  int idx = tile_pixel_index(tid, prm.width, prm.height);
  if (idx < 0) return;
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  auto    rng = rngs[idx];
  auto    acc = image[idx];
  for (int s = 0; s < nsamples; s++) {
    auto ray = sample_camera(sc, prm, idx, rng);
    auto rad = guard_radiance(shade_path<S>(sc, prm, ray, rng, st));
    acc      = acc + rad;
  }
  image[idx] = acc;
  rngs[idx]  = rng;
}

// first-hit queries for ray batches (parity gate for the traversal)
__global__ void __launch_bounds__(mega_block)
    k_intersect(const DScene sc, const float* __restrict__ rays, int n, int instance,
        int* __restrict__ ids, float* __restrict__ uvd) {
  extern __shared__ int stack_mem[];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  auto    r   = rays + 8 * (size_t)i;
  auto    ray = ray3{{r[0], r[1], r[2]}, {r[3], r[4], r[5]}, r[6], r[7]};
  auto    h   = instance >= 0 ? intersect_instance(sc, instance, ray, st)
                              : intersect_scene(sc, ray, st);
  ids[3 * i + 0] = h.hit ? 1 : 0;
  ids[3 * i + 1] = h.hit ? h.instance : -1;
  ids[3 * i + 2] = h.hit ? h.element : -1;
  uvd[3 * i + 0] = h.hit ? h.uv.x : 0;
  uvd[3 * i + 1] = h.hit ? h.uv.y : 0;
  uvd[3 * i + 2] = h.hit ? h.distance : 0;
}

__global__ void __launch_bounds__(mega_block)
    k_spheretrace(const DScene sc, const float* __restrict__ rays, int n, int maxiter,
        int* __restrict__ ids, float* __restrict__ dn) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  auto r   = rays + 8 * (size_t)i;
  auto ray = ray3{{r[0], r[1], r[2]}, {r[3], r[4], r[5]}, r[6], r[7]};
  auto st  = spheretrace(sc, ray, maxiter);
  auto nrm = f3{0, 0, 0};
  if (st.hit) {
    auto pos = ray_point(ray, st.dist);
    nrm      = st.instance != invalidid
                   ? eval_sdf_normal_grid(sc, ld_vol_instance(sc, st.instance), pos, st.dist)
                   : eval_sdf_normal_fn(ld_sdf(sc, st.sdf), pos, st.dist);
  }
  ids[3 * i + 0] = st.hit ? 1 : 0;
  ids[3 * i + 1] = st.hit ? st.instance : -1;
  ids[3 * i + 2] = st.hit ? st.sdf : -1;
  dn[4 * i + 0]  = st.hit ? st.dist : 0;
  dn[4 * i + 1] = nrm.x, dn[4 * i + 2] = nrm.y, dn[4 * i + 3] = nrm.z;
}

__global__ void k_advance_rngs(rng_t* rngs, int n, unsigned long long delta) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  auto r = rngs[i];
  rng_advance(r, delta);
  rngs[i] = r;
}

// dst[i] += sum over peers of peers[k][i]; peers are P2P-mapped device pointers
struct peer_list {
  const float* p[16];
};
__global__ void k_reduce_images(float* __restrict__ dst, peer_list peers, int npeers, size_t n4) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= n4) return;
  auto acc = reinterpret_cast<float4*>(dst)[i];
  for (int k = 0; k < npeers; k++) {
    auto v = reinterpret_cast<const float4*>(peers.p[k])[i];
    acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
  }
  reinterpret_cast<float4*>(dst)[i] = acc;
}

// ---------------------------------------------------------------------------
// launch helpers
// ---------------------------------------------------------------------------
template <int S>
static cudaError_t launch_mega(const vpt_scene* scene, vpt_state* state, const DParams& prm,
    int nsamples) {
  int    npix   = state->width * state->height;
  int    slots  = tile_slot_count(state->width, state->height);
  int    blocks = (slots + mega_block - 1) / mega_block;
  size_t smem   = (size_t)scene->flat.stack_depth * vpt_stack_words * mega_block * sizeof(int);
  auto   kernel = k_render_mega<S>;
  if (smem > 48 * 1024) {
    auto e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  kernel<<<blocks, mega_block, smem>>>(scene->flat.d, prm, state->image, state->rngs, npix,
      nsamples);
  return cudaGetLastError();
}

static cudaError_t dispatch_mega(const vpt_scene* scene, vpt_state* state, const DParams& prm,
    int nsamples) {
  switch (prm.shader) {
    case shader_volpathtrace: return launch_mega<shader_volpathtrace>(scene, state, prm, nsamples);
    case shader_pathtrace: return launch_mega<shader_pathtrace>(scene, state, prm, nsamples);
    case shader_naive: return launch_mega<shader_naive>(scene, state, prm, nsamples);
    case shader_eyelight: return launch_mega<shader_eyelight>(scene, state, prm, nsamples);
    case shader_normal: return launch_mega<shader_normal>(scene, state, prm, nsamples);
    case shader_texcoord: return launch_mega<shader_texcoord>(scene, state, prm, nsamples);
    case shader_color: return launch_mega<shader_color>(scene, state, prm, nsamples);
    case shader_implicit: return launch_mega<shader_implicit>(scene, state, prm, nsamples);
    default: return launch_mega<shader_implicit_normal>(scene, state, prm, nsamples);
  }
}

// get_render + rgb_to_srgb + float_to_byte (yocto_color.h:207-231), one thread per pixel
__device__ __forceinline__ float out_rgb_to_srgb(float rgb) {
  return (rgb <= 0.0031308f) ? 12.92f * rgb : (1 + 0.055f) * vpt::vpow(rgb, 1 / 2.4f) - 0.055f;
}
__device__ __forceinline__ unsigned char out_float_to_byte(float a) {
  return (unsigned char)vpt::vclamp((int)(a * 256), 0, 255);
}
// `gain` = exp2(exposure), evaluated by the host libm as the reference does (std::exp2,
// yocto_color.h:308), 1 for exposure 0; filmic = the ACES fit of tonemap_filmic
// (yocto_color.h:274-281), expression for expression, no FMA contraction
__global__ void k_output_srgb8(const vpt::f4* __restrict__ image, uchar4* __restrict__ out,
    size_t pixels, float scale, float gain, int apply_gain, int filmic) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= pixels) return;
  auto p   = image[idx] * scale;
  auto rgb = vpt::f3{p.x, p.y, p.z};
  if (apply_gain) rgb = rgb * gain;
  if (filmic) {
    auto hdr = rgb * 0.6f;
    auto ldr = (hdr * hdr * 2.51f + hdr * 0.03f) / (hdr * hdr * 2.43f + hdr * 0.59f + 0.14f);
    rgb = {vpt::vmax(0.0f, ldr.x), vpt::vmax(0.0f, ldr.y), vpt::vmax(0.0f, ldr.z)};
  }
  out[idx] = make_uchar4(out_float_to_byte(out_rgb_to_srgb(rgb.x)),
      out_float_to_byte(out_rgb_to_srgb(rgb.y)), out_float_to_byte(out_rgb_to_srgb(rgb.z)),
      out_float_to_byte(p.w));
}

// ---------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------
extern "C" {

// debug builds only (-DVPT_GUARD): loop-guard flags, zeros otherwise
int vpt_cuda_debug_flags(int32_t out[16]) {
  for (int i = 0; i < 16; i++) out[i] = 0;
#if defined(VPT_GUARD)
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(out, vpt_guard_flags, sizeof(int) * 16);
#endif
  return VPT_OK;
}

int vpt_cuda_device_count(void) {
  int  n = 0;
  auto e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

int vpt_cuda_set_device(int device) {
  if (auto r = require_device()) return r;
  VPT_CUDA_TRY(cudaSetDevice(device));
  return VPT_OK;
}

int vpt_cuda_current_device(void) {
  int  device = 0;
  auto e      = cudaGetDevice(&device);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return device;
}

const char* vpt_cuda_last_error(void) { return g_last_error.c_str(); }

int vpt_cuda_host_register(void* ptr, size_t bytes) {
  if (!ptr || !bytes) return set_error(VPT_ERR_INVALID, "null argument");
  if (auto r = require_device()) return r;
  VPT_CUDA_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
  return VPT_OK;
}

int vpt_cuda_host_unregister(void* ptr) {
  if (!ptr) return set_error(VPT_ERR_INVALID, "null argument");
  VPT_CUDA_TRY(cudaHostUnregister(ptr));
  return VPT_OK;
}

int vpt_cuda_upload_scene(const vpt_scene_desc* desc, vpt_scene** out) {
  if (!desc || !out) return set_error(VPT_ERR_INVALID, "null argument");
  *out = nullptr;
  if (auto r = require_device()) return r;
  auto scene = new (std::nothrow) vpt_scene();
  if (!scene) return set_error(VPT_ERR_OUT_OF_MEMORY, "host allocation failed");
  cudaGetDevice(&scene->device);
  std::string error;
  auto        r = flatten_scene(*desc, scene->flat, error);
  if (r != VPT_OK) {
    delete scene;
    return set_error(r, error);
  }
  *out = scene;
  return VPT_OK;
}

// ---- make_lights on the device (yocto_pathtrace.cpp:983-1049) ------------------------------
// weights: one thread per CDF entry
__global__ void __launch_bounds__(256)
    k_light_weights(const DScene sc, float* __restrict__ cdf, int light) {
  auto l0 = ld4(sc.lights + 2 * light), l1 = ld4(sc.lights + 2 * light + 1);
  int  n  = l1.x;
  int  idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  float w = 0;
  if (l0.x != invalidid) {
    // element area in object space: triangle_area / quad_area (yocto_geometry.h:351-362)
    auto in = ld4(sc.inst_frame + 4 * l0.x);
    int  shape = f2i(in.w);
    auto s0 = ld4(sc.shapes + 2 * shape), s1 = ld4(sc.shapes + 2 * shape + 1);
    auto q  = ld4(sc.elems + s1.y + idx);
    auto p0 = ld3(sc.vpos + 3 * (size_t)(s0.y + q.x)), p1 = ld3(sc.vpos + 3 * (size_t)(s0.y + q.y));
    auto p2 = ld3(sc.vpos + 3 * (size_t)(s0.y + q.z));
    if (s0.x == kind_triangles) {
      w = length(cross(p1 - p0, p2 - p0)) / 2;
    } else {
      auto p3 = ld3(sc.vpos + 3 * (size_t)(s0.y + q.w));
      w = length(cross(p1 - p0, p3 - p0)) / 2 + length(cross(p3 - p2, p1 - p2)) / 2;
    }
  } else if (l0.y != invalidid) {
    // texel weight: max(rgba) * sin(theta) (yocto_pathtrace.cpp:1019-1027)
    auto ev  = ld4(sc.environments + 9 * l0.y + 8);
    auto tex = ld4(sc.textures + f2i(ev.w));
    int  i = idx % tex.x, j = idx / tex.x;
    auto th = (j + 0.5f) * pif / tex.y;
    auto v  = lookup_texture(sc, tex, i, j, false);
    w = vmax(vmax(vmax(v.x, v.y), v.z), v.w) * vsin(th);
  }
  cdf[l0.w + idx] = w;
}
// Running sums in index order, exactly the reference's float additions: the chain of dependent
// FADDs cannot be split (float addition does not associate), so ONE thread per light walks it;
// the other 255 threads of its block stage the values through shared memory in coalesced
// tiles, which takes the DRAM / L2 latency off the chain (a lone thread reading HBM 16 bytes
// at a time ran 34 ns per element; staged: the chain's own 4-cycle adds plus LDS).
constexpr int scan_tile = 4096;
__global__ void __launch_bounds__(256) k_light_scan(const DScene sc, float* __restrict__ cdf) {
  __shared__ __align__(16) float tile[scan_tile];
  int  light = blockIdx.x;
  auto l0 = ld4(sc.lights + 2 * light), l1 = ld4(sc.lights + 2 * light + 1);
  if (l0.z != invalidid) return;  // sdf lights hold one area, set by the host
  float* c   = cdf + l0.w;
  int    n   = l1.x;
  float  acc = 0;  // thread 0 only
  for (int base = 0; base < n; base += scan_tile) {
    int m = n - base < scan_tile ? n - base : scan_tile;
    for (int i = threadIdx.x; i < m; i += blockDim.x) tile[i] = c[base + i];
    __syncthreads();
    if (threadIdx.x == 0) {
      int i = 0;
      if (base == 0) acc = tile[0], i = 1;  // elements_cdf[0] is the first weight itself
      for (; i < m && (i & 15); i++) tile[i] = acc = acc + tile[i];
      // 16 values per step through registers: 4 LDS.128, the 16 dependent adds, 4 STS.128
      for (; i + 16 <= m; i += 16) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = *reinterpret_cast<const float4*>(tile + i + 4 * k);
#pragma unroll
        for (int k = 0; k < 4; k++) {
          v[k].x = acc + v[k].x, v[k].y = v[k].x + v[k].y;
          v[k].z = v[k].y + v[k].z, v[k].w = v[k].z + v[k].w;
          acc = v[k].w;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) *reinterpret_cast<float4*>(tile + i + 4 * k) = v[k];
      }
      for (; i < m; i++) tile[i] = acc = acc + tile[i];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < m; i += blockDim.x) c[base + i] = tile[i];
    __syncthreads();
  }
}

// the "running sums never decrease" flag of every light's CDF (lights record .y, see
// lights.cuh sample_discrete): one block per light
__global__ void __launch_bounds__(256) k_light_sorted(const DScene sc, i4* __restrict__ lights) {
  int  l  = blockIdx.x;
  auto l0 = ld4(sc.lights + 2 * l), l1 = ld4(sc.lights + 2 * l + 1);
  int  ok = 1;
  for (int k = 1 + threadIdx.x; k < l1.x; k += blockDim.x)
    if (!(sc.cdf[l0.w + k] >= sc.cdf[l0.w + k - 1])) ok = 0;
  ok = __syncthreads_and(ok);
  if (threadIdx.x == 0) lights[2 * l + 1].y = ok && l1.x > 0 ? 1 : 0;
}

int vpt_cuda_upload_scene_device_lights(const vpt_scene_desc* desc, vpt_scene** out,
    vpt_lights_stats* stats) {
  if (!desc || !out) return set_error(VPT_ERR_INVALID, "null argument");
  *out = nullptr;
  if (auto r = require_device()) return r;
  // the light list, in the reference's order; CDF ranges are reserved, not filled
  auto is_zero3 = [](const float* v) { return v[0] == 0 && v[1] == 0 && v[2] == 0; };
  std::vector<vpt_light_desc> lights;
  std::vector<float>          sdf_area(desc->num_sdfs > 0 ? desc->num_sdfs : 0);
  for (int i = 0; i < desc->num_instances; i++) {
    auto& inst = desc->instances[i];
    if (inst.material < 0 || inst.material >= desc->num_materials || inst.shape < 0 ||
        inst.shape >= desc->num_shapes)
      return set_error(VPT_ERR_INVALID, "instance references a missing shape/material");
    if (is_zero3(desc->materials[inst.material].emission)) continue;
    auto& sh = desc->shapes[inst.shape];
    if (sh.num_triangles <= 0 && sh.num_quads <= 0) continue;
    if (sh.num_triangles > 0 && sh.num_quads > 0)
      return set_error(VPT_ERR_UNSUPPORTED,
          "emissive shape with both triangles and quads: build its lights on the host");
    lights.push_back({i, -1, -1, nullptr, sh.num_triangles > 0 ? sh.num_triangles : sh.num_quads});
  }
  for (int e = 0; e < desc->num_environments; e++) {
    auto& ev = desc->environments[e];
    if (is_zero3(ev.emission)) continue;
    int n = 0;
    if (ev.emission_tex >= 0) {
      if (ev.emission_tex >= desc->num_textures)
        return set_error(VPT_ERR_INVALID, "environment references a missing texture");
      n = desc->textures[ev.emission_tex].width * desc->textures[ev.emission_tex].height;
    }
    lights.push_back({-1, e, -1, nullptr, n});
  }
  for (int s = 0; s < desc->num_sdfs; s++) {
    auto& sd = desc->sdfs[s];
    if (sd.material < 0 || sd.material >= desc->num_materials)
      return set_error(VPT_ERR_INVALID, "sdf references a missing material");
    if (is_zero3(desc->materials[sd.material].emission)) continue;
    sdf_area[s] = sd.whd[0] * sd.whd[1];  // flat xy box assumed (yocto_pathtrace.cpp:1044)
    lights.push_back({-1, -1, s, &sdf_area[s], 1});
  }
  auto with_lights       = *desc;
  with_lights.num_lights = (int)lights.size();
  with_lights.lights     = lights.data();
  auto r = vpt_cuda_upload_scene(&with_lights, out);
  if (r != VPT_OK) return r;
  auto&       sc  = (*out)->flat.d;
  auto        cdf = const_cast<float*>(sc.cdf);
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0), cudaEventCreate(&e1), cudaEventCreate(&e2);
  uint64_t total = 0;
  cudaEventRecord(e0);
  for (int l = 0; l < (int)lights.size(); l++) {
    if (lights[l].sdf >= 0 || lights[l].num_elements <= 0) continue;
    total += lights[l].num_elements;
    k_light_weights<<<(lights[l].num_elements + 255) / 256, 256>>>(sc, cdf, l);
  }
  cudaEventRecord(e1);
  if (!lights.empty()) k_light_scan<<<(int)lights.size(), 256>>>(sc, cdf);
  if (!lights.empty()) k_light_sorted<<<(int)lights.size(), 256>>>(sc, const_cast<i4*>(sc.lights));
  cudaEventRecord(e2);
  auto e = cudaEventSynchronize(e2);
  if (e == cudaSuccess) e = cudaGetLastError();
  float w_ms = 0, s_ms = 0;
  cudaEventElapsedTime(&w_ms, e0, e1), cudaEventElapsedTime(&s_ms, e1, e2);
  cudaEventDestroy(e0), cudaEventDestroy(e1), cudaEventDestroy(e2);
  if (e != cudaSuccess) {
    vpt_cuda_free_scene(*out);
    *out = nullptr;
    return cuda_fail(e, "device make_lights");
  }
  if (stats) *stats = {(int32_t)lights.size(), total, w_ms, s_ms};
  return VPT_OK;
}

// ---- shape BVH on the device (bvh_build.cuh) ------------------------------------
int vpt_cuda_build_shape_bvh(const vpt_shape_desc* shape, vpt_bvh_node* nodes, size_t node_capacity,
    int32_t* primitives, size_t prim_capacity, int32_t* num_nodes, vpt_bvh_build_stats* stats) {
  if (!shape || !nodes || !primitives || !num_nodes) return set_error(VPT_ERR_INVALID, "null argument");
  if (auto r = require_device()) return r;
  if (shape->num_points > 0 || shape->num_lines > 0)
    return set_error(VPT_ERR_UNSUPPORTED, "point / line primitives");
  // same precedence as the reference: triangles, then quads (yocto_bvh.cpp:532-560)
  int  verts = shape->num_triangles > 0 ? 3 : 4;
  int  n     = shape->num_triangles > 0 ? shape->num_triangles : shape->num_quads;
  auto elems = shape->num_triangles > 0 ? shape->triangles : shape->quads;
  *num_nodes = 0;
  if (n <= 0) return VPT_OK;
  if (!shape->positions || !elems) return set_error(VPT_ERR_INVALID, "null geometry");
  if (prim_capacity < (size_t)n || node_capacity < 1)
    return set_error(VPT_ERR_INVALID, "output buffers too small");
  for (size_t k = 0; k < (size_t)n * verts; k++)
    if (elems[k] < 0 || elems[k] >= shape->num_positions)
      return set_error(VPT_ERR_INVALID, "vertex index out of range");
  // one arena: positions | elems | boxes | centroids | keys x2 | ids x2 | children | ranges |
  // parents x2 | axis | node boxes | leaf boxes | arrived | alive | rank | out nodes | bounds
  struct arena_t {
    char*  base = nullptr;
    size_t used = 0;
    void*  take(size_t bytes) {
      used = (used + 255) / 256 * 256;
      auto p = base ? base + used : nullptr;
      used += bytes;
      return p;
    }
  } arena;
  size_t sort_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (unsigned long long*)nullptr,
      (unsigned long long*)nullptr, (int*)nullptr, (int*)nullptr, n);
  size_t scan_bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int*)nullptr, (int*)nullptr, n);
  float *d_pos, *d_cent;
  int *  d_elems, *d_ids, *d_sorted, *d_pint, *d_pleaf, *d_arrived, *d_alive, *d_rank;
  unsigned long long *d_keys, *d_keys2;
  int2 *   d_children, *d_ranges;
  lbvh_box *d_boxes, *d_nboxes, *d_lboxes;
  signed char*  d_axis;
  vpt_bvh_node* d_out;
  unsigned*     d_bounds;
  void *        d_sort, *d_scan;
  auto layout = [&]() {
    arena.used = 0;
    d_pos      = (float*)arena.take(sizeof(float) * 3 * (size_t)shape->num_positions);
    d_elems    = (int*)arena.take(sizeof(int) * (size_t)n * verts);
    d_boxes    = (lbvh_box*)arena.take(sizeof(lbvh_box) * (size_t)n);
    d_cent     = (float*)arena.take(sizeof(float) * 3 * (size_t)n);
    d_keys     = (unsigned long long*)arena.take(8 * (size_t)n);
    d_keys2    = (unsigned long long*)arena.take(8 * (size_t)n);
    d_ids      = (int*)arena.take(4 * (size_t)n);
    d_sorted   = (int*)arena.take(4 * (size_t)n);
    d_children = (int2*)arena.take(8 * (size_t)n);
    d_ranges   = (int2*)arena.take(8 * (size_t)n);
    d_pint     = (int*)arena.take(4 * (size_t)n);
    d_pleaf    = (int*)arena.take(4 * (size_t)n);
    d_axis     = (signed char*)arena.take((size_t)n);
    d_nboxes   = (lbvh_box*)arena.take(sizeof(lbvh_box) * (size_t)n);
    d_lboxes   = (lbvh_box*)arena.take(sizeof(lbvh_box) * (size_t)n);
    d_arrived  = (int*)arena.take(4 * (size_t)n);
    d_alive    = (int*)arena.take(4 * (size_t)n);
    d_rank     = (int*)arena.take(4 * (size_t)n);
    d_out      = (vpt_bvh_node*)arena.take(sizeof(vpt_bvh_node) * (2 * (size_t)n + 1));
    d_bounds   = (unsigned*)arena.take(6 * sizeof(unsigned));
    d_sort     = arena.take(sort_bytes);
    d_scan     = arena.take(scan_bytes);
  };
  layout();
  size_t total = arena.used + 256;
  VPT_CUDA_TRY(cudaMalloc(&arena.base, total));
  layout();
  cudaEvent_t ev[4];
  for (auto& e : ev) cudaEventCreate(&e);
  auto done = [&](int code) {
    for (auto& e : ev) cudaEventDestroy(e);
    cudaFree(arena.base);
    return code;
  };
#define VPT_LBVH_TRY(call)                                  \
  do {                                                      \
    cudaError_t e_ = (call);                                \
    if (e_ != cudaSuccess) return done(cuda_fail(e_, #call)); \
  } while (0)
  int grid = (n + lbvh_block - 1) / lbvh_block;
  VPT_LBVH_TRY(cudaMemcpy(d_pos, shape->positions, sizeof(float) * 3 * (size_t)shape->num_positions,
      cudaMemcpyHostToDevice));
  VPT_LBVH_TRY(cudaMemcpy(d_elems, elems, sizeof(int) * (size_t)n * verts, cudaMemcpyHostToDevice));
  unsigned init_bounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0, 0, 0};
  VPT_LBVH_TRY(cudaMemcpy(d_bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice));
  VPT_LBVH_TRY(cudaMemset(d_arrived, 0, 4 * (size_t)n));
  VPT_LBVH_TRY(cudaMemset(d_pint, 0xff, 4 * (size_t)n));
  VPT_LBVH_TRY(cudaMemset(d_pleaf, 0xff, 4 * (size_t)n));
  cudaEventRecord(ev[0]);
  k_lbvh_boxes<<<grid, lbvh_block>>>(d_pos, d_elems, verts, n, d_boxes, d_cent, d_bounds);
  k_lbvh_keys<<<grid, lbvh_block>>>(d_cent, d_bounds, n, d_keys, d_ids);
  cudaEventRecord(ev[1]);
  VPT_LBVH_TRY(cub::DeviceRadixSort::SortPairs(d_sort, sort_bytes, d_keys, d_keys2, d_ids, d_sorted, n));
  cudaEventRecord(ev[2]);
  int alive_total = 0;
  if (n > 1) {
    k_lbvh_hierarchy<<<grid, lbvh_block>>>(d_keys2, n, d_children, d_ranges, d_pint, d_pleaf, d_axis);
    k_lbvh_fit<<<grid, lbvh_block>>>(d_boxes, d_sorted, n, d_children, d_pint, d_pleaf, d_nboxes,
        d_lboxes, d_arrived);
    k_lbvh_alive<<<grid, lbvh_block>>>(d_ranges, n, d_alive);
    VPT_LBVH_TRY(cub::DeviceScan::ExclusiveSum(d_scan, scan_bytes, d_alive, d_rank, n - 1));
    k_lbvh_emit<<<grid, lbvh_block>>>(d_children, d_ranges, d_axis, d_alive, d_rank, d_nboxes,
        d_lboxes, n, d_out);
    int last[2] = {0, 0};
    VPT_LBVH_TRY(cudaMemcpy(&last[0], d_rank + (n - 2), 4, cudaMemcpyDeviceToHost));
    VPT_LBVH_TRY(cudaMemcpy(&last[1], d_alive + (n - 2), 4, cudaMemcpyDeviceToHost));
    alive_total = last[0] + last[1];
  }
  cudaEventRecord(ev[3]);
  VPT_LBVH_TRY(cudaEventSynchronize(ev[3]));
  VPT_LBVH_TRY(cudaGetLastError());
  VPT_LBVH_TRY(cudaMemcpy(primitives, d_sorted, 4 * (size_t)n, cudaMemcpyDeviceToHost));
  int count = 1 + 2 * alive_total;
  if ((size_t)count > node_capacity) return done(set_error(VPT_ERR_INVALID, "node buffer too small"));
  if (alive_total == 0) {
    // <= 4 primitives: one leaf over all of them (its box is the union of the primitive boxes)
    auto boxes = std::vector<lbvh_box>(n);
    VPT_LBVH_TRY(cudaMemcpy(boxes.data(), d_boxes, sizeof(lbvh_box) * (size_t)n, cudaMemcpyDeviceToHost));
    vpt_bvh_node root{};
    for (int k = 0; k < 3; k++) root.bbox_min[k] = flt_max, root.bbox_max[k] = -flt_max;
    for (auto& b : boxes)
      for (int k = 0; k < 3; k++) {
        root.bbox_min[k] = b.lo[k] < root.bbox_min[k] ? b.lo[k] : root.bbox_min[k];
        root.bbox_max[k] = b.hi[k] > root.bbox_max[k] ? b.hi[k] : root.bbox_max[k];
      }
    root.start = 0, root.num = (int16_t)n, root.axis = 0, root.internal = 0;
    nodes[0] = root;
  } else {
    VPT_LBVH_TRY(cudaMemcpy(nodes, d_out, sizeof(vpt_bvh_node) * (size_t)count, cudaMemcpyDeviceToHost));
  }
#undef VPT_LBVH_TRY
  *num_nodes = count;
  if (stats) {
    stats->primitives = n, stats->nodes = count;
    cudaEventElapsedTime(&stats->keys_ms, ev[0], ev[1]);
    cudaEventElapsedTime(&stats->sort_ms, ev[1], ev[2]);
    cudaEventElapsedTime(&stats->hierarchy_ms, ev[2], ev[3]);
  }
  return done(VPT_OK);
}

// ---- Catmull-Clark subdivision, vertex normals, displacement (tesselate.cuh) ----
}  // extern "C"
namespace {
template <typename T>
struct dev_buf {  // scene-preparation code: plain cudaMalloc'ed scratch, freed on scope exit
  T*     p = nullptr;
  size_t n = 0;
  dev_buf() = default;
  explicit dev_buf(size_t count) { alloc(count); }
  dev_buf(const dev_buf&) = delete;
  dev_buf& operator=(const dev_buf&) = delete;
  cudaError_t alloc(size_t count) {
    release();
    n = count;
    return cudaMalloc(&p, sizeof(T) * (count ? count : 1));
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
  }
  void swap(dev_buf& o) { std::swap(p, o.p), std::swap(n, o.n); }
  ~dev_buf() { release(); }
};
inline int tess_grid(size_t n) { return (int)((n + tess_block - 1) / tess_block); }

template <typename K, typename V>
cudaError_t sort_pairs(const K* kin, K* kout, const V* vin, V* vout, int n, int end_bit) {
  size_t bytes = 0;
  auto   e = cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, n, 0, end_bit);
  if (e != cudaSuccess) return e;
  dev_buf<char> tmp;
  e = tmp.alloc(bytes);
  if (e != cudaSuccess) return e;
  return cub::DeviceRadixSort::SortPairs(tmp.p, bytes, kin, kout, vin, vout, n, 0, end_bit);
}
cudaError_t exclusive_sum(const int* in, int* out, int n) {
  size_t bytes = 0;
  auto   e = cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, n);
  if (e != cudaSuccess) return e;
  dev_buf<char> tmp;
  e = tmp.alloc(bytes);
  if (e != cudaSuccess) return e;
  return cub::DeviceScan::ExclusiveSum(tmp.p, bytes, in, out, n);
}
int bits_for(unsigned max_value) {
  int b = 1;
  while (b < 32 && (max_value >> b)) b++;
  return b;
}

// one level of tesselate_catmullclark on device buffers: (quads, verts) -> (tquads, avert)
template <int D>
cudaError_t catmullclark_level(dev_buf<int4>& quads, dev_buf<float>& verts, int& nf, int& nv,
    bool lock_boundary) {
#define TRY(x)                       \
  do {                               \
    cudaError_t e_ = (x);            \
    if (e_ != cudaSuccess) return e_; \
  } while (0)
  int ns = 4 * nf;
  dev_buf<unsigned long long> keys, keys_sorted;
  dev_buf<int> pos, pos_sorted, face_count, face_offset, head, run;
  TRY(keys.alloc(ns));
  TRY(keys_sorted.alloc(ns));
  TRY(pos.alloc(ns));
  TRY(pos_sorted.alloc(ns));
  TRY(face_count.alloc(nf + 1));
  TRY(face_offset.alloc(nf + 1));
  TRY(head.alloc(ns + 1));
  TRY(run.alloc(ns + 1));
  TRY(cudaMemset(face_count.p, 0, sizeof(int) * (nf + 1)));
  TRY(cudaMemset(head.p, 0, sizeof(int) * (ns + 1)));
  k_tess_sides<<<tess_grid(ns), tess_block>>>(quads.p, nf, keys.p, pos.p, face_count.p);
  TRY((sort_pairs(keys.p, keys_sorted.p, pos.p, pos_sorted.p, ns, 64)));
  k_tess_heads<<<tess_grid(ns), tess_block>>>(keys_sorted.p, ns, head.p);
  TRY(exclusive_sum(head.p, run.p, ns + 1));       // run[ns] = number of unique edges
  TRY(exclusive_sum(face_count.p, face_offset.p, nf + 1));  // face_offset[nf] = new quads
  int ne = 0, nq = 0;
  TRY(cudaMemcpy(&ne, run.p + ns, sizeof(int), cudaMemcpyDeviceToHost));
  TRY(cudaMemcpy(&nq, face_offset.p + nf, sizeof(int), cudaMemcpyDeviceToHost));
  dev_buf<int> first_pos, uid, run_start, first_sorted, order, nfaces, index_of_unique, side_edge;
  dev_buf<int2> edges;
  TRY(first_pos.alloc(ne));
  TRY(uid.alloc(ne));
  TRY(run_start.alloc(ne));
  TRY(first_sorted.alloc(ne));
  TRY(order.alloc(ne));
  TRY(nfaces.alloc(ne));
  TRY(index_of_unique.alloc(ne));
  TRY(side_edge.alloc(ns));
  TRY(edges.alloc(ne));
  k_tess_uniques<<<tess_grid(ns), tess_block>>>(keys_sorted.p, pos_sorted.p, head.p, run.p, ns,
      first_pos.p, uid.p, run_start.p);
  // edges numbered in the order the reference's insert_edge first meets them
  TRY((sort_pairs((const unsigned*)first_pos.p, (unsigned*)first_sorted.p, uid.p, order.p, ne,
      bits_for((unsigned)ns))));
  k_tess_edges<<<tess_grid(ne), tess_block>>>(keys_sorted.p, order.p, run_start.p, ne, ns, edges.p,
      nfaces.p, index_of_unique.p);
  k_tess_side_edges<<<tess_grid(ns), tess_block>>>(keys_sorted.p, pos_sorted.p, head.p, run.p,
      index_of_unique.p, ns, side_edge.p);
  int ntv = nv + ne + nf;
  dev_buf<float> tverts, avert;
  dev_buf<int4>  tquads;
  dev_buf<signed char> val;
  dev_buf<int>   acount;
  TRY(tverts.alloc((size_t)D * ntv));
  TRY(avert.alloc((size_t)D * ntv));
  TRY(tquads.alloc(nq));
  TRY(val.alloc(ntv));
  TRY(acount.alloc(ntv));
  k_tess_vertices<D><<<tess_grid(ntv), tess_block>>>(verts.p, nv, edges.p, ne, quads.p, nf, tverts.p);
  k_tess_faces<<<tess_grid(nf), tess_block>>>(quads.p, nf, side_edge.p, face_offset.p, nv, ne, tquads.p);
  TRY(cudaMemset(val.p, 2, ntv));
  TRY(cudaMemset(avert.p, 0, sizeof(float) * D * ntv));
  TRY(cudaMemset(acount.p, 0, sizeof(int) * ntv));
  k_tess_boundary<D><<<tess_grid(ne), tess_block>>>(edges.p, nfaces.p, ne, nv, lock_boundary ? 1 : 0,
      tverts.p, val.p, avert.p, acount.p, 0);
  k_tess_boundary<D><<<tess_grid(ne), tess_block>>>(edges.p, nfaces.p, ne, nv, lock_boundary ? 1 : 0,
      tverts.p, val.p, avert.p, acount.p, 1);
  int nentries = 4 * nq;
  dev_buf<unsigned> vkey, vkey_sorted;
  dev_buf<int>      entry, entry_sorted;
  TRY(vkey.alloc(nentries));
  TRY(vkey_sorted.alloc(nentries));
  TRY(entry.alloc(nentries));
  TRY(entry_sorted.alloc(nentries));
  k_tess_corners<<<tess_grid(nentries), tess_block>>>(tquads.p, nq, vkey.p, entry.p);
  TRY((sort_pairs(vkey.p, vkey_sorted.p, entry.p, entry_sorted.p, nentries, bits_for((unsigned)ntv))));
  k_tess_average<D><<<tess_grid(ntv), tess_block>>>(tquads.p, vkey_sorted.p, entry_sorted.p, nentries,
      tverts.p, ntv, val.p, avert.p, acount.p);
  TRY(cudaGetLastError());
  TRY(cudaDeviceSynchronize());
  quads.swap(tquads), verts.swap(avert);
  nf = nq, nv = ntv;
  return cudaSuccess;
#undef TRY
}
}  // namespace
extern "C" {

void vpt_cuda_free_host(void* p) { free(p); }

int vpt_cuda_catmullclark(const int32_t* quads, int nquads, const float* verts, int nverts, int dim,
    int lock_boundary, int levels, int32_t** out_quads, int32_t* out_nquads, float** out_verts,
    int32_t* out_nverts) {
  if (!out_quads || !out_nquads || !out_verts || !out_nverts || nquads < 0 || nverts < 0 ||
      (dim != 2 && dim != 3) || levels < 0 || (nquads > 0 && !quads) || (nverts > 0 && !verts))
    return set_error(VPT_ERR_INVALID, "bad argument");
  *out_quads = nullptr, *out_verts = nullptr, *out_nquads = 0, *out_nverts = 0;
  if (auto r = require_device()) return r;
  for (size_t k = 0; k < (size_t)nquads * 4; k++)
    if (quads[k] < 0 || quads[k] >= nverts) return set_error(VPT_ERR_INVALID, "vertex index out of range");
  dev_buf<int4>  dq;
  dev_buf<float> dv;
  VPT_CUDA_TRY(dq.alloc(nquads));
  VPT_CUDA_TRY(dv.alloc((size_t)dim * nverts));
  VPT_CUDA_TRY(cudaMemcpy(dq.p, quads, sizeof(int4) * (size_t)nquads, cudaMemcpyHostToDevice));
  VPT_CUDA_TRY(cudaMemcpy(dv.p, verts, sizeof(float) * dim * (size_t)nverts, cudaMemcpyHostToDevice));
  int nf = nquads, nv = nverts;
  for (int level = 0; level < levels && nf > 0; level++) {
    if ((long long)nf * 16 > 0x7fffffffLL) return set_error(VPT_ERR_UNSUPPORTED, "mesh too large");
    auto e = dim == 3 ? catmullclark_level<3>(dq, dv, nf, nv, lock_boundary != 0)
                      : catmullclark_level<2>(dq, dv, nf, nv, lock_boundary != 0);
    if (e != cudaSuccess) return cuda_fail(e, "catmullclark level");
  }
  auto hq = (int32_t*)malloc(sizeof(int4) * (size_t)(nf ? nf : 1));
  auto hv = (float*)malloc(sizeof(float) * dim * (size_t)(nv ? nv : 1));
  if (!hq || !hv) {
    free(hq), free(hv);
    return set_error(VPT_ERR_OUT_OF_MEMORY, "host allocation failed");
  }
  auto e = cudaMemcpy(hq, dq.p, sizeof(int4) * (size_t)nf, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(hv, dv.p, sizeof(float) * dim * (size_t)nv, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    free(hq), free(hv);
    return cuda_fail(e, "catmullclark download");
  }
  *out_quads = hq, *out_nquads = nf, *out_verts = hv, *out_nverts = nv;
  return VPT_OK;
}

int vpt_cuda_vertex_normals(const int32_t* elems, int nelems, int verts_per, const float* positions,
    int nverts, float* normals) {
  if (!normals || nelems < 0 || nverts < 0 || (verts_per != 3 && verts_per != 4) ||
      (nelems > 0 && !elems) || (nverts > 0 && !positions))
    return set_error(VPT_ERR_INVALID, "bad argument");
  if (auto r = require_device()) return r;
  if (nverts == 0) return VPT_OK;
  int n = verts_per * nelems;
  for (int k = 0; k < n; k++)
    if (elems[k] < 0 || elems[k] >= nverts) return set_error(VPT_ERR_INVALID, "vertex index out of range");
  dev_buf<int> de, entry, entry_sorted;
  dev_buf<unsigned> vkey, vkey_sorted;
  dev_buf<float> dp, dn;
  VPT_CUDA_TRY(de.alloc(n));
  VPT_CUDA_TRY(entry.alloc(n));
  VPT_CUDA_TRY(entry_sorted.alloc(n));
  VPT_CUDA_TRY(vkey.alloc(n));
  VPT_CUDA_TRY(vkey_sorted.alloc(n));
  VPT_CUDA_TRY(dp.alloc(3 * (size_t)nverts));
  VPT_CUDA_TRY(dn.alloc(3 * (size_t)nverts));
  VPT_CUDA_TRY(cudaMemcpy(de.p, elems, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
  VPT_CUDA_TRY(cudaMemcpy(dp.p, positions, sizeof(float) * 3 * (size_t)nverts, cudaMemcpyHostToDevice));
  if (n > 0) {
    k_tess_elem_corners<<<tess_grid(n), tess_block>>>(de.p, nelems, verts_per, vkey.p, entry.p);
    VPT_CUDA_TRY((sort_pairs(vkey.p, vkey_sorted.p, entry.p, entry_sorted.p, n, 32)));
  }
  k_tess_normals<<<tess_grid(nverts), tess_block>>>(de.p, verts_per, vkey_sorted.p, entry_sorted.p, n,
      dp.p, nverts, dn.p);
  VPT_CUDA_TRY(cudaGetLastError());
  VPT_CUDA_TRY(cudaMemcpy(normals, dn.p, sizeof(float) * 3 * (size_t)nverts, cudaMemcpyDeviceToHost));
  return VPT_OK;
}

int vpt_cuda_displace(float* positions, const float* normals, const float* texcoords, int nverts,
    const vpt_texture_desc* texture, float displacement) {
  if (!positions || !normals || !texcoords || !texture || nverts < 0)
    return set_error(VPT_ERR_INVALID, "bad argument");
  if (auto r = require_device()) return r;
  if (nverts == 0) return VPT_OK;
  size_t texels = (size_t)texture->width * texture->height;
  if (texture->width < 0 || texture->height < 0 || (texels > 0 && !texture->pixelsf && !texture->pixelsb))
    return set_error(VPT_ERR_INVALID, "bad texture");
  // a one-texture scene view: record 0, its texel pool, the host-built sRGB table
  dev_buf<float> dp, dn, dt, lut;
  dev_buf<f4>    texf;
  dev_buf<unsigned char> texb;
  dev_buf<i4>    rec;
  VPT_CUDA_TRY(dp.alloc(3 * (size_t)nverts));
  VPT_CUDA_TRY(dn.alloc(3 * (size_t)nverts));
  VPT_CUDA_TRY(dt.alloc(2 * (size_t)nverts));
  VPT_CUDA_TRY(lut.alloc(256));
  VPT_CUDA_TRY(rec.alloc(1));
  VPT_CUDA_TRY(texf.alloc(texture->pixelsf ? texels : 0));
  VPT_CUDA_TRY(texb.alloc(texture->pixelsf ? 0 : 4 * texels));
  float table[256];
  for (int b = 0; b < 256; b++) {
    float srgb = b / 255.0f;
    table[b]   = ((double)srgb <= 0.04045) ? srgb / 12.92f : powf((srgb + 0.055f) / (1.0f + 0.055f), 2.4f);
  }
  i4 record = {texture->width, texture->height, (texture->linear ? 1 : 0) | (texture->pixelsf ? 2 : 0), 0};
  VPT_CUDA_TRY(cudaMemcpy(rec.p, &record, sizeof(record), cudaMemcpyHostToDevice));
  VPT_CUDA_TRY(cudaMemcpy(lut.p, table, sizeof(table), cudaMemcpyHostToDevice));
  if (texture->pixelsf)
    VPT_CUDA_TRY(cudaMemcpy(texf.p, texture->pixelsf, sizeof(f4) * texels, cudaMemcpyHostToDevice));
  else if (texels)
    VPT_CUDA_TRY(cudaMemcpy(texb.p, texture->pixelsb, 4 * texels, cudaMemcpyHostToDevice));
  VPT_CUDA_TRY(cudaMemcpy(dp.p, positions, sizeof(float) * 3 * (size_t)nverts, cudaMemcpyHostToDevice));
  VPT_CUDA_TRY(cudaMemcpy(dn.p, normals, sizeof(float) * 3 * (size_t)nverts, cudaMemcpyHostToDevice));
  VPT_CUDA_TRY(cudaMemcpy(dt.p, texcoords, sizeof(float) * 2 * (size_t)nverts, cudaMemcpyHostToDevice));
  DScene sc{};
  sc.textures = rec.p, sc.texf = texf.p, sc.texb = texb.p, sc.srgb_lut = lut.p;
  k_tess_displace<<<tess_grid(nverts), tess_block>>>(sc, texture->pixelsf ? 0 : 1, displacement, dn.p,
      dt.p, nverts, dp.p);
  VPT_CUDA_TRY(cudaGetLastError());
  VPT_CUDA_TRY(cudaMemcpy(positions, dp.p, sizeof(float) * 3 * (size_t)nverts, cudaMemcpyDeviceToHost));
  return VPT_OK;
}

int vpt_cuda_scene_num_lights(const vpt_scene* scene) { return scene ? scene->flat.d.num_lights : 0; }

int vpt_cuda_download_light(const vpt_scene* scene, int light, int32_t ids[3], float* cdf,
    size_t capacity, int32_t* num_elements) {
  if (!scene || !ids || !num_elements || light < 0 || light >= scene->flat.d.num_lights)
    return set_error(VPT_ERR_INVALID, "bad light index");
  i4 rec[2];
  VPT_CUDA_TRY(cudaMemcpy(rec, scene->flat.d.lights + 2 * light, sizeof(rec), cudaMemcpyDeviceToHost));
  ids[0] = rec[0].x, ids[1] = rec[0].y, ids[2] = rec[0].z;
  *num_elements = rec[1].x;
  if (cdf && rec[1].x > 0) {
    if (capacity < (size_t)rec[1].x) return set_error(VPT_ERR_INVALID, "cdf buffer too small");
    VPT_CUDA_TRY(cudaMemcpy(cdf, scene->flat.d.cdf + rec[0].w, sizeof(float) * rec[1].x,
        cudaMemcpyDeviceToHost));
  }
  return VPT_OK;
}

void vpt_cuda_free_scene(vpt_scene* scene) {
  if (!scene) return;
  scene->flat.pool.release();
  delete scene;
}

size_t vpt_cuda_scene_bytes(const vpt_scene* scene) { return scene ? scene->flat.pool.bytes : 0; }

int vpt_cuda_create_state(int width, int height, const uint64_t* rngs, vpt_state** out) {
  if (!out || width <= 0 || height <= 0) return set_error(VPT_ERR_INVALID, "bad state size");
  *out = nullptr;
  if (auto r = require_device()) return r;
  auto state = new (std::nothrow) vpt_state();
  if (!state) return set_error(VPT_ERR_OUT_OF_MEMORY, "host allocation failed");
  state->width = width, state->height = height;
  cudaGetDevice(&state->device);
  size_t n = (size_t)width * height;
  auto   e = cudaMalloc(&state->image, n * sizeof(f4));
  if (e == cudaSuccess) e = cudaMalloc(&state->rngs, n * sizeof(rng_t));
  if (e == cudaSuccess) e = cudaEventCreate(&state->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&state->ev1);
  if (e != cudaSuccess) {
    vpt_cuda_free_state(state);
    return cuda_fail(e, "state allocation");
  }
  auto r = vpt_cuda_reset_state(state, rngs);
  if (r != VPT_OK) {
    vpt_cuda_free_state(state);
    return r;
  }
  *out = state;
  return VPT_OK;
}

int vpt_cuda_reset_state(vpt_state* state, const uint64_t* rngs) {
  if (!state) return set_error(VPT_ERR_INVALID, "null state");
  size_t n = (size_t)state->width * state->height;
  VPT_CUDA_TRY(cudaMemsetAsync(state->image, 0, n * sizeof(f4)));
  if (rngs) {
    VPT_CUDA_TRY(cudaMemcpyAsync(state->rngs, rngs, n * sizeof(rng_t), cudaMemcpyHostToDevice));
  } else {
    VPT_CUDA_TRY(cudaMemsetAsync(state->rngs, 0, n * sizeof(rng_t)));
  }
  state->samples = 0;
  state->wf.live = false;
  return VPT_OK;
}

int vpt_cuda_advance_rngs(vpt_state* state, uint64_t delta) {
  if (!state) return set_error(VPT_ERR_INVALID, "null state");
  int n = state->width * state->height;
  k_advance_rngs<<<(n + 255) / 256, 256>>>(state->rngs, n, delta);
  VPT_CUDA_TRY(cudaGetLastError());
  return VPT_OK;
}

void vpt_cuda_free_state(vpt_state* state) {
  if (!state) return;
  cudaFree(state->image);
  cudaFree(state->rngs);
  if (state->reduced) cudaFree(state->reduced);
  if (state->staging) cudaFree(state->staging);
  if (state->out8) cudaFree(state->out8);
  if (state->pinned8) cudaFreeHost(state->pinned8);
  state->wf.release();
  if (state->ev0) cudaEventDestroy(state->ev0);
  if (state->ev1) cudaEventDestroy(state->ev1);
  delete state;
}

int vpt_cuda_state_samples(const vpt_state* state) { return state ? state->samples : 0; }

int vpt_cuda_trace_samples(vpt_scene* scene, vpt_state* state, const vpt_params* params,
    int nsamples) {
  if (!scene || !state || !params) return set_error(VPT_ERR_INVALID, "null argument");
  if (params->shader < 0 || params->shader >= VPT_SHADER_COUNT)
    return set_error(VPT_ERR_INVALID, "sampler unknown");  // get_shader, yocto_pathtrace.cpp:948
  if (params->camera < 0 || params->camera >= scene->flat.num_cameras)
    return set_error(VPT_ERR_INVALID, "camera index out of range");
  // the wavefront packs the bounce counter in 16 bits (the CLI's range is [1, 128])
  if (params->bounces < 1 || params->bounces > 65535)
    return set_error(VPT_ERR_INVALID, "bounces out of range [1, 65535]");
  state->stats = vpt_trace_stats{};
  int todo = std::min(nsamples, params->samples - state->samples);
  if (todo <= 0) return VPT_OK;  // yocto_pathtrace.cpp:1055

  DParams prm{};
  prm.camera              = params->camera;
  prm.shader              = params->shader;
  prm.samples             = params->samples;
  prm.bounces             = params->bounces;
  prm.noimplicit_mis      = params->noimplicit_mis;
  prm.spheretrace_maxiter = params->spheretrace_maxiter;
  prm.width = state->width, prm.height = state->height;

  bool path_shader = prm.shader == shader_volpathtrace || prm.shader == shader_pathtrace ||
                     prm.shader == shader_naive || prm.shader == shader_implicit;
  // auto: the four Monte-Carlo shaders run as a wavefront (persistent extend with lane refill
  // measures faster than the megakernel for BVH traversal and for sphere tracing alike)
  bool wavefront = params->pipeline == VPT_PIPELINE_WAVEFRONT ||
                   (params->pipeline == VPT_PIPELINE_AUTO && path_shader);

  VPT_CUDA_TRY(cudaEventRecord(state->ev0));
  int launches = 0;
  if (wavefront && path_shader) {
    auto e = run_wavefront(scene->flat, state->wf, prm, state->image, state->rngs, todo,
        &launches, &state->stats);
    if (e != cudaSuccess) return cuda_fail(e, "wavefront pipeline");
  } else {
    auto e = dispatch_mega(scene, state, prm, todo);
    if (e != cudaSuccess) return cuda_fail(e, "megakernel launch");
    launches = 1;
  }
  VPT_CUDA_TRY(cudaEventRecord(state->ev1));
  VPT_CUDA_TRY(cudaEventSynchronize(state->ev1));
  VPT_CUDA_TRY(cudaGetLastError());
  float ms = 0;
  VPT_CUDA_TRY(cudaEventElapsedTime(&ms, state->ev0, state->ev1));
  state->samples += todo;
  state->stats.device_ms       = ms;
  state->stats.kernel_launches = launches;
  state->stats.paths           = (uint64_t)todo * state->width * state->height;
  if (state->stats.top_kernel_launches == 0) {  // megakernel: the call is the kernel
    state->stats.top_kernel_ms       = ms;
    state->stats.top_kernel_launches = 1;
    state->stats.top_kernel_units    = state->stats.paths;
  }
  return VPT_OK;
}

int vpt_cuda_get_trace_stats(const vpt_state* state, vpt_trace_stats* stats) {
  if (!state || !stats) return set_error(VPT_ERR_INVALID, "null argument");
  *stats = state->stats;
  return VPT_OK;
}

// Sum of the root's and its peers' accumulation images on the root's device (SURVEY 8(b)/(e):
// "vpt_cuda_download_image performs the reduce if multi-GPU"). One kernel reads every peer's
// image through P2P loads over NVLink (k_reduce_images); a peer the root cannot map is copied
// over first. The root's own image is left alone, so downloads stay idempotent. Returns the
// image to read (state->image when there are no peers).
static int reduced_image(vpt_state* state, const f4** out) {
  *out = state->image;
  if (state->peers.empty()) return VPT_OK;
  size_t pixels = (size_t)state->width * state->height;
  int    current = 0;
  VPT_CUDA_TRY(cudaGetDevice(&current));
  if (current != state->device) VPT_CUDA_TRY(cudaSetDevice(state->device));
  if (!state->reduced) VPT_CUDA_TRY(cudaMalloc(&state->reduced, pixels * sizeof(f4)));
  VPT_CUDA_TRY(cudaMemcpy(state->reduced, state->image, pixels * sizeof(f4), cudaMemcpyDeviceToDevice));
  peer_list direct{};
  int       ndirect = 0;
  auto flush = [&]() -> cudaError_t {
    if (!ndirect) return cudaSuccess;
    k_reduce_images<<<(unsigned)((pixels + 255) / 256), 256>>>((float*)state->reduced, direct,
        ndirect, pixels);
    ndirect = 0;
    return cudaGetLastError();
  };
  for (auto peer : state->peers) {
    bool mapped = peer->device == state->device;
    if (!mapped) {
      int can = 0;
      cudaDeviceCanAccessPeer(&can, state->device, peer->device);
      if (can) {
        auto e = cudaDeviceEnablePeerAccess(peer->device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError(), e = cudaSuccess;
        mapped = e == cudaSuccess;
        if (!mapped) cudaGetLastError();
      }
    }
    if (mapped) {
      direct.p[ndirect++] = (const float*)peer->image;
      if (ndirect == 16) VPT_CUDA_TRY(flush());
    } else {
      VPT_CUDA_TRY(flush());
      if (!state->staging) VPT_CUDA_TRY(cudaMalloc(&state->staging, pixels * sizeof(f4)));
      VPT_CUDA_TRY(cudaMemcpyPeer(state->staging, state->device, peer->image, peer->device,
          pixels * sizeof(f4)));
      direct.p[ndirect++] = (const float*)state->staging;
      VPT_CUDA_TRY(flush());
    }
  }
  VPT_CUDA_TRY(flush());
  VPT_CUDA_TRY(cudaDeviceSynchronize());
  if (current != state->device) VPT_CUDA_TRY(cudaSetDevice(current));
  *out = state->reduced;
  return VPT_OK;
}

int vpt_cuda_state_add_peer(vpt_state* root, vpt_state* peer) {
  if (!root || !peer || root == peer) return set_error(VPT_ERR_INVALID, "bad argument");
  if (root->width != peer->width || root->height != peer->height)
    return set_error(VPT_ERR_INVALID, "image should have the same size");
  if (!peer->peers.empty()) return set_error(VPT_ERR_INVALID, "a peer cannot have peers of its own");
  for (auto p : root->peers)
    if (p == peer) return VPT_OK;
  root->peers.push_back(peer);
  return VPT_OK;
}

int vpt_cuda_state_remove_peers(vpt_state* root) {
  if (!root) return set_error(VPT_ERR_INVALID, "null state");
  root->peers.clear();
  return VPT_OK;
}

int vpt_cuda_state_total_samples(const vpt_state* state) {
  if (!state) return 0;
  int total = state->samples;
  for (auto p : state->peers) total += p->samples;
  return total;
}

int vpt_cuda_download_image(vpt_state* state, float* rgba, size_t n) {
  if (!state || !rgba) return set_error(VPT_ERR_INVALID, "null argument");
  size_t want = (size_t)state->width * state->height * 4;
  if (n != want) return set_error(VPT_ERR_INVALID, "image should have the same size");
  const f4* image = nullptr;
  if (auto r = reduced_image(state, &image)) return r;
  VPT_CUDA_TRY(cudaMemcpy(rgba, image, want * sizeof(float), cudaMemcpyDeviceToHost));
  return VPT_OK;
}

// exposure + filmic tonemap of the reference's display path (tonemap, yocto_color.h / the
// viewer's params.exposure / params.filmic, ypathtrace.cpp:159-189): rgb * 2^exposure, optional
// ACES-fitted filmic curve, then rgb_to_srgb + float_to_byte as save_image does
int vpt_cuda_download_tonemapped8(vpt_state* state, int samples, float exposure, int filmic,
    uint8_t* rgba, size_t n);

int vpt_cuda_download_srgb8(vpt_state* state, int samples, uint8_t* rgba, size_t n) {
  return vpt_cuda_download_tonemapped8(state, samples, 0.0f, 0, rgba, n);
}

int vpt_cuda_download_tonemapped8(vpt_state* state, int samples, float exposure, int filmic,
    uint8_t* rgba, size_t n) {
  if (auto r = require_device()) return r;
  if (!state || !rgba) return set_error(VPT_ERR_INVALID, "null argument");
  size_t pixels = (size_t)state->width * state->height;
  if (n != pixels * 4) return set_error(VPT_ERR_INVALID, "image should have the same size");
  if (samples <= 0) samples = vpt_cuda_state_total_samples(state);
  if (samples <= 0) return set_error(VPT_ERR_INVALID, "nothing rendered yet");
  const f4* image = nullptr;
  if (auto r = reduced_image(state, &image)) return r;
  // device and pinned host buffers are kept with the state: a preview loop calls this per frame
  if (!state->out8) VPT_CUDA_TRY(cudaMalloc(&state->out8, pixels * sizeof(uchar4)));
  if (!state->pinned8) VPT_CUDA_TRY(cudaMallocHost(&state->pinned8, pixels * sizeof(uchar4)));
  k_output_srgb8<<<(unsigned)((pixels + 255) / 256), 256>>>(image, state->out8, pixels,
      1.0f / (float)samples, exp2f(exposure), exposure != 0 ? 1 : 0, filmic ? 1 : 0);
  VPT_CUDA_TRY(cudaGetLastError());
  VPT_CUDA_TRY(cudaMemcpy(state->pinned8, state->out8, pixels * sizeof(uchar4), cudaMemcpyDeviceToHost));
  memcpy(rgba, state->pinned8, pixels * sizeof(uchar4));
  return VPT_OK;
}

int vpt_cuda_download_rngs(vpt_state* state, uint64_t* rngs, size_t n) {
  if (!state || !rngs) return set_error(VPT_ERR_INVALID, "null argument");
  size_t want = (size_t)state->width * state->height * 2;
  if (n != want) return set_error(VPT_ERR_INVALID, "rng buffer should have the same size");
  VPT_CUDA_TRY(cudaMemcpy(rngs, state->rngs, want * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  return VPT_OK;
}

void* vpt_cuda_image_device_ptr(vpt_state* state) { return state ? state->image : nullptr; }

int vpt_cuda_export_image(vpt_state* state, void* device_dst, size_t n) {
  if (!state || !device_dst) return set_error(VPT_ERR_INVALID, "null argument");
  size_t want = (size_t)state->width * state->height * 4;
  if (n != want) return set_error(VPT_ERR_INVALID, "image should have the same size");
  VPT_CUDA_TRY(cudaMemcpy(device_dst, state->image, want * sizeof(float), cudaMemcpyDeviceToDevice));
  return VPT_OK;
}

int vpt_cuda_import_image(vpt_state* state, const void* device_src, size_t n) {
  if (!state || !device_src) return set_error(VPT_ERR_INVALID, "null argument");
  size_t want = (size_t)state->width * state->height * 4;
  if (n != want) return set_error(VPT_ERR_INVALID, "image should have the same size");
  VPT_CUDA_TRY(cudaMemcpy(state->image, device_src, want * sizeof(float), cudaMemcpyDeviceToDevice));
  return VPT_OK;
}

int vpt_cuda_intersect(vpt_scene* scene, const float* rays, int n, int instance, int32_t* ids,
    float* uvd) {
  if (!scene || !rays || !ids || !uvd || n < 0) return set_error(VPT_ERR_INVALID, "bad argument");
  if (instance >= scene->flat.d.num_instances)
    return set_error(VPT_ERR_INVALID, "instance out of range");
  if (n == 0) return VPT_OK;
  float* d_rays = nullptr;
  int*   d_ids  = nullptr;
  float* d_uvd  = nullptr;
  auto   e      = cudaMalloc(&d_rays, (size_t)n * 8 * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&d_ids, (size_t)n * 3 * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&d_uvd, (size_t)n * 3 * sizeof(float));
  if (e == cudaSuccess)
    e = cudaMemcpy(d_rays, rays, (size_t)n * 8 * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    size_t smem = (size_t)scene->flat.stack_depth * vpt_stack_words * mega_block * sizeof(int);
    if (smem > 48 * 1024)
      e = cudaFuncSetAttribute(k_intersect, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) {
      k_intersect<<<(n + mega_block - 1) / mega_block, mega_block, smem>>>(scene->flat.d, d_rays,
          n, instance, d_ids, d_uvd);
      e = cudaGetLastError();
    }
  }
  if (e == cudaSuccess) e = cudaMemcpy(ids, d_ids, (size_t)n * 3 * sizeof(int), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(uvd, d_uvd, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(d_rays), cudaFree(d_ids), cudaFree(d_uvd);
  if (e != cudaSuccess) return cuda_fail(e, "vpt_cuda_intersect");
  return VPT_OK;
}

}  // extern "C" (the scratch array and its kernel have C++ linkage)

// ---- first-hit queries without any allocation -------------------------------------------------
// vpt_cuda_intersect stages its ray batch and results in three cudaMalloc'ed buffers per call.
// Some callers cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
// that forbid cudaMalloc): this variant stages through ONE file-scope device array instead.
// Every GPU that loads the library holds the array until the process exits.
//   * the array name is a host-side stand-in in host code, NOT a device address: the address
//     comes from cudaGetSymbolAddress, once per device, with that device current;
//   * all calls share the array: the kernel overwrites every record of [0, n) it is asked for
//     and nothing has to start at zero, but two calls that overlap on the GPU would race, so the
//     entry point serialises host threads of this process (it is blocking anyway).
constexpr int      noalloc_max_rays  = VPT_NOALLOC_MAX_RAYS;
constexpr size_t   noalloc_ray_bytes = 8 * sizeof(float), noalloc_ids_bytes = 3 * sizeof(int),
                   noalloc_uvd_bytes = 3 * sizeof(float);
constexpr size_t   noalloc_bytes =
    (size_t)noalloc_max_rays * (noalloc_ray_bytes + noalloc_ids_bytes + noalloc_uvd_bytes);
static __device__ __align__(256) unsigned char g_noalloc_scratch[noalloc_bytes];  // 3.5 MiB

// the scratch variant of k_intersect: {rays | ids | uvd} regions of one array, sized for
// noalloc_max_rays whatever n is, so region offsets never depend on the call
__global__ void __launch_bounds__(mega_block)
    k_intersect_scratch(const DScene sc, unsigned char* __restrict__ scratch, int n, int instance) {
  extern __shared__ int stack_mem[];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  auto rays = reinterpret_cast<const float*>(scratch);
  auto ids  = reinterpret_cast<int*>(scratch + (size_t)noalloc_max_rays * noalloc_ray_bytes);
  auto uvd  = reinterpret_cast<float*>(
      scratch + (size_t)noalloc_max_rays * (noalloc_ray_bytes + noalloc_ids_bytes));
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  auto    r   = rays + 8 * (size_t)i;
  auto    ray = ray3{{r[0], r[1], r[2]}, {r[3], r[4], r[5]}, r[6], r[7]};
  auto    h   = instance >= 0 ? intersect_instance(sc, instance, ray, st)
                              : intersect_scene(sc, ray, st);
  ids[3 * i + 0] = h.hit ? 1 : 0;
  ids[3 * i + 1] = h.hit ? h.instance : -1;
  ids[3 * i + 2] = h.hit ? h.element : -1;
  uvd[3 * i + 0] = h.hit ? h.uv.x : 0;
  uvd[3 * i + 1] = h.hit ? h.uv.y : 0;
  uvd[3 * i + 2] = h.hit ? h.distance : 0;
}

static std::mutex     g_noalloc_mutex;
static unsigned char* g_noalloc_ptr[64] = {};  // device address of the array, per device

extern "C" {

int vpt_cuda_intersect_noalloc(vpt_scene* scene, const float* rays, int n, int instance,
    int32_t* ids, float* uvd) {
  if (!scene || !rays || !ids || !uvd || n < 0) return set_error(VPT_ERR_INVALID, "bad argument");
  if (instance >= scene->flat.d.num_instances)
    return set_error(VPT_ERR_INVALID, "instance out of range");
  if (n > noalloc_max_rays)
    return set_error(VPT_ERR_INVALID, "vpt_cuda_intersect_noalloc: " + std::to_string(n) +
                                          " rays exceed the static scratch of " +
                                          std::to_string(noalloc_max_rays) +
                                          " rays (split the batch or use vpt_cuda_intersect)");
  if (n == 0) return VPT_OK;
  auto lock   = std::lock_guard<std::mutex>(g_noalloc_mutex);
  int  device = 0;
  VPT_CUDA_TRY(cudaGetDevice(&device));
  if (device < 0 || device >= 64) return set_error(VPT_ERR_UNSUPPORTED, "device index above 63");
  if (device != scene->device)
    return set_error(VPT_ERR_INVALID, "the scene lives on another device than the current one");
  if (!g_noalloc_ptr[device]) {
    // set-up, once per device, with that device current
    void* ptr = nullptr;
    VPT_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
    if (!ptr) return set_error(VPT_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[device] = static_cast<unsigned char*>(ptr);
  }
  auto   base  = g_noalloc_ptr[device];
  auto   d_ids = base + (size_t)noalloc_max_rays * noalloc_ray_bytes;
  auto   d_uvd = base + (size_t)noalloc_max_rays * (noalloc_ray_bytes + noalloc_ids_bytes);
  size_t smem  = (size_t)scene->flat.stack_depth * vpt_stack_words * mega_block * sizeof(int);
  if (smem > 48 * 1024)
    VPT_CUDA_TRY(cudaFuncSetAttribute(k_intersect_scratch,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  VPT_CUDA_TRY(cudaMemcpy(base, rays, (size_t)n * noalloc_ray_bytes, cudaMemcpyHostToDevice));
  k_intersect_scratch<<<(n + mega_block - 1) / mega_block, mega_block, smem>>>(scene->flat.d, base,
      n, instance);
  VPT_CUDA_TRY(cudaGetLastError());
  VPT_CUDA_TRY(cudaMemcpy(ids, d_ids, (size_t)n * noalloc_ids_bytes, cudaMemcpyDeviceToHost));
  VPT_CUDA_TRY(cudaMemcpy(uvd, d_uvd, (size_t)n * noalloc_uvd_bytes, cudaMemcpyDeviceToHost));
  return VPT_OK;
}

int vpt_cuda_spheretrace(vpt_scene* scene, const float* rays, int n, int maxiter, int32_t* ids,
    float* dn) {
  if (!scene || !rays || !ids || !dn || n < 0) return set_error(VPT_ERR_INVALID, "bad argument");
  if (n == 0) return VPT_OK;
  float* d_rays = nullptr;
  int*   d_ids  = nullptr;
  float* d_dn   = nullptr;
  auto   e      = cudaMalloc(&d_rays, (size_t)n * 8 * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(&d_ids, (size_t)n * 3 * sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&d_dn, (size_t)n * 4 * sizeof(float));
  if (e == cudaSuccess)
    e = cudaMemcpy(d_rays, rays, (size_t)n * 8 * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    k_spheretrace<<<(n + mega_block - 1) / mega_block, mega_block>>>(scene->flat.d, d_rays, n,
        maxiter, d_ids, d_dn);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(ids, d_ids, (size_t)n * 3 * sizeof(int), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(dn, d_dn, (size_t)n * 4 * sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(d_rays), cudaFree(d_ids), cudaFree(d_dn);
  if (e != cudaSuccess) return cuda_fail(e, "vpt_cuda_spheretrace");
  return VPT_OK;
}

int vpt_cuda_reduce_images(float* dst, const float* const* peers, int npeers, size_t count) {
  if (!dst || (npeers > 0 && !peers) || npeers < 0 || npeers > 16 || (count & 3))
    return set_error(VPT_ERR_INVALID, "bad argument");
  if (npeers == 0 || count == 0) return VPT_OK;
  peer_list pl{};
  for (int k = 0; k < npeers; k++) pl.p[k] = peers[k];
  size_t n4 = count / 4;
  k_reduce_images<<<(unsigned)((n4 + 255) / 256), 256>>>(dst, pl, npeers, n4);
  VPT_CUDA_TRY(cudaGetLastError());
  VPT_CUDA_TRY(cudaDeviceSynchronize());
  return VPT_OK;
}

}  // extern "C"
