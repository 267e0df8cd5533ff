// wavefront.cuh — the path loop as a wavefront pipeline.
//
// One path slot per pixel (slot order = 8x4 pixel tiles). A slot owns its
// pixel's PCG32 stream and runs that pixel's samples one after the other, so
// the random streams replay exactly as in the reference / the megakernel and
// both pipelines produce bit-identical images. Per bounce ("iteration"):
//
//   k_wf_extend    closest hit (BVH) or sphere-trace for every queued slot
//   k_wf_shade     shade_hit: emission, opacity, BSDF / phase / light sampling;
//                  paths that need the MIS light pdf go to the pdf queue, the
//                  rest run bounce_tail and go to the next extend queue
//   k_wf_lightpdf  sample_lights_pdf (instance-targeted traversals, sdf march,
//                  environment texel pdf), weight update, bounce_tail
//
// A path that ends is accumulated into the image by its owner thread and the
// slot regenerates the pixel's next camera sample in place (path regeneration),
// so many samples run per launch without a host round trip. Queues hold slot
// ids and are compacted with warp-aggregated appends: __ballot_sync + __popc
// give each lane its rank, one atomicAdd per warp reserves the range.
// Per-slot state is one contiguous 160-byte record (see wf_view).
//
// Between extend and shade the queue is counting-sorted by shade class (miss,
// inside-a-volume, surface material id): k_wf_extend takes a rank inside its
// class (one atomicAdd per distinct class per warp, __match_any_sync),
// k_wf_sort scatters the slot ids to class_base + rank. Warps of k_wf_shade
// then run one material's code instead of all of them (ncu r01: 7.9 of 32
// lanes and 15 warps stalled on instruction fetch per issue before sorting).
// The order inside a class is arbitrary; pixels do not interact, so results
// are unaffected.
//
// Scheduling (none of it changes a pixel's sequence of operations):
//   * pipes: mesh shaders run two independent wavefronts over interleaved
//     tiles on two streams (wf_pipe), so that the latency-bound kernels of one
//     fill the issue slots the other's draining extend leaves idle;
//   * parked rays: a sphere-tracing warp that only drains hands its unfinished
//     rays to the next iteration's launch (wf_bit_susp), see below;
//   * the host looks at the counters every wf_batch iterations, two batches in
//     flight per pipe, and hands the last slots to k_wf_drain.
#pragma once

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>
#include <cuda_runtime.h>

#include "../../include/vpt_cuda.h"
#include "shaders.cuh"
#include "tiling.cuh"
#include "upload.cuh"

namespace vpt {

constexpr int wf_block   = 128;
// independent wavefronts (streams) per render call: 1 .. 3 by image size, see run_wavefront_shader
constexpr int wf_pipes_min_slots = 1 << 16;  // smaller images run one pipe
constexpr int wf_pipes_two_slots = 300000, wf_pipes_one_slots = 1200000;  // see run_wavefront_shader
constexpr int wf_batch   = 8;  // iterations enqueued between two looks at the counters
constexpr int wf_timed_every = 8;  // with graphs: every 8th batch of pipe 0 carries the timed extend
constexpr int wf_classes = 64;  // 0 miss, 1 in-volume, 2.. material id (clamped), last: suspended
constexpr int wf_class_susp = wf_classes - 1;

struct wf_counters {
  int                count_ext[2];  // extend queue sizes (double buffered)
  int                count_pdf;
  int                pad;
  unsigned long long rays_extend, rays_lightpdf, paths_done;
  int                class_count[wf_classes];
  int                ext_head;  // next unclaimed entry of the extend queue (persistent extend)
  int                pdf_head, pad2;  // next unclaimed entry of the light-pdf queue (persistent light pdf)
  int                count_shade;    // entries of the class-sorted queue that k_wf_shade runs
  int                susp_count[2];  // suspended-ray records in use, per pool
  unsigned long long rays_resumed;   // queue entries that were continuations, not new rays
};

// One 128-byte record per slot, 128-byte aligned = two 64-byte DRAM bursts, four 32-byte
// sectors laid out by who writes them (ncu r01f / r02a: the former 160-byte record straddled
// 3-4 bursts, k_wf_shade moved 586 B per slot for 320 B of state and k_wf_extend 157 B for 56):
//   sector 0  the next ray + flags + remaining samples: written by shade / tail, READ by extend
//   sector 1  the hit: written whole by extend (no partial-sector fill), read by shade
//   sector 2  radiance and weight: shade / tail
//   sector 3  pending MIS terms and the pixel's generator: shade / tail
// The medium of a path inside a volume (volpathtrace) lives in a separate 64-byte record per
// slot (`vol`), touched only when a path enters a volume or scatters inside one.
// Keeping a slot's state contiguous matters because the shade queue is class-sorted, i.e.
// slots are visited in scattered order: nine separate SoA arrays cost nine half-used sectors
// per path (ncu r01: 1.0 GB of DRAM traffic per k_wf_shade launch against 0.2 GB of state).
enum : int {
  wf_ray0 = 0,   // o.xyz, d.x
  wf_ray1 = 1,   // d.yz, flags, remaining samples
  wf_hit  = 2,   // instance, element, uv
  wf_hit1 = 3,   // hit distance, hit flag, light pdf (k_wf_lightpdf), 0
  wf_rad  = 4,   // radiance.xyz, weight.x
  wf_w1   = 5,   // weight.yz, vol_anisotropy, 0
  wf_mis  = 6,   // pending f.xyz, bsdf pdf
  wf_rng  = 7,   // the pixel's PCG32 state while the slot is live
  wf_rec  = 8,
  wf_vol0 = 0, wf_vol1 = 1, wf_vol2 = 2,  // `vol` record: density, scattering, emission
  wf_vol_rec = 4
};

// device pointers of one state's wavefront buffers
struct wf_view {
  f4* rec;
  f4* vol;
  __device__ __forceinline__ f4& at(int field, int slot) const {
    return rec[(size_t)slot * wf_rec + field];
  }
  __device__ __forceinline__ f4& vol_at(int field, int slot) const {
    return vol[(size_t)slot * wf_vol_rec + field];
  }
  int *       queue_ext[2], *queue_pdf;
  int *       queue_shade, *sort_key;  // class-sorted slots; (class << 24 | rank) per entry
  wf_counters* counters;
  // suspended traversals (k_wf_extend_persistent): two pools of `susp_f4`-f4 records,
  // written by the extend launch of one iteration and read by the next one
  f4* susp[2];
  int susp_f4;
};

// One pipe = an independent wavefront over an interleaved share of the slots (tile t belongs
// to pipe t % pipes), with its own queues, counters and CUDA stream. Several pipes let the
// latency-bound kernels of one (shade, light pdf, the draining end of extend) run in the issue
// slots the other leaves idle; pixels do not interact, so results do not depend on the split.
constexpr int wf_max_pipes = 4;
struct wf_pipe {
  wf_view      view{};
  cudaStream_t stream = nullptr;
  wf_counters* host_counters[2] = {nullptr, nullptr};  // pinned, one per batch in flight
  cudaEvent_t  poll[2] = {nullptr, nullptr};
  int*         queues = nullptr;
  size_t       susp_bytes = 0;
  // one batch of wf_batch iterations as a CUDA graph (kernel arguments are baked in: `graph_key`
  // holds their bytes, a different scene / state / shader / size re-captures)
  static constexpr int graph_classes = 7;  // grids of full, 1/2, ... 1/64 of the pipe's slots
  cudaGraphExec_t   graph[graph_classes] = {};
  std::vector<char> graph_key[graph_classes];
  // host-side progress
  int active = 0, cur = 0, inflight = 0, batches = 0;
  int polled_cur[2] = {0, 0};
  bool draining = false;
};

struct wavefront_buffers {
  bool        live     = false;
  int         capacity = 0, pipes = 0;
  f4*         rec = nullptr;
  f4*         vol = nullptr;
  wf_pipe     pipe[wf_max_pipes];
  cudaEvent_t k0 = nullptr, k1 = nullptr;
  void release() {
    if (!capacity) return;
    cudaFree(rec);
    if (vol) cudaFree(vol);
    for (auto& p : pipe) {
      if (p.queues) cudaFree(p.queues);
      if (p.view.counters) cudaFree(p.view.counters);
      if (p.view.susp[0]) cudaFree(p.view.susp[0]);
      for (int i = 0; i < 2; i++) {
        if (p.host_counters[i]) cudaFreeHost(p.host_counters[i]);
        if (p.poll[i]) cudaEventDestroy(p.poll[i]);
      }
      for (auto g : p.graph)
        if (g) cudaGraphExecDestroy(g);
      if (p.stream) cudaStreamDestroy(p.stream);
    }
    if (k0) cudaEventDestroy(k0);
    if (k1) cudaEventDestroy(k1);
    *this = wavefront_buffers{};
  }
  // `susp_records` suspended-ray records of `susp_f4` f4 each, per pool and pipe
  cudaError_t reserve_suspended(int susp_records, int susp_f4) {
    size_t bytes = sizeof(f4) * (size_t)susp_f4 * (size_t)susp_records;
    for (int i = 0; i < pipes; i++) {
      auto& p = pipe[i];
      p.view.susp_f4 = susp_f4;
      if (bytes > p.susp_bytes) {
        if (p.view.susp[0]) cudaFree(p.view.susp[0]);
        p.view.susp[0] = p.view.susp[1] = nullptr, p.susp_bytes = 0;
        auto e = cudaMalloc(&p.view.susp[0], 2 * bytes);
        if (e != cudaSuccess) return e;
        p.susp_bytes = bytes;
      }
      p.view.susp[1] = p.view.susp[0] + (size_t)susp_f4 * susp_records;
    }
    return cudaSuccess;
  }
  // slots of pipe i: whole tiles (32 slots), tile t -> pipe t % npipes
  static int pipe_slots(int slots, int npipes, int i) {
    int tiles = slots / 32;
    return ((tiles - i + npipes - 1) / npipes) * 32;
  }
  cudaError_t reserve(int slots, int npipes) {
    if (capacity >= slots && pipes == npipes) return cudaSuccess;
    release();
    auto e = reserve_all(slots, npipes);
    // a half-built set of buffers must never satisfy the early-out above
    if (e != cudaSuccess) release();
    return e;
  }
  cudaError_t reserve_all(int slots, int npipes) {
    auto e = cudaMalloc(&rec, sizeof(f4) * wf_rec * (size_t)slots);  // 256-byte aligned
    if (e != cudaSuccess) return e;
    capacity = slots, pipes = npipes;  // from here on release() frees what exists
    e = cudaMalloc(&vol, sizeof(f4) * wf_vol_rec * (size_t)slots);
    if (e != cudaSuccess) return e;
    for (int i = 0; i < npipes; i++) {
      auto&  p = pipe[i];
      size_t n = (size_t)pipe_slots(slots, npipes, 0);
      e = cudaMalloc(&p.queues, sizeof(int) * 5 * n);
      if (e != cudaSuccess) return e;
      auto q = p.queues;
      p.view.rec = rec, p.view.vol = vol;
      p.view.queue_ext[0] = q, p.view.queue_ext[1] = q + n, p.view.queue_pdf = q + 2 * n;
      p.view.queue_shade = q + 3 * n, p.view.sort_key = q + 4 * n;
      e = cudaMalloc(&p.view.counters, sizeof(wf_counters));
      for (int k = 0; k < 2 && e == cudaSuccess; k++) {
        e = cudaMallocHost(&p.host_counters[k], sizeof(wf_counters));
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p.poll[k], cudaEventDisableTiming);
      }
      // blocking streams: ordered against the legacy default stream the caller's events use
      if (e == cudaSuccess) e = cudaStreamCreate(&p.stream);
      if (e != cudaSuccess) return e;
    }
    e = cudaEventCreate(&k0);
    if (e == cudaSuccess) e = cudaEventCreate(&k1);
    return e;
  }
};

#if defined(__CUDACC__)

// ---- per-slot state <-> registers -------------------------------------------
constexpr int wf_bit_hit = 1 << 16, wf_bit_vol = 1 << 17;
constexpr int wf_bit_susp = 1 << 19;  // the slot's ray is a suspended traversal to be resumed

// `store_vol`: the volume fields only change when a path enters a volume
template <int S>
__device__ __forceinline__ void wf_store_path(const wf_view& w, int slot, const path_t& p,
    int remaining, bool store_vol) {
  int flags = (p.bounce & 0xffff) | (p.hit ? wf_bit_hit : 0) | (p.in_volume ? wf_bit_vol : 0);
  w.at(wf_ray0, slot) = {p.o.x, p.o.y, p.o.z, p.d.x};
  w.at(wf_ray1, slot) = {p.d.y, p.d.z, i2f(flags), i2f(remaining)};
  w.at(wf_rad, slot)  = {p.radiance.x, p.radiance.y, p.radiance.z, p.weight.x};
  w.at(wf_w1, slot)   = {p.weight.y, p.weight.z, p.vol_anisotropy, 0};
  if (S == shader_volpathtrace && store_vol) {
    w.vol_at(wf_vol0, slot) = {p.vol_density.x, p.vol_density.y, p.vol_density.z,
        p.vol_scattering.x};
    w.vol_at(wf_vol1, slot) = {p.vol_scattering.y, p.vol_scattering.z, p.vol_emission.x,
        p.vol_emission.y};
    w.vol_at(wf_vol2, slot) = {p.vol_emission.z, 0, 0, 0};
  }
}

// volume fields are only meaningful (and only loaded) while the path is inside one
template <int S>
__device__ __forceinline__ void wf_load_path(const wf_view& w, int slot, path_t& p,
    int& remaining) {
  auto a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot), a2 = w.at(wf_rad, slot),
       a3 = w.at(wf_w1, slot);
  p.o = {a0.x, a0.y, a0.z}, p.d = {a0.w, a1.x, a1.y};
  int flags = f2i(a1.z);
  remaining = f2i(a1.w);
  p.bounce = flags & 0xffff, p.hit = (flags & wf_bit_hit) != 0;
  p.in_volume = (flags & wf_bit_vol) != 0;
  p.radiance = {a2.x, a2.y, a2.z}, p.weight = {a2.w, a3.x, a3.y};
  p.vol_anisotropy = a3.z;
  p.vol_density = p.vol_scattering = p.vol_emission = {0, 0, 0};
  if (S == shader_volpathtrace && p.in_volume) {
    auto v0 = w.vol_at(wf_vol0, slot), v1 = w.vol_at(wf_vol1, slot), v2 = w.vol_at(wf_vol2, slot);
    p.vol_density = {v0.x, v0.y, v0.z}, p.vol_scattering = {v0.w, v1.x, v1.y};
    p.vol_emission = {v1.z, v1.w, v2.x};
  }
}

__device__ __forceinline__ rng_t wf_load_rng(const wf_view& w, int slot) {
  return *reinterpret_cast<const rng_t*>(&w.at(wf_rng, slot));
}
__device__ __forceinline__ void wf_store_rng(const wf_view& w, int slot, const rng_t& rng) {
  *reinterpret_cast<rng_t*>(&w.at(wf_rng, slot)) = rng;
}

// warp-aggregated append of `slot` to a queue for the lanes with `push` set
__device__ __forceinline__ void wf_push(int* queue, int* count, bool push, int slot) {
  unsigned mask = __ballot_sync(0xffffffffu, push);
  if (!mask) return;
  int lane   = threadIdx.x & 31;
  int leader = __ffs(mask) - 1;
  int base   = 0;
  if (lane == leader) base = atomicAdd(count, __popc(mask));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (push) queue[base + __popc(mask & ((1u << lane) - 1))] = slot;
}

// A finished path: accumulate (owner thread, no atomics), then regenerate the
// pixel's next camera sample in place. Returns false when the slot is done.
template <int S>
__device__ __forceinline__ bool wf_finish(const DScene& sc, const DParams& prm, int idx,
    path_t& p, rng_t& rng, f4* image, int& remaining) {
  auto rad   = guard_radiance(path_result(S, p));
  image[idx] = image[idx] + rad;
  remaining -= 1;
  if (remaining <= 0) return false;
  auto ray = sample_camera(sc, prm, idx, rng);
  path_init(p, ray);
  return true;
}


// ---- suspended traversals ------------------------------------------------------
// Once the extend queue is exhausted a persistent warp only drains: its lanes finish one by
// one and the warp keeps issuing for the few that are left (ncu r01: 16.7 of 32 lanes held a
// ray). Instead, when at most `wf_suspend` lanes are still walking, the warp parks them: the
// walk's registers and stack go to a record in HBM, the slot goes to the NEXT iteration's
// extend queue flagged wf_bit_susp, and that launch resumes the walk where it stopped, packed
// into full warps with the new rays. The ray's own sequence of operations is unchanged, so
// hits stay bit-identical; only the iteration in which the slot reaches shade moves.
// Record: 7 f4 of registers, then the stack entries two per f4.
constexpr int wf_susp_head = 7;
inline int wf_susp_record_f4(int stack_depth) { return wf_susp_head + (stack_depth + 1) / 2; }

__device__ __forceinline__ void wf_suspend_ray(f4* rec, const trav_t& t, stack_t st) {
  int bits = (t.plain ? 1 : 0) | (t.wplain ? 2 : 0) | (t.cull ? 4 : 0) | (t.h.hit ? 8 : 0) |
             (t.wide ? 16 : 0) | (t.wwide ? 32 : 0) | (t.oct << 8) | (t.woct << 12);
  rec[0] = {t.o.x, t.o.y, t.o.z, t.d.x};
  rec[1] = {t.d.y, t.d.z, t.dinv.x, t.dinv.y};
  rec[2] = {t.dinv.z, t.wdinv.x, t.wdinv.y, t.wdinv.z};
  rec[3] = {t.tmin, t.tmax, t.omax, i2f(bits)};
  rec[4] = {i2f(t.kind), i2f(t.inst), i2f(t.cur), i2f(t.h.instance)};
  rec[5] = {i2f(t.h.element), t.h.uv.x, t.h.uv.y, t.h.distance};
  for (int i = 0; i < t.cur; i += 2)
    rec[wf_susp_head + (i >> 1)] = {i2f(st.ref(i)), st.t0(i), i + 1 < t.cur ? i2f(st.ref(i + 1)) : 0,
        i + 1 < t.cur ? st.t0(i + 1) : 0};
}
__device__ __forceinline__ void wf_resume_ray(const f4* rec, trav_t& t, stack_t st) {
  auto a = rec[0], b = rec[1], c = rec[2], d = rec[3], e = rec[4], f = rec[5];
  t.o = {a.x, a.y, a.z}, t.d = {a.w, b.x, b.y};
  t.dinv = {b.z, b.w, c.x}, t.wdinv = {c.y, c.z, c.w};
  t.tmin = d.x, t.tmax = d.y, t.omax = d.z;
  int bits = f2i(d.w);
  t.plain = bits & 1, t.wplain = bits & 2, t.cull = bits & 4;
  t.wide = bits & 16, t.wwide = bits & 32, t.oct = (bits >> 8) & 7, t.woct = (bits >> 12) & 7;
  t.kind = f2i(e.x), t.inst = f2i(e.y), t.cur = f2i(e.z);
  t.h = hit_t{f2i(e.w), f2i(f.x), {f.y, f.z}, f.w, (bits & 8) != 0};
  for (int i = 0; i < t.cur; i += 2) {
    auto s = rec[wf_susp_head + (i >> 1)];
    st.push(i, f2i(s.x), s.y);
    if (i + 1 < t.cur) st.push(i + 1, f2i(s.z), s.w);
  }
}

// ---- leaf tests shared by the warp ----------------------------------------------------
// After a descent only some lanes hold a leaf, each with 1..4 primitives; tested lane by
// lane the warp runs the (long) primitive test 3-4 times at ~5 of 32 lanes (ncu r01c). Here
// the (lane, primitive) pairs of the whole warp are laid out densely, one per lane: a lane
// fetches the owner's ray with shuffles, tests ONE primitive against the owner's tmax, and
// every owner then folds its primitives' distances in stored order with `t <= tmax`, exactly
// as the sequential loop does (each test sees a tmax no smaller than the sequential one, so
// it reports a superset of hits with the same distances; the fold rejects the surplus and
// the later of two equal distances still wins).
// Measured on B200: no gain (03_volume 181.5 ms per 64 spp with it, 177.8 ms without: the ten
// shuffles per pass, the fold and 58 bytes of spills cost what the denser tests save), so the
// per-lane loop stays the default and this path is opt-in (-DVPT_WF_COOP_LEAVES); it passes the
// same bit-exact parity suite.
#if defined(VPT_WF_COOP_LEAVES)
__device__ __forceinline__ void wf_test_leaves_warp(const DScene& sc, trav_t& t, int leaf,
    unsigned char* tab) {
  const unsigned full = 0xffffffffu;
  if (__ballot_sync(full, leaf != 0) == 0) return;
  int lane = threadIdx.x & 31;
  int n    = leaf != 0 ? (leaf & 7) : 0;
  int incl = n;
  for (int o = 1; o < 32; o <<= 1) {
    int v = __shfl_up_sync(full, incl, o);
    if (lane >= o) incl += v;
  }
  int total = __shfl_sync(full, incl, 31);
  int nmax  = __reduce_max_sync(full, n);
  int excl  = incl - n;
  for (int k = 0; k < n; k++) tab[excl + k] = (unsigned char)(lane | (k << 5));
  __syncwarp();
  int lstart = leaf >> 3;
  for (int base = 0; base < total; base += 32) {  // warp-uniform
    int  item  = base + lane;
    bool valid = item < total;
    int  code  = valid ? tab[item] : lane;
    int  src = code & 31, k = code >> 5;
    f3    o  = {__shfl_sync(full, t.o.x, src), __shfl_sync(full, t.o.y, src), __shfl_sync(full, t.o.z, src)};
    f3    d  = {__shfl_sync(full, t.d.x, src), __shfl_sync(full, t.d.y, src), __shfl_sync(full, t.d.z, src)};
    float tmin = __shfl_sync(full, t.tmin, src), tmax = __shfl_sync(full, t.tmax, src);
    int   kind = __shfl_sync(full, t.kind, src), ls = __shfl_sync(full, lstart, src);
    float rt = __int_as_float(0x7fc00000);  // NaN: no hit (fails `rt <= tmax`)
    f2    ruv = {0, 0};
    int   relem = -1;
    if (valid) {
      float dist;
      if (kind == kind_quads) {
        auto p = sc.prims + ls + 4 * k;
        auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2), q = ld4(p + 3);
        if (intersect_quad(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), xyz(q), ruv, dist))
          rt = dist, relem = f2i(a.w);
      } else {
        auto p = sc.prims + ls + 3 * k;
        auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2);
        if (intersect_triangle(o, d, tmin, tmax, xyz(a), xyz(b), xyz(c), ruv, dist))
          rt = dist, relem = f2i(a.w);
      }
    }
    // owners fold the items of this pass that are theirs, in stored order
    int win = -1;
    for (int k2 = 0; k2 < nmax; k2++) {  // warp-uniform
      int   it   = excl + k2;
      bool  mine = k2 < n && it >= base && it < base + 32;
      float tt   = __shfl_sync(full, rt, mine ? it - base : lane);
      if (mine && tt <= t.tmax) t.tmax = tt, win = it - base;
    }
    int   from = win >= 0 ? win : lane;
    float u = __shfl_sync(full, ruv.x, from), v = __shfl_sync(full, ruv.y, from);
    int   el = __shfl_sync(full, relem, from);
    if (win >= 0) {
      t.h.hit = true, t.h.instance = t.inst, t.h.element = el;
      t.h.uv = {u, v}, t.h.distance = t.tmax;
    }
  }
  __syncwarp();  // the table is rewritten next round
}
#endif

// ---- kernels -----------------------------------------------------------------
template <int S>
__global__ void __launch_bounds__(wf_block)
    k_wf_init(const DScene sc, const DParams prm, wf_view w, rng_t* __restrict__ rngs, int slots,
        int nsamples, int pipe, int npipes) {
  int  tid    = blockIdx.x * blockDim.x + threadIdx.x;
  int  slot   = (((tid >> 5) * npipes + pipe) << 5) | (tid & 31);  // this pipe's tiles
  bool active = false;
  if (slot < slots) {
    int idx = tile_pixel_index(slot, prm.width, prm.height);
    if (idx >= 0) {
      auto   rng = rngs[idx];
      path_t p;
      path_init(p, sample_camera(sc, prm, idx, rng));
      wf_store_path<S>(w, slot, p, nsamples, false);
      wf_store_rng(w, slot, rng);
      active = true;
    }
  }
  wf_push(w.queue_ext[0], &w.counters->count_ext[0], active, slot);
}

template <int S>
__global__ void __launch_bounds__(wf_block)
    k_wf_extend(const DScene sc, const DParams prm, wf_view w, int cur) {
  extern __shared__ int stack_mem[];
  int  tid   = blockIdx.x * blockDim.x + threadIdx.x;
  int  count = w.counters->count_ext[cur];
  if (tid == 0) {
    // the queues written later in this iteration start empty
    w.counters->count_ext[cur ^ 1] = 0;
    w.counters->count_pdf          = 0;
    w.counters->rays_extend += (unsigned long long)count;
  }
  if (tid >= count) return;
  int     slot = w.queue_ext[cur][tid];
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  auto    a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot);
  path_t  p;
  p.o = {a0.x, a0.y, a0.z}, p.d = {a0.w, a1.x, a1.y};
  auto isec  = find_hit<S>(sc, prm, p, st);
  w.at(wf_hit, slot)  = {i2f(isec.instance), i2f(isec.element), isec.uv.x, isec.uv.y};
  w.at(wf_hit1, slot) = {isec.distance, i2f(isec.hit ? 1 : 0), 0, 0};
  // shade class and the rank inside it
  int cls = 0;
  if (isec.hit) {
    int material;
    if (is_sdf_shader<S>())
      material = isec.instance != invalidid ? f2i(ld4(sc.vol_instances + vol_inst_f4 * isec.instance + 2).w)
                                            : f2i(ld4(sc.sdfs + 10 * isec.element + 1).w);
    else
      material = f2i(ld4(sc.inst_trav + 4 * isec.instance + 3).w);
    cls = 2 + (material < wf_classes - 4 ? material : wf_classes - 4);
    if (S == shader_volpathtrace && (f2i(a1.z) & wf_bit_vol)) cls = 1;
  }
  unsigned peers  = __match_any_sync(__activemask(), cls);
  int      lane   = threadIdx.x & 31;
  int      leader = __ffs(peers) - 1;
  int      base   = 0;
  if (lane == leader) base = atomicAdd(&w.counters->class_count[cls], __popc(peers));
  base = __shfl_sync(peers, base, leader);
  w.sort_key[tid] = (cls << 24) | (base + __popc(peers & ((1u << lane) - 1)));
}

// Persistent form of k_wf_extend. The grid is sized to the machine (one resident set of
// warps); a warp claims queue entries for its idle lanes with one atomicAdd, runs rounds
// (mesh shaders: trav_t::round = descend to a leaf and process it; SDF shaders: a few
// sphere-tracing steps), retires the lanes whose walk is over and refills them as soon as
// `wf_refill` lanes idle. A fixed ray-per-lane launch keeps a warp busy until its slowest ray
// ends (ncu r01: 14.8 of 32 lanes live per traversal round, 10 of 32 in the SDF march).
// Per ray the walk is the same sequence of operations, so hits are bit-identical.
#if !defined(VPT_WF_REFILL)
#define VPT_WF_REFILL 8
#endif
#if !defined(VPT_EXT_BLOCKS)
#define VPT_EXT_BLOCKS 6
#endif
constexpr int wf_refill    = VPT_WF_REFILL;  // BVH walks
constexpr int wf_refill_sdf = 4;             // sphere tracing (B200: 07_sdfunction 147.6 -> 143.9 ms per 16 spp)
#if !defined(VPT_WF_SUSPEND_AT)
#define VPT_WF_SUSPEND_AT 20
#endif
constexpr int wf_suspend   = VPT_WF_SUSPEND_AT;  // 0: never park (drain in place)
#if !defined(VPT_WF_SDF_STEPS)
#define VPT_WF_SDF_STEPS 4
#endif
constexpr int wf_sdf_steps = VPT_WF_SDF_STEPS;

// ROOMY = 1: the same kernel compiled for one resident block fewer per SM (96 instead of 80
// registers). Measured on B200: triangle meshes walk faster with the registers (05_head1ss
// 203.1 -> 193.9 ms per 64 spp, 04_head1 33.0 -> 31.4, 01_surface 52.2 -> 50.7), the all-quad
// 03_volume and sphere tracing with the residency (170.9 vs 172.4, 135.5 vs 140.2 ms); running
// the 80-register build at fewer blocks changes nothing, so it is the registers. The host picks
// by the scene's triangle count (wf_roomy_triangles); results do not depend on the choice.
constexpr long long wf_roomy_triangles = 2048;
template <int S, int ROOMY>
__global__ void __launch_bounds__(wf_block, VPT_EXT_BLOCKS - ROOMY)
    k_wf_extend_persistent(const DScene sc, const DParams prm, wf_view w, int cur, int suspend_at) {
  extern __shared__ int stack_mem[];
#if defined(VPT_WF_COOP_LEAVES)
  __shared__ unsigned char leaf_tab[wf_block / 32][32 * 8];
#endif
  constexpr bool sdf  = is_sdf_shader<S>();
  const unsigned full = 0xffffffffu;
  int     count = w.counters->count_ext[cur];
  int     lane  = threadIdx.x & 31;
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // count_ext[cur ^ 1] was reset by the previous iteration's k_wf_lightpdf: parked
    // rays are appended to it while this kernel still runs
    w.counters->count_pdf = 0;
    w.counters->rays_extend += (unsigned long long)count;
  }
  trav_t   t;
  sphere_t sp;
  ray3     ray     = make_ray({0, 0, 0}, {0, 0, 1});
  int      qidx    = 0, slot = 0, flags = 0;
  bool     has_ray = false;
  bool     more    = true;   // warp-uniform: the queue may still hold unclaimed entries
  bool     running = true;   // warp-uniform, the only exit
  int      resumed = 0;
#if defined(VPT_EXT_ELECT)
  int held = 0, held_arg = 0;  // the operation this lane's ray holds (trav_t::op_*)
#if !defined(VPT_ELECT_POPS)
#define VPT_ELECT_POPS 3
#endif
#endif
  // Measured on B200: parking pays for sphere tracing (07_sdfunction 136 -> 100 ms
  // per 16 spp), not for BVH walks, whose parked state is 20x larger and whose pixels' tail
  // then lasts more iterations (03_volume 209 -> 217 ms); -DVPT_WF_SUSPEND_MESH opts in.
#if defined(VPT_WF_SUSPEND_MESH)
  const int suspend_mesh = suspend_at;
#else
  constexpr int suspend_mesh = 0;
#endif
  t.cur       = 0;
  sp.finished = true;
  // (Leaf postponement - a lane holds the leaf it reached until >= N lanes of the warp hold
  // one - was measured on B200 and dropped: with <= 8 pops per round nearly every walking lane
  // ends its round on a leaf anyway, so N = 4 ... 28 changed neither the instruction count
  // (145.0 M vs 144.7-145.2 M per launch, profiles/r02a_*) nor the step time (180.5 ms).)
  while (running) {
    unsigned idle = __ballot_sync(full, !has_ray);
    if (more && (__popc(idle) >= (sdf ? wf_refill_sdf : wf_refill))) {
      int base = 0;
      if (lane == 0) base = atomicAdd(&w.counters->ext_head, __popc(idle));
      base = __shfl_sync(full, base, 0);
      more = base + __popc(idle) < count;
      int mine = base + __popc(idle & ((1u << lane) - 1));
      if (!has_ray && mine < count) {
        qidx    = mine;
        slot    = w.queue_ext[cur][qidx];
        auto a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot);
        flags   = f2i(a1.z);
        ray     = make_ray({a0.x, a0.y, a0.z}, {a0.w, a1.x, a1.y});
        if constexpr (sdf) {
          sp.begin(ray);
          if (flags & wf_bit_susp) {
            auto s = w.at(wf_hit, slot);
            sp.t = s.x, sp.i = f2i(s.y);
          }
        } else {
          if (flags & wf_bit_susp)
            wf_resume_ray(w.susp[cur] + (size_t)f2i(w.at(wf_hit, slot).x) * w.susp_f4, t, st);
          else
            t.begin(sc, -1, ray, st);
        }
        resumed += (flags & wf_bit_susp) ? 1 : 0;
        flags &= ~wf_bit_susp;
        has_ray = true;
      }
    }
    if constexpr (sdf) {
      for (int k = 0; k < wf_sdf_steps; k++)
        if (has_ray && !sp.done()) sp.step(sc, ray, prm.spheretrace_maxiter);
    } else {
#if defined(VPT_EXT_ELECT)
      // Election (tests/hostsim/warpsim "elect"): every lane pops until it holds a node visit,
      // an instance entry or a mesh leaf; the warp then runs the operation most lanes hold, and
      // any other that VPT_EXT_ELECT lanes hold. The others keep theirs for a later round.
      if (has_ray && held == 0 && !t.done()) held = t.fetch(sc, ray, st, VPT_ELECT_POPS, held_arg);
      {
        int cn = __popc(__ballot_sync(full, has_ray && held == trav_t::op_node));
        int ci = __popc(__ballot_sync(full, has_ray && held == trav_t::op_instance));
        int cl = __popc(__ballot_sync(full, has_ray && held == trav_t::op_leaf));
        int best = cn > ci ? (cn > cl ? cn : cl) : (ci > cl ? ci : cl);
        if (has_ray && held == trav_t::op_node && (cn == best || cn >= VPT_EXT_ELECT))
          t.visit_node(sc, held_arg, st), held = 0;
        if (has_ray && held == trav_t::op_instance && (ci == best || ci >= VPT_EXT_ELECT))
          t.enter_instance(sc, ray, held_arg, st), held = 0;
        if (has_ray && held == trav_t::op_leaf && (cl == best || cl >= VPT_EXT_ELECT))
          t.test_leaf(sc, held_arg), held = 0;
      }
#elif !defined(VPT_WF_COOP_LEAVES)
      if (has_ray && !t.done()) t.round(sc, ray, st);
#else
      int leaf = 0;
      if (has_ray && !t.done()) leaf = t.descend(sc, ray, st);
      wf_test_leaves_warp(sc, t, leaf, leaf_tab[threadIdx.x >> 5]);
#endif
    }
#if defined(VPT_EXT_ELECT)
    bool     retire   = has_ray && (sdf ? sp.done() : (t.done() && held == 0));
#else
    bool     retire   = has_ray && (sdf ? sp.done() : t.done());
#endif
    unsigned retiring = __ballot_sync(full, retire);
    if (retire) {
      isect_t isec;
      if constexpr (sdf)
        isec = {sp.r.hit, sp.r.instance, sp.r.sdf, {0, 0}, sp.r.dist};
      else
        isec = {t.h.hit, t.h.instance, t.h.element, t.h.uv, t.h.distance};
      w.at(wf_hit, slot)  = {i2f(isec.instance), i2f(isec.element), isec.uv.x, isec.uv.y};
      w.at(wf_hit1, slot) = {isec.distance, i2f(isec.hit ? 1 : 0), 0, 0};
      // (a resumed ray's park flag in sector 0 is cleared by shade, which rewrites the flags
      // of every slot it sees before the next extend launch reads them)
      int cls = 0;
      if (isec.hit) {
        int material;
        if constexpr (sdf)
          material = isec.instance != invalidid
                         ? f2i(ld4(sc.vol_instances + vol_inst_f4 * isec.instance + 2).w)
                         : f2i(ld4(sc.sdfs + 10 * isec.element + 1).w);
        else
          material = f2i(ld4(sc.inst_trav + 4 * isec.instance + 3).w);
        cls = 2 + (material < wf_classes - 4 ? material : wf_classes - 4);
        if (S == shader_volpathtrace && (flags & wf_bit_vol)) cls = 1;
      }
      unsigned peers  = __match_any_sync(retiring, cls);
      int      leader = __ffs(peers) - 1;
      int      base   = 0;
      if (lane == leader) base = atomicAdd(&w.counters->class_count[cls], __popc(peers));
      base = __shfl_sync(peers, base, leader);
      w.sort_key[qidx] = (cls << 24) | (base + __popc(peers & ((1u << lane) - 1)));
      has_ray = false;
    }
    unsigned walking = __ballot_sync(full, has_ray);
    if (!more && walking && __popc(walking) <= (sdf ? suspend_at : suspend_mesh)) {

      // park what is left of this warp for the next iteration's launch
      int n = __popc(walking), leader = __ffs(walking) - 1;
      int rank = __popc(walking & ((1u << lane) - 1));
      int rec = 0, at = 0, order = 0;
      if (lane == leader) {
        if constexpr (!sdf) rec = atomicAdd(&w.counters->susp_count[cur ^ 1], n);
        at    = atomicAdd(&w.counters->count_ext[cur ^ 1], n);
        order = atomicAdd(&w.counters->class_count[wf_class_susp], n);
      }
      rec = __shfl_sync(full, rec, leader), at = __shfl_sync(full, at, leader);
      order = __shfl_sync(full, order, leader);
      if (has_ray) {
        if constexpr (sdf) {
          w.at(wf_hit, slot) = {sp.t, i2f(sp.i), 0, 0};
        } else {
          wf_suspend_ray(w.susp[cur ^ 1] + (size_t)(rec + rank) * w.susp_f4, t, st);
          w.at(wf_hit, slot).x = i2f(rec + rank);
        }
        w.at(wf_ray1, slot).z          = i2f(flags | wf_bit_susp);
        w.queue_ext[cur ^ 1][at + rank] = slot;
        w.sort_key[qidx]               = (wf_class_susp << 24) | (order + rank);
        has_ray = false;
      }
    }
    running = more || __any_sync(full, has_ray);
  }
  for (int o = 16; o > 0; o >>= 1) resumed += __shfl_down_sync(full, resumed, o);
  if (lane == 0 && resumed) atomicAdd(&w.counters->rays_resumed, (unsigned long long)resumed);
}

// ---- Pooled extend (mesh shaders) -------------------------------------------------
// k_wf_extend_persistent keeps one ray per lane in registers, so at any instruction the lanes of
// a warp are spread over the walk's operations (node visit / instance entry / leaf test / pops):
// ncu r02c counts 7.7 of 32 lanes per instruction, and tests/hostsim/warpsim reproduces it from
// the rays alone. Here a warp owns a POOL of wf_pool_k * 32 rays whose state lives in shared
// memory. Every round it counts how many of its rays hold each kind of operation, and for every
// kind that enough rays hold (else for the one most rays hold) it compacts up to 32 of those
// rays onto its lanes, runs that one operation for all of them, pops each ray's next stack
// entry to find what it holds next, and writes the changed fields back. Per ray the operations
// and their order are those of trav_t::descend (the same member functions run them): hits are
// bit-identical, only the SIMT schedule differs.
// Shared memory per ray: 19 words of state + `limit` stack entries; deeper entries (the
// uploader's depth bound is a worst case) go to a global spill column.
#if !defined(VPT_POOL_K)
#define VPT_POOL_K 3
#endif
#if !defined(VPT_POOL_BLOCKS)
#define VPT_POOL_BLOCKS 3
#endif
#if !defined(VPT_POOL_RUN_AT)
#define VPT_POOL_RUN_AT 24
#endif
constexpr int wf_pool_k = VPT_POOL_K;  // rays per lane of a warp's pool
enum : int {
  pf_ox, pf_oy, pf_oz, pf_dx, pf_dy, pf_dz, pf_ix, pf_iy, pf_iz,  // ray in the current space, 1 / d
  pf_wix, pf_wiy, pf_wiz,                                        // 1 / d of the world-space ray
  pf_tmax, pf_inst, pf_hinst, pf_arg, pf_qidx, pf_slot, pf_ctl,
  pf_fields
};
// pf_ctl: what the ray holds (3 bits) | stack height (7) | kind + 1 (2) | plain | wide | oct (3)
//         | wplain | wwide | woct (3) | cull | hit | in-volume
constexpr int pc_pop = 0,    // nothing yet: pop
              pc_node = 1,   // a node to visit
              pc_leaf = 2,   // a mesh leaf to test
              pc_inst = 3,   // an instance to enter
              pc_done = 4,   // walk over: retire
              pc_sleaf = 5,  // a scene leaf: queue its instances
              pc_idle = 7;
constexpr int pc_hit = 1 << 23, pc_vol = 1 << 24;
__device__ __forceinline__ int pool_pack(const trav_t& t, int op, int keep) {
  return op | (t.cur << 3) | ((t.kind + 1) << 10) | (t.plain ? 1 << 12 : 0) | (t.wide ? 1 << 13 : 0) |
         (t.oct << 14) | (t.wplain ? 1 << 17 : 0) | (t.wwide ? 1 << 18 : 0) | (t.woct << 19) |
         (t.cull ? 1 << 22 : 0) | (keep & (pc_hit | pc_vol));
}
__device__ __forceinline__ void pool_unpack(trav_t& t, int ctl) {
  t.cur = (ctl >> 3) & 127, t.kind = ((ctl >> 10) & 3) - 1;
  t.plain = (ctl >> 12) & 1, t.wide = (ctl >> 13) & 1, t.oct = (ctl >> 14) & 7;
  t.wplain = (ctl >> 17) & 1, t.wwide = (ctl >> 18) & 1, t.woct = (ctl >> 19) & 7;
  t.cull = (ctl >> 22) & 1;
  t.tmin = 1e-4f;  // make_ray
}

template <int S>
__global__ void __launch_bounds__(wf_block, VPT_POOL_BLOCKS)
    k_wf_extend_pool(const DScene sc, const DParams prm, wf_view w, int cur, int limit) {
  extern __shared__ int pool_mem[];
  constexpr int      K    = wf_pool_k;
  constexpr int      BR   = wf_block * K;  // rays per block
  const unsigned     full = 0xffffffffu;
  float*             state = reinterpret_cast<float*>(pool_mem);       // [pf_fields][BR]
  int*               stack = pool_mem + pf_fields * BR;                // [limit][BR] entries
  unsigned char*     ids   = reinterpret_cast<unsigned char*>(stack + 2 * limit * BR) + (threadIdx.x >> 5) * 32 * K;
  const int          lane  = threadIdx.x & 31;
  const int          r0    = (threadIdx.x >> 5) * 32 * K;  // this warp's rays: [r0, r0 + 32 K)
  const unsigned     lt    = (1u << lane) - 1;
  const int          count = w.counters->count_ext[cur];
  int*               spill = reinterpret_cast<int*>(w.susp[0]);
  const int          ostride = gridDim.x * BR;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    w.counters->count_pdf = 0;
    w.counters->rays_extend += (unsigned long long)count;
  }
#define PF(f, r) state[(f) * BR + (r)]
#define PI(f, r) reinterpret_cast<int*>(state)[(f) * BR + (r)]
  // What the ray holds next: pops until one of the operations turns up (at most 3 entries; a
  // ray that still has nothing goes to the pop round). Stale entries are dropped and a sentinel
  // restores the world-space ray on the way. Writes pf_arg and pf_ctl.
  auto next_op = [&](trav_t& t, int r, const pool_stack_t& st, int ctl) {
    int op = pc_pop, arg = 0;
    for (int it = 0; it < 3 && op == pc_pop && t.cur != 0; it++) {
      auto  en = st.get(--t.cur);
      int   e  = en.ref;
      float te = en.t0;
      if (e >= 0) {
        if (te <= t.tmax * box_slack) op = pc_node, arg = e;
      } else {
        int x = ~e, tag = x & 3;
        if (tag == ref_leaf) {
          if (te <= t.tmax * box_slack)
            op = t.kind < 0 ? pc_sleaf : pc_leaf, arg = x >> 2;
        } else if (tag == ref_instance) {
          op = pc_inst, arg = x >> 2;
        } else {
          // sentinel: back to the world-space ray (kept in the slot record)
          int  slot = PI(pf_slot, r);
          auto a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot);
          PF(pf_ox, r) = a0.x, PF(pf_oy, r) = a0.y, PF(pf_oz, r) = a0.z;
          PF(pf_dx, r) = a0.w, PF(pf_dy, r) = a1.x, PF(pf_dz, r) = a1.y;
          PF(pf_ix, r) = PF(pf_wix, r), PF(pf_iy, r) = PF(pf_wiy, r), PF(pf_iz, r) = PF(pf_wiz, r);
          t.plain = t.wplain, t.wide = t.wwide, t.oct = t.woct;
          t.kind  = -1;
        }
      }
    }
    if (op == pc_pop && t.cur == 0) op = pc_done;
    PI(pf_arg, r) = arg;
    PI(pf_ctl, r) = pool_pack(t, op, ctl);
  };
  for (int k = 0; k < K; k++) PI(pf_ctl, r0 + k * 32 + lane) = pc_idle;
  __syncwarp();
  bool more = count > 0;
  while (true) {
    // what every ray of the pool holds: one 8-bit counter per kind, two warp reductions
    int      cat[K];
    unsigned lo = 0, hi = 0;
#pragma unroll
    for (int k = 0; k < K; k++) {
      cat[k] = PI(pf_ctl, r0 + k * 32 + lane) & 7;
      lo += cat[k] < 4 ? 1u << (8 * cat[k]) : 0u;
      hi += cat[k] >= 4 ? 1u << (8 * (cat[k] - 4)) : 0u;
    }
    lo = __reduce_add_sync(full, lo), hi = __reduce_add_sync(full, hi);
    if (!more) hi &= 0x00ffffffu;  // nothing left to refill with
    if ((lo | hi) == 0) break;
    // the kinds to run this round: every kind that VPT_POOL_RUN_AT rays hold (byte-parallel
    // compare of the eight counters), else those that half as many hold, else the fullest
    auto at_least = [&](unsigned t) {
      unsigned a = ((lo | 0x80808080u) - t * 0x01010101u) & 0x80808080u;
      unsigned b = ((hi | 0x80808080u) - t * 0x01010101u) & 0x80808080u;
      // bit 7 of byte c -> bit c (lo), bit c + 4 (hi)
      a = (a >> 7) * 0x00204081u >> 21 & 15u;  // gathers bits 0, 8, 16, 24 into bits 0..3
      b = (b >> 7) * 0x00204081u >> 21 & 15u;
      return a | (b << 4);
    };
    unsigned run = at_least(VPT_POOL_RUN_AT);
    if (run == 0) run = at_least(VPT_POOL_RUN_AT / 2);
    if (run == 0) {
      int best = 0, best_kind = 0;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        int n = (int)(((c < 4 ? lo : hi) >> (8 * (c & 3))) & 255u);
        if (n > best) best = n, best_kind = c;
      }
      run = 1u << best_kind;
    }
    while (run) {
      const int pick = __ffs(run) - 1;
      run &= run - 1;
      // compact the rays holding it onto the lanes
      int n = 0;
#pragma unroll
      for (int k = 0; k < K; k++) {
        unsigned b = __ballot_sync(full, cat[k] == pick);
        if (cat[k] == pick) ids[n + __popc(b & lt)] = (unsigned char)(k * 32 + lane);
        n += __popc(b);
      }
      __syncwarp();
      int take = n < 32 ? n : 32;
      int qbase = 0;
      if (pick == pc_idle) {
        if (lane == 0) qbase = atomicAdd(&w.counters->ext_head, take);
        qbase = __shfl_sync(full, qbase, 0);
        more  = qbase + take < count;
        take  = count - qbase < take ? (count - qbase > 0 ? count - qbase : 0) : take;
      }
      const bool act = lane < take;
      const int  r   = r0 + (act ? ids[lane] : 0);
      const auto st  = pool_stack_t{stack + 2 * r, BR, spill + 2 * ((size_t)blockIdx.x * BR + r), ostride, limit};
      if (pick == pc_done) {  // write the hit, count the shade class; the entry takes a new ray
        unsigned retiring = __ballot_sync(full, act);
        if (act) {
          int  ctl  = PI(pf_ctl, r);
          int  slot = PI(pf_slot, r);
          bool hit  = (ctl & pc_hit) != 0;
          if (!hit) w.at(wf_hit, slot) = {i2f(-1), i2f(-1), 0, 0};
          w.at(wf_hit1, slot) = {hit ? PF(pf_tmax, r) : 0.0f, i2f(hit ? 1 : 0), 0, 0};
          int cls = 0;
          if (hit) {
            int material = f2i(ld4(sc.inst_trav + 4 * PI(pf_hinst, r) + 3).w);
            cls = 2 + (material < wf_classes - 4 ? material : wf_classes - 4);
            if (S == shader_volpathtrace && (ctl & pc_vol)) cls = 1;
          }
          unsigned peers  = __match_any_sync(retiring, cls);
          int      leader = __ffs(peers) - 1;
          int      base   = 0;
          if (lane == leader) base = atomicAdd(&w.counters->class_count[cls], __popc(peers));
          base = __shfl_sync(peers, base, leader);
          w.sort_key[PI(pf_qidx, r)] = (cls << 24) | (base + __popc(peers & lt));
          PI(pf_ctl, r) = pc_idle;
        }
        if (more) {
          if (lane == 0) qbase = atomicAdd(&w.counters->ext_head, take);
          qbase = __shfl_sync(full, qbase, 0);
          more  = qbase + take < count;
          take  = count - qbase < take ? (count - qbase > 0 ? count - qbase : 0) : take;
        } else {
          take = 0;
        }
      }
      if (pick == pc_done || pick == pc_idle) {
        if (lane < take) {
          trav_t t;
          int  qidx = qbase + lane;
          int  slot = w.queue_ext[cur][qidx];
          auto a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot);
          auto ray = make_ray({a0.x, a0.y, a0.z}, {a0.w, a1.x, a1.y});
          t.begin(sc, -1, ray, st);
          PF(pf_ox, r) = t.o.x, PF(pf_oy, r) = t.o.y, PF(pf_oz, r) = t.o.z;
          PF(pf_dx, r) = t.d.x, PF(pf_dy, r) = t.d.y, PF(pf_dz, r) = t.d.z;
          PF(pf_ix, r) = t.dinv.x, PF(pf_iy, r) = t.dinv.y, PF(pf_iz, r) = t.dinv.z;
          PF(pf_wix, r) = t.dinv.x, PF(pf_wiy, r) = t.dinv.y, PF(pf_wiz, r) = t.dinv.z;
          PF(pf_tmax, r) = t.tmax;
          PI(pf_inst, r) = -1, PI(pf_hinst, r) = -1;
          PI(pf_qidx, r) = qidx, PI(pf_slot, r) = slot;
          next_op(t, r, st, (f2i(a1.z) & wf_bit_vol) ? pc_vol : 0);
        }
      } else if (act) {
        trav_t t;
        int    ctl = 0;
        {
          ctl = PI(pf_ctl, r);
          pool_unpack(t, ctl);
          t.tmax = PF(pf_tmax, r);
          if (pick == pc_node) {
            t.o    = {PF(pf_ox, r), PF(pf_oy, r), PF(pf_oz, r)};
            t.dinv = {PF(pf_ix, r), PF(pf_iy, r), PF(pf_iz, r)};
            t.visit_node(sc, PI(pf_arg, r), st);
          } else if (pick == pc_sleaf) {
            // the ray is the world-space ray at the scene level
            t.o    = {PF(pf_ox, r), PF(pf_oy, r), PF(pf_oz, r)};
            t.dinv = {PF(pf_ix, r), PF(pf_iy, r), PF(pf_iz, r)};
            t.omax = fmaxf(fmaxf(fabsf(t.o.x), fabsf(t.o.y)), fabsf(t.o.z));
            t.queue_instances(sc, ray3{}, PI(pf_arg, r), st);
          } else if (pick == pc_inst) {
            auto ray = ray3{{PF(pf_ox, r), PF(pf_oy, r), PF(pf_oz, r)},
                {PF(pf_dx, r), PF(pf_dy, r), PF(pf_dz, r)}, t.tmin, 0};
            t.o = ray.o, t.d = ray.d, t.dinv = {0, 0, 0};
            int enter = PI(pf_arg, r);
            t.enter_instance(sc, ray, enter, st);
            if (t.kind >= 0) {
              PF(pf_ox, r) = t.o.x, PF(pf_oy, r) = t.o.y, PF(pf_oz, r) = t.o.z;
              PF(pf_dx, r) = t.d.x, PF(pf_dy, r) = t.d.y, PF(pf_dz, r) = t.d.z;
              PF(pf_ix, r) = t.dinv.x, PF(pf_iy, r) = t.dinv.y, PF(pf_iz, r) = t.dinv.z;
              PI(pf_inst, r) = enter;
            }
          } else if (pick == pc_leaf) {
            t.o    = {PF(pf_ox, r), PF(pf_oy, r), PF(pf_oz, r)};
            t.d    = {PF(pf_dx, r), PF(pf_dy, r), PF(pf_dz, r)};
            t.inst = PI(pf_inst, r);
            t.h    = hit_t{-1, -1, {0, 0}, 0, false};
            t.test_leaf(sc, PI(pf_arg, r));
            if (t.h.hit) {
              PF(pf_tmax, r)  = t.h.distance;
              PI(pf_hinst, r) = t.h.instance;
              w.at(wf_hit, PI(pf_slot, r)) = {i2f(t.h.instance), i2f(t.h.element), t.h.uv.x, t.h.uv.y};
              ctl |= pc_hit;
            }
          }
          // pc_pop: nothing to run, just look further down the stack
        }
        next_op(t, r, st, ctl);
      }
      __syncwarp();
    }
  }
#undef PF
#undef PI
}

// scatter the extend queue into class order (counting sort, second pass)
__global__ void __launch_bounds__(wf_block) k_wf_sort(wf_view w, int cur) {
  __shared__ int base[wf_classes];
  if (threadIdx.x < 32) {
    // exclusive scan of the 64 class counts by one warp
    int a = w.counters->class_count[2 * threadIdx.x], b = w.counters->class_count[2 * threadIdx.x + 1];
    int sum = a + b, incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
      int v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((int)threadIdx.x >= o) incl += v;
    }
    base[2 * threadIdx.x]     = incl - sum;
    base[2 * threadIdx.x + 1] = incl - sum + a;
  }
  __syncthreads();
  int tid   = blockIdx.x * blockDim.x + threadIdx.x;
  int count = w.counters->count_ext[cur];
  // parked rays sort last and are not shaded in this iteration
  if (tid == 0) w.counters->count_shade = count - w.counters->class_count[wf_class_susp];
  if (tid >= count) return;
  int key = w.sort_key[tid];
  w.queue_shade[base[key >> 24] + (key & 0xffffff)] = w.queue_ext[cur][tid];
}

#if !defined(VPT_SHADE_BLOCKS)
#define VPT_SHADE_BLOCKS 1
#endif
template <int S>
__global__ void __launch_bounds__(wf_block, VPT_SHADE_BLOCKS)
    k_wf_shade(const DScene sc, const DParams prm, wf_view w, f4* __restrict__ image,
        rng_t* __restrict__ rngs, int cur) {
  int  tid      = blockIdx.x * blockDim.x + threadIdx.x;
  int  count    = w.counters->count_shade;
  bool to_ext   = false, to_pdf = false;
  int  slot     = 0;
  if (blockIdx.x == 0 && threadIdx.x < wf_classes) w.counters->class_count[threadIdx.x] = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    w.counters->ext_head = 0;
    w.counters->pdf_head = 0;
    w.counters->susp_count[cur] = 0;  // that pool was consumed by this iteration's extend
  }
  if (tid < count) {
    slot    = w.queue_shade[tid];
    int idx = tile_pixel_index(slot, prm.width, prm.height);
    path_t p;
    int    remaining;
    wf_load_path<S>(w, slot, p, remaining);
    auto a4 = w.at(wf_hit, slot), a5 = w.at(wf_hit1, slot);
    auto isec = isect_t{f2i(a5.y) != 0, f2i(a4.x), f2i(a4.y), {a4.z, a4.w}, a5.x};
    auto      rng     = wf_load_rng(w, slot);
    bool      was_vol = p.in_volume;
    pending_t pd;
    auto      r = shade_hit<S>(sc, prm, p, rng, isec, pd);
    if (r == step_lightpdf) {
      w.at(wf_mis, slot) = {pd.f.x, pd.f.y, pd.f.z, pd.bpdf};
      to_pdf             = true;
    } else {
      bool alive = r == step_skip;
      if (r == step_continue) alive = bounce_tail<S>(prm, p, rng);
      if (!alive) alive = wf_finish<S>(sc, prm, idx, p, rng, image, remaining);
      if (!alive) rngs[idx] = rng;  // slot done: hand the generator back to the state
      to_ext = alive;
    }
    wf_store_path<S>(w, slot, p, remaining, p.in_volume && !was_vol);
    wf_store_rng(w, slot, rng);
  }
  wf_push(w.queue_ext[cur ^ 1], &w.counters->count_ext[cur ^ 1], to_ext, slot);
  wf_push(w.queue_pdf, &w.counters->count_pdf, to_pdf, slot);
}

// Light pdf in two kernels. k_wf_lightpdf computes sample_lights_pdf along the slot's next ray
// (instance-targeted walks, sdf marches, environment texel pdf) and leaves the value in the
// slot's hit field; it carries no path state, so it fits in half the registers of the fused
// form (ncu r02a: 149 registers, 3 blocks per SM, 18 % of the warp slots, long-scoreboard
// bound). k_wf_pdf_tail then applies the MIS weight and runs bounce_tail / path regeneration.
#if !defined(VPT_PDF_BLOCKS)
#define VPT_PDF_BLOCKS 6
#endif
// ROOMY: as for k_wf_extend_persistent (96 instead of 80 registers; the instance walks of a
// triangle-mesh light gain the same way: 04_head1 837 -> 870, 02_rollingteapot 346 -> 353 Mpaths/s)
template <int S, int ROOMY>
__global__ void __launch_bounds__(wf_block, VPT_PDF_BLOCKS - ROOMY)
    k_wf_lightpdf(const DScene sc, const DParams prm, wf_view w, int cur) {
  extern __shared__ int stack_mem[];
  int tid   = blockIdx.x * blockDim.x + threadIdx.x;
  int count = w.counters->count_pdf;
  if (tid == 0) w.counters->rays_lightpdf += (unsigned long long)count;
  if (tid >= count) return;
  int     slot = w.queue_pdf[tid];
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  // after shade_hit the next ray is (position, incoming): the pdf is taken along it
  auto a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot);
  auto lpdf = sample_lights_pdf(sc, f3{a0.x, a0.y, a0.z}, f3{a0.w, a1.x, a1.y},
      prm.spheretrace_maxiter, st);
  w.at(wf_hit1, slot).z = lpdf;
}

// Persistent form of k_wf_lightpdf, same scheme as k_wf_extend_persistent: the grid is one
// resident set of warps, a warp claims queue entries for its idle lanes, and sample_lights_pdf
// (lights.cuh) runs as a per-lane state machine in rounds - `advance` walks the light list and
// sums the closed-form terms (environment pdf) until a light needs an instance-targeted walk
// or a single-field march, `walk` runs one trav_t::round of the current walk (up to 100 chained
// walks per mesh light), `march` a few sphere-tracing steps - so that a lane whose query is over
// takes the next queue entry instead of idling until the longest query of its warp ends
// (ncu r02g: the one-thread-per-ray kernel ran 13 of 32 lanes on 03_volume, 10.8 on
// 07_sdfunction). Per ray the arithmetic and its order are those of sample_lights_pdf.
constexpr int wf_pdf_march_steps = 8;
#if !defined(VPT_PDFP_BLOCKS)
#define VPT_PDFP_BLOCKS VPT_PDF_BLOCKS
#endif
template <int S>
__global__ void __launch_bounds__(wf_block, VPT_PDFP_BLOCKS)
    k_wf_lightpdf_persistent(const DScene sc, const DParams prm, wf_view w, int cur) {
  extern __shared__ int stack_mem[];
  constexpr bool sdf  = is_sdf_shader<S>();
  const unsigned full = 0xffffffffu;
  const int      count = w.counters->count_pdf;
  const int      lane  = threadIdx.x & 31;
  stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
  if (blockIdx.x == 0 && threadIdx.x == 0) w.counters->rays_lightpdf += (unsigned long long)count;
  enum : int { advance = 0, walk = 1, march = 2 };
  bool   has = false, more = count > 0, running = true;
  int    slot = 0, li = 0, bounce = 0, phase = advance, target = 0, mi = 0;
  float  pdf = 0, lpdf = 0, area = 0, mt = 0;
  f3     position{0, 0, 0}, direction{0, 0, 1}, next_position{0, 0, 0};
  trav_t t;
  t.cur = 0;
  while (running) {
    unsigned idle = __ballot_sync(full, !has);
    if (more && __popc(idle) >= (sdf ? wf_refill_sdf : wf_refill)) {
      int base = 0;
      if (lane == 0) base = atomicAdd(&w.counters->pdf_head, __popc(idle));
      base = __shfl_sync(full, base, 0);
      more = base + __popc(idle) < count;
      int mine = base + __popc(idle & ((1u << lane) - 1));
      if (!has && mine < count) {
        slot    = w.queue_pdf[mine];
        // after shade_hit the next ray is (position, incoming): the pdf is taken along it
        auto a0 = w.at(wf_ray0, slot), a1 = w.at(wf_ray1, slot);
        position = {a0.x, a0.y, a0.z}, direction = {a0.w, a1.x, a1.y};
        pdf = 0, li = 0, phase = advance, has = true;
      }
    }
    if (has && phase == advance) {
      while (li < sc.num_lights && phase == advance) {
        auto l0 = ld4(sc.lights + 2 * li), l1 = ld4(sc.lights + 2 * li + 1);
        if (l0.x != invalidid) {
          area = ld1(sc.cdf + l0.w + l1.x - 1);
          lpdf = 0, next_position = position, bounce = 0, target = l0.x;
          t.begin(sc, target, make_ray(next_position, direction), st);
          phase = walk;
        } else if (l0.z != invalidid) {
          area = ld1(sc.cdf + l0.w + l1.x - 1);
          target = l0.z, mt = 1e-4f, mi = 0;  // make_ray's tmin
          phase  = march;
        } else {
          if (l0.y != invalidid) {
            auto e  = sc.environments + 9 * l0.y;
            auto em = ld4(e + 8);
            if (f2i(em.w) != invalidid) {
              auto tex      = ld4(sc.textures + f2i(em.w));
              auto wl       = transform_direction(ld_frame(e + 4), direction);
              auto texcoord = environment_texcoord(wl);
              auto i    = vclamp((int)(texcoord.x * tex.x), 0, tex.x - 1);
              auto j    = vclamp((int)(texcoord.y * tex.y), 0, tex.y - 1);
              auto prob = sample_discrete_pdf(sc.cdf + l0.w, j * tex.x + i) / ld1(sc.cdf + l0.w + l1.x - 1);
              auto angle = (2 * pif / tex.x) * (pif / tex.y) * vsin(pif * (j + 0.5f) / tex.y);
              pdf += prob / angle;
            } else {
              pdf += 1 / (4 * pif);
            }
          }
          li++;
        }
      }
    }
    if (has && phase == walk) {
      auto ray = make_ray(next_position, direction);
      if (!t.done()) t.round(sc, ray, st);
      if (t.done()) {
        if (t.h.hit) {
          auto in        = ld_instance(sc, target);
          auto lposition = eval_position(sc, in, t.h.element, t.h.uv);
          auto lnormal   = eval_element_normal(sc, in, t.h.element);
          lpdf += distance_squared(lposition, position) / (vabs(dot(lnormal, direction)) * area);
          next_position = lposition + direction * 1e-3f;
          bounce++;
        }
        if (!t.h.hit || bounce == 100) {
          pdf += lpdf;
          li++, phase = advance;
        } else {
          t.begin(sc, target, make_ray(next_position, direction), st);
        }
      }
    }
    if (has && phase == march) {
      auto ray = make_ray(position, direction);
      auto s   = ld_sdf(sc, target);
      bool fin = false, hit = false;
      for (int k = 0; k < wf_pdf_march_steps && !fin; k++) {
        if (mi < prm.spheretrace_maxiter && mt < ray.tmax) {
          auto res = eval_sdf_fn(s, transform_point(s.frame, ray_point(ray, mt)));
          if (vabs(res) < (flt_eps * mt)) hit = true, fin = true;
          else mt += res, ++mi;
        } else {
          fin = true;
        }
      }
      if (fin) {
        if (hit) {
          auto lposition = ray_point(ray, mt);
          // the reference evaluates the normal at the shading position (quirk kept)
          auto lnormal = eval_sdf_normal_fn(s, position, mt);
          pdf += distance_squared(lposition, position) / (vabs(dot(lnormal, direction)) * area);
        }
        li++, phase = advance;
      }
    }
    if (has && phase == advance && li >= sc.num_lights) {
      w.at(wf_hit1, slot).z = sc.num_lights > 0 ? pdf * ((float)1 / (float)sc.num_lights) : 0.0f;
      has = false;
    }
    running = more || __any_sync(full, has);
  }
}

template <int S>
__global__ void __launch_bounds__(wf_block, VPT_SHADE_BLOCKS)
    k_wf_pdf_tail(const DScene sc, const DParams prm, wf_view w, f4* __restrict__ image,
        rng_t* __restrict__ rngs, int cur) {
  int  tid    = blockIdx.x * blockDim.x + threadIdx.x;
  int  count  = w.counters->count_pdf;
  bool to_ext = false;
  int  slot   = 0;
  if (tid == 0) w.counters->count_ext[cur] = 0;  // consumed; the next iteration fills it
  if (tid < count) {
    slot    = w.queue_pdf[tid];
    int idx = tile_pixel_index(slot, prm.width, prm.height);
    path_t  p;
    int     remaining;
    wf_load_path<S>(w, slot, p, remaining);
    auto a5   = w.at(wf_mis, slot);
    auto pd   = pending_t{p.o, p.d, {a5.x, a5.y, a5.z}, a5.w};
    auto lpdf = w.at(wf_hit1, slot).z;
    apply_lightpdf(p, pd, lpdf);
    auto rng   = wf_load_rng(w, slot);
    bool alive = bounce_tail<S>(prm, p, rng);
    if (!alive) alive = wf_finish<S>(sc, prm, idx, p, rng, image, remaining);
    if (!alive) rngs[idx] = rng;
    to_ext = alive;
    wf_store_path<S>(w, slot, p, remaining, false);
    wf_store_rng(w, slot, rng);
  }
  wf_push(w.queue_ext[cur ^ 1], &w.counters->count_ext[cur ^ 1], to_ext, slot);
}

// Drain: below `wf_drain_threshold()` active slots an iteration no longer fills
// the machine and its cost is the latency of three dependent launches, so the
// remaining slots are finished by ONE launch in which each thread runs its slot
// to the end (rest of the current path, then the pixel's remaining samples).
// Same state machine, same order of arithmetic and RNG draws per pixel.
template <int S>
__global__ void __launch_bounds__(wf_block)
    k_wf_drain(const DScene sc, const DParams prm, wf_view w, f4* __restrict__ image,
        rng_t* __restrict__ rngs, int cur) {
  extern __shared__ int stack_mem[];
  int tid   = blockIdx.x * blockDim.x + threadIdx.x;
  int count = w.counters->count_ext[cur];
  unsigned long long n_ext = 0, n_pdf = 0;
  if (tid < count) {
    int     slot = w.queue_ext[cur][tid];
    int     idx  = tile_pixel_index(slot, prm.width, prm.height);
    stack_t st{stack_mem + vpt_stack_words * threadIdx.x, (int)blockDim.x};
    path_t  p;
    int     remaining;
    wf_load_path<S>(w, slot, p, remaining);
    auto rng   = wf_load_rng(w, slot);
    auto alive = true;
    bool parked = (f2i(w.at(wf_ray1, slot).z) & wf_bit_susp) != 0;
    while (alive) {  // single exit
      isect_t isec;
      if (parked) {
        // the slot's ray was parked by the last extend launch: finish that walk
        auto ray = make_ray(p.o, p.d);
        if constexpr (is_sdf_shader<S>()) {
          sphere_t sp;
          sp.begin(ray);
          auto s = w.at(wf_hit, slot);
          sp.t = s.x, sp.i = f2i(s.y);
          while (!sp.done()) sp.step(sc, ray, prm.spheretrace_maxiter);
          isec = {sp.r.hit, sp.r.instance, sp.r.sdf, {0, 0}, sp.r.dist};
        } else {
          trav_t t;
          wf_resume_ray(w.susp[cur] + (size_t)f2i(w.at(wf_hit, slot).x) * w.susp_f4, t, st);
          while (!t.done()) t.round(sc, ray, st);
          isec = {t.h.hit, t.h.instance, t.h.element, t.h.uv, t.h.distance};
        }
        parked = false;
        n_ext -= 1;  // counted when it was first queued
      } else {
        isec = find_hit<S>(sc, prm, p, st);
      }
      pending_t pd;
      auto      r = shade_hit<S>(sc, prm, p, rng, isec, pd);
      n_ext += 1;
      if (r == step_lightpdf) {
        auto lpdf = sample_lights_pdf(sc, pd.position, pd.incoming, prm.spheretrace_maxiter, st);
        apply_lightpdf(p, pd, lpdf);
        n_pdf += 1;
      }
      alive = r == step_skip;
      if (r == step_continue || r == step_lightpdf) alive = bounce_tail<S>(prm, p, rng);
      if (!alive) alive = wf_finish<S>(sc, prm, idx, p, rng, image, remaining);
    }
    rngs[idx] = rng;
  }
  for (int o = 16; o > 0; o >>= 1) {
    n_ext += __shfl_down_sync(0xffffffffu, n_ext, o);
    n_pdf += __shfl_down_sync(0xffffffffu, n_pdf, o);
  }
  if ((threadIdx.x & 31) == 0 && n_ext) {
    atomicAdd(&w.counters->rays_extend, n_ext);
    atomicAdd(&w.counters->rays_lightpdf, n_pdf);
  }
}

// active slots at or below which the wavefront hands over to k_wf_drain;
// VPT_WF_DRAIN overrides (0 disables), for experiments and the parity tests
inline int wf_drain_threshold(int slots) {
  const char* env = getenv("VPT_WF_DRAIN");
  if (env) return atoi(env);
  int t = slots / 8;
  return t < 32768 ? t : 32768;
}

// VPT_WF_TRACE=<file>: per-iteration kernel times from CUDA events (a debugging
// aid; the events perturb the run, so a traced call is not a bench value)
struct wf_trace {
  std::vector<cudaEvent_t> events;
  std::vector<int>         grids;
  const char*              path;
  static std::unique_ptr<wf_trace> open() {
    const char* env = getenv("VPT_WF_TRACE");
    if (!env || !*env) return nullptr;
    auto t  = std::make_unique<wf_trace>();
    t->path = env;
    return t;
  }
  void mark(int grid) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e);
    events.push_back(e);
    grids.push_back(grid);
  }
  void dump() {
    cudaDeviceSynchronize();
    if (auto f = fopen(path, "a")) {
      fprintf(f, "iteration,grid,extend_us,shade_us,lightpdf_us\n");
      for (size_t i = 0; i + 3 < events.size(); i += 4) {
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, events[i], events[i + 1]);
        cudaEventElapsedTime(&b, events[i + 1], events[i + 2]);
        cudaEventElapsedTime(&c, events[i + 2], events[i + 3]);
        fprintf(f, "%zu,%d,%.1f,%.1f,%.1f\n", i / 4, grids[i], a * 1e3f, b * 1e3f, c * 1e3f);
      }
      fclose(f);
    }
    for (auto e : events) cudaEventDestroy(e);
  }
};

// ---- host driver ---------------------------------------------------------------
template <int S>
cudaError_t run_wavefront_shader(const flat_scene& scene, wavefront_buffers& wf,
    const DParams& prm, f4* image, rng_t* rngs, int nsamples, int* launches,
    vpt_trace_stats* stats) {
  int slots = tile_slot_count(prm.width, prm.height);
  if (slots >= (1 << 24)) return cudaErrorInvalidValue;  // sort_key packs the rank in 24 bits
  auto trace = wf_trace::open();
  // pipes: VPT_WF_PIPES overrides; small images, traced runs and sphere tracing (whose extend
  // keeps the issue slots 88 % busy on its own: B200, 07_sdfunction 140 vs 136 Mpaths/s) use one
  const char* pipes_env = getenv("VPT_WF_PIPES");  // read per call: bench.py samples one pipe alone
  // by image size (B200, 64 / 16 spp calls): a 1.5 M-pixel wavefront fills the machine on its own
  // (05_head1ss 186.7 ms with one pipe, 194.6 with two; 03_volume at 1920x800 87.0 vs 89.0), around
  // 0.2 M pixels a third pipe still hides latency (02_rollingteapot 37.4 vs 38.6 ms, 01_surface
  // 49.7 vs 51.0), in between two (03_volume 1280x533: 171.0 ms with two, 177.9 / 175.4 with one / three)
  int by_size = slots >= wf_pipes_one_slots ? 1 : (slots >= wf_pipes_two_slots ? 2 : 3);
  int npipes = pipes_env ? atoi(pipes_env)
                         : (slots >= wf_pipes_min_slots && !is_sdf_shader<S>() ? by_size : 1);
  npipes     = npipes < 1 ? 1 : (npipes > wf_max_pipes ? wf_max_pipes : npipes);
  if (trace) npipes = 1;
  auto e = wf.reserve(slots, npipes);
  if (e != cudaSuccess) return e;
  size_t smem = (size_t)scene.stack_depth * vpt_stack_words * wf_block * sizeof(int);
  // the extend build: see k_wf_extend_persistent (VPT_WF_ROOMY=0 / 1 overrides the scene rule)
  auto ext_kernel = k_wf_extend_persistent<S, 0>;
  auto pdf_kernel = k_wf_lightpdf<S, 0>;
  int  roomy      = 0;
  if constexpr (!is_sdf_shader<S>()) {
    const char* roomy_env = getenv("VPT_WF_ROOMY");  // read per call
    roomy = roomy_env ? (atoi(roomy_env) != 0) : scene.num_triangles >= wf_roomy_triangles;
    if (roomy) ext_kernel = k_wf_extend_persistent<S, 1>, pdf_kernel = k_wf_lightpdf<S, 1>;
  }
  if (smem > 48 * 1024) {
    e = cudaFuncSetAttribute(k_wf_extend<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(ext_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(pdf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(k_wf_drain<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  // persistent extend: as many blocks as the device keeps resident
  bool persistent = !getenv("VPT_WF_NO_PERSISTENT");
  int  resident   = 0;
  if (persistent) {
    int per_sm = 0, device = 0, sms = 0;
    cudaGetDevice(&device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ext_kernel, wf_block, smem);
    if (e != cudaSuccess) return e;
    resident = per_sm * sms;
    // VPT_WF_EXT_BLOCKS=<n>: resident extend blocks per SM (experiments: leaves room for the
    // other pipe's kernels next to a running extend)
    const char* blocks_env = getenv("VPT_WF_EXT_BLOCKS");  // read per call
    if (blocks_env && atoi(blocks_env) > 0 && atoi(blocks_env) < per_sm) resident = atoi(blocks_env) * sms;
    if (resident <= 0) persistent = false;
  }
  // pooled extend (mesh shaders; VPT_WF_POOL=0 keeps the ray-per-lane kernel): shared memory
  // per block = ray state + as many stack entries as fit, the rest of the depth bound spills
  bool pooled = persistent && !is_sdf_shader<S>() && !trace;
  {
    const char* pool_env = getenv("VPT_WF_POOL");  // read per call (tests run both kernels)
    if (!pool_env || atoi(pool_env) == 0) pooled = false;
  }
  int    pool_limit = 0, pool_resident = 0;
  size_t pool_smem  = 0;
  if (pooled) {
    constexpr int BR = wf_block * wf_pool_k;
    int device = 0, sms = 0, smem_sm = 0;
    cudaGetDevice(&device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device);
    const char* pb_env = getenv("VPT_WF_POOL_BLOCKS");  // resident blocks per SM to size for
    const char* pl_env = getenv("VPT_WF_POOL_LIMIT");   // tests: stack entries kept in shared memory
    int    want   = pb_env && atoi(pb_env) > 0 ? atoi(pb_env) : VPT_POOL_BLOCKS;
    size_t fixed  = sizeof(int) * pf_fields * BR + 32 * wf_pool_k * (wf_block / 32);
    size_t budget = (size_t)smem_sm / want - 1024 - 256;
    pool_limit    = budget > fixed ? (int)((budget - fixed) / (sizeof(int) * 2 * BR)) : 0;
    if (pool_limit > scene.stack_depth) pool_limit = scene.stack_depth;
    if (pl_env && atoi(pl_env) > 0 && atoi(pl_env) < pool_limit) pool_limit = atoi(pl_env);
    if (pool_limit < 1) pooled = false;
    if (pooled) {
      pool_smem = fixed + sizeof(int) * 2 * BR * (size_t)pool_limit;
      e = cudaFuncSetAttribute(k_wf_extend_pool<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pool_smem);
      if (e != cudaSuccess) return e;
      int per_sm = 0;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wf_extend_pool<S>, wf_block, pool_smem);
      if (e != cudaSuccess) return e;
      pool_resident = per_sm * sms;
      if (pool_resident <= 0) pooled = false;
    }
    if (pooled) {
      int spill = scene.stack_depth - pool_limit;
      e = wf.reserve_suspended(pool_resident * BR, spill > 0 ? (spill + 1) / 2 : 1);
      if (e != cudaSuccess) return e;
    }
  }
  // persistent light pdf: opt-in (VPT_WF_PDF_PERSISTENT=1). Measured on B200, bit-identical and
  // slower than one thread per queued ray on every scene (03_volume 181.3 vs 170.6 ms per 64 spp,
  // 07_sdfunction 140.5 vs 136.5 ms per 16 spp; 96 / 128 registers: 177.6 / 180.5 ms): the state
  // machine's bookkeeping and spills cost more than the refilled lanes give back
  int pdf_resident = 0;
  {
    const char* pp_env = getenv("VPT_WF_PDF_PERSISTENT");  // read per call (tests run both kernels)
    if (persistent && pp_env && atoi(pp_env) != 0) {
      int per_sm = 0, device = 0, sms = 0;
      cudaGetDevice(&device);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
      if (smem > 48 * 1024) {
        e = cudaFuncSetAttribute(k_wf_lightpdf_persistent<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
      }
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_wf_lightpdf_persistent<S>, wf_block, smem);
      if (e != cudaSuccess) return e;
      pdf_resident = per_sm * sms;
    }
  }
  // lanes still walking at or below which a warp parks its rays once the queue is empty
  const char* susp_env = getenv("VPT_WF_SUSPEND");  // read per call (tests sweep it)
  int suspend_at = susp_env ? atoi(susp_env) : wf_suspend;
  // parked BVH walks need record pools (sphere tracing parks into the slot record itself)
#if defined(VPT_WF_SUSPEND_MESH)
  if (persistent && !is_sdf_shader<S>()) {
    e = wf.reserve_suspended(resident * wf_block, wf_susp_record_f4(scene.stack_depth));
    if (e != cudaSuccess) return e;
  }
#endif
  *launches = 0;
  int graph_count = 0, *graph_launches = &graph_count;
  // the first extend launch of every batch of pipe 0 is timed with CUDA events: its ray count
  // is known exactly, so the pairs give the dominant kernel's rays per second (with two
  // pipes the other pipe's kernels share the device during that time, as they do in the run)
  double             top_ms = 0;
  unsigned long long top_rays = 0;
  int                top_launches = 0, timed_active = 0, timed_batch = 0;
  bool               timed_pending = false;

  for (int i = 0; i < npipes; i++) {
    auto& p  = wf.pipe[i];
    int   ps = wavefront_buffers::pipe_slots(slots, npipes, i);
    e = cudaMemsetAsync(p.view.counters, 0, sizeof(wf_counters), p.stream);
    if (e != cudaSuccess) return e;
    if (ps > 0)
      k_wf_init<S><<<(ps + wf_block - 1) / wf_block, wf_block, 0, p.stream>>>(scene.d, prm, p.view,
          rngs, slots, nsamples, i, npipes);
    *launches += 1;
    p.active = ps, p.cur = 0, p.inflight = 0, p.batches = 0, p.draining = false;
  }
  int drain_at = wf_drain_threshold(slots) / npipes;

  // one batch = wf_batch iterations, then a snapshot of the counters; two batches per pipe
  // are kept in flight so that a stream never runs dry while the host reads a snapshot
  // the 5 kernels of one iteration at grid `grid` (extend: at most the resident set)
  auto iteration = [&](wf_pipe& p, int grid, bool timed) {
    auto& w = p.view;
    if (timed) cudaEventRecord(wf.k0, p.stream);
    if (trace) trace->mark(grid);
    if (pooled) {
      int blocks = (grid + wf_pool_k - 1) / wf_pool_k;
      k_wf_extend_pool<S><<<blocks < pool_resident ? blocks : pool_resident, wf_block, pool_smem, p.stream>>>(
          scene.d, prm, w, p.cur, pool_limit);
    } else if (persistent)
      ext_kernel<<<grid < resident ? grid : resident, wf_block, smem, p.stream>>>(scene.d, prm, w, p.cur,
          suspend_at);
    else
      k_wf_extend<S><<<grid, wf_block, smem, p.stream>>>(scene.d, prm, w, p.cur);
    if (timed) cudaEventRecord(wf.k1, p.stream);
    k_wf_sort<<<grid, wf_block, 0, p.stream>>>(w, p.cur);
    if (trace) trace->mark(grid);
    k_wf_shade<S><<<grid, wf_block, 0, p.stream>>>(scene.d, prm, w, image, rngs, p.cur);
    if (trace) trace->mark(grid);
    if (pdf_resident > 0)
      k_wf_lightpdf_persistent<S><<<grid < pdf_resident ? grid : pdf_resident, wf_block, smem, p.stream>>>(
          scene.d, prm, w, p.cur);
    else
      pdf_kernel<<<grid, wf_block, smem, p.stream>>>(scene.d, prm, w, p.cur);
    k_wf_pdf_tail<S><<<grid, wf_block, 0, p.stream>>>(scene.d, prm, w, image, rngs, p.cur);
    if (trace) trace->mark(grid);
    p.cur ^= 1;
  };
  // A batch as ONE graph launch. Every kernel reads its true work count from device memory
  // and a block without work exits at once, so a graph captured at a grid no smaller than the
  // batch needs can be replayed for any batch (of this call and of later calls on the same
  // scene / state). Graphs are captured lazily for grids of 1, 1/2, ... 1/64 of the pipe's
  // slots and a batch takes the smallest that covers its active slots (B200: with the full
  // grid alone the ~200 tail iterations of 03_volume paid 3 ms for empty blocks). The host
  // submits 1 graph + 1 counter copy per batch instead of 40 kernels. The batch that
  // carries the timed extend launch (one in wf_timed_every on pipe 0) is enqueued kernel by
  // kernel, as before. wf_batch is even, so every batch starts on queue parity 0.
  static_assert(wf_batch % 2 == 0, "a graph assumes every batch starts on the same queue parity");
  auto batch_graph = [&](int i, int need) -> cudaGraphExec_t {
    auto& p = wf.pipe[i];
    struct key_t {
      DScene d; DParams prm; wf_view view; f4* image; rng_t* rngs;
      int suspend_at, resident, grid, shader, persistent; size_t smem;
    } key;
    memset(&key, 0, sizeof(key));
    int full = (wavefront_buffers::pipe_slots(slots, npipes, i) + wf_block - 1) / wf_block;
    int cls  = 0;
    while (cls + 1 < wf_pipe::graph_classes && ((full + (2 << cls) - 1) >> (cls + 1)) >= need) cls++;
    full = (full + (1 << cls) - 1) >> cls;
    if (full < need) return nullptr;
    key.d = scene.d, key.prm = prm, key.view = p.view, key.image = image, key.rngs = rngs;
    key.suspend_at = pooled ? pool_limit : suspend_at, key.resident = pooled ? pool_resident : resident;
    key.grid = full, key.shader = S;
    key.persistent = (persistent ? 1 : 0) | (pooled ? 2 : 0) | (roomy ? 4 : 0) | (pdf_resident << 3), key.smem = pooled ? pool_smem : smem;
    auto& exec = p.graph[cls];
    auto& kept = p.graph_key[cls];
    if (exec && kept.size() == sizeof(key) && !memcmp(kept.data(), &key, sizeof(key))) return exec;
    if (exec) cudaGraphExecDestroy(exec), exec = nullptr;
    if (p.cur != 0) return nullptr;
    cudaGraph_t g = nullptr;
    if (cudaStreamBeginCapture(p.stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    for (int it = 0; it < wf_batch; it++) iteration(p, full, false);
    auto e = cudaStreamEndCapture(p.stream, &g);
    if (e == cudaSuccess) e = cudaGraphInstantiate(&exec, g, 0);
    if (g) cudaGraphDestroy(g);
    if (e != cudaSuccess) {
      cudaGetLastError();
      exec = nullptr;
      return nullptr;
    }
    kept.assign((const char*)&key, (const char*)&key + sizeof(key));
    return exec;
  };
  const char* graph_env = getenv("VPT_WF_GRAPH");  // 0: kernel-by-kernel submission everywhere
  bool use_graph = !(graph_env && atoi(graph_env) == 0) && !trace;

  // one batch = wf_batch iterations, then a snapshot of the counters; two batches per pipe
  // are kept in flight so that a stream never runs dry while the host reads a snapshot
  auto enqueue = [&](int i) -> cudaError_t {
    auto& p    = wf.pipe[i];
    auto& w    = p.view;
    int   grid = (p.active + wf_block - 1) / wf_block;
    bool  timed = i == 0 && !timed_pending && (!use_graph || p.batches % wf_timed_every == 0);
    auto  exec  = use_graph && !timed && p.cur == 0 ? batch_graph(i, grid) : nullptr;
    if (exec) {
      auto e = cudaGraphLaunch(exec, p.stream);
      if (e != cudaSuccess) return e;
      *graph_launches += 1;
    } else {
      for (int it = 0; it < wf_batch; it++) iteration(p, grid, it == 0 && timed);
    }
    *launches += 5 * wf_batch;
    if (timed) timed_pending = true, timed_active = p.active, timed_batch = p.batches;
    int  k = p.batches & 1;
    auto e = cudaMemcpyAsync(p.host_counters[k], w.counters, sizeof(wf_counters),
        cudaMemcpyDeviceToHost, p.stream);
    if (e == cudaSuccess) e = cudaEventRecord(p.poll[k], p.stream);
    p.polled_cur[k] = p.cur;
    p.batches += 1, p.inflight += 1;
    return e;
  };
  auto drain = [&](int i) -> cudaError_t {
    auto& p    = wf.pipe[i];
    int   grid = (p.active + wf_block - 1) / wf_block;
    k_wf_drain<S><<<grid, wf_block, smem, p.stream>>>(scene.d, prm, p.view, image, rngs, p.cur);
    *launches += 1;
    int  k = p.batches & 1;
    auto e = cudaMemcpyAsync(p.host_counters[k], p.view.counters, sizeof(wf_counters),
        cudaMemcpyDeviceToHost, p.stream);
    if (e == cudaSuccess) e = cudaEventRecord(p.poll[k], p.stream);
    p.batches += 1, p.inflight += 1;
    p.draining = true;
    return e;
  };
  int  depth = npipes > 1 ? 2 : 1;
  bool busy  = true;
  while (busy) {
    busy = false;
    for (int i = 0; i < npipes; i++) {
      auto& p = wf.pipe[i];
      if (p.active <= 0 || p.draining) continue;
      if (p.active <= drain_at) {
        // few slots left: once the batches in flight are done, one launch finishes them
        if (p.inflight == 0) e = drain(i);
      } else {
        while (e == cudaSuccess && p.inflight < depth) e = enqueue(i);
      }
      if (e != cudaSuccess) return e;
    }
    for (int i = 0; i < npipes; i++) {
      auto& p = wf.pipe[i];
      if (p.inflight == 0) continue;
      busy       = true;
      int batch  = p.batches - p.inflight;  // oldest batch in flight
      int k      = batch & 1;
      e          = cudaEventSynchronize(p.poll[k]);
      if (e != cudaSuccess) return e;
      p.inflight -= 1;
      if (p.draining) {
        p.active = 0;
        continue;
      }
      if (i == 0 && timed_pending && batch == timed_batch) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, wf.k0, wf.k1) == cudaSuccess)
          top_ms += ms, top_rays += (unsigned long long)timed_active, top_launches += 1;
        timed_pending = false;
      }
      int count = p.host_counters[k]->count_ext[p.polled_cur[k]];
      p.active  = count < p.active ? count : p.active;
    }
  }
  if (trace) trace->dump();
  stats->top_kernel_ms       = (float)top_ms;
  stats->top_kernel_launches = top_launches;
  stats->top_kernel_units    = top_rays;
  stats->graph_launches      = (uint64_t)graph_count;
  stats->rays_extend = stats->rays_lightpdf = 0;
  for (int i = 0; i < npipes; i++) {
    if (wf.pipe[i].batches == 0) continue;  // a pipe without slots never wrote its counters
    auto c = wf.pipe[i].host_counters[(wf.pipe[i].batches - 1) & 1];
    stats->rays_extend += c->rays_extend - c->rays_resumed;
    stats->rays_lightpdf += c->rays_lightpdf;
  }
  return cudaGetLastError();
}

inline cudaError_t run_wavefront(const flat_scene& scene, wavefront_buffers& wf,
    const DParams& prm, f4* image, rng_t* rngs, int nsamples, int* launches,
    vpt_trace_stats* stats) {
  switch (prm.shader) {
    case shader_volpathtrace:
      return run_wavefront_shader<shader_volpathtrace>(scene, wf, prm, image, rngs, nsamples,
          launches, stats);
    case shader_pathtrace:
      return run_wavefront_shader<shader_pathtrace>(scene, wf, prm, image, rngs, nsamples,
          launches, stats);
    case shader_naive:
      return run_wavefront_shader<shader_naive>(scene, wf, prm, image, rngs, nsamples, launches,
          stats);
    case shader_implicit:
      return run_wavefront_shader<shader_implicit>(scene, wf, prm, image, rngs, nsamples,
          launches, stats);
    default: return cudaErrorNotSupported;
  }
}

#endif  // __CUDACC__

}  // namespace vpt
