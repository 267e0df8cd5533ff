// bvh_build.cuh — shape BVH construction on the device.
//
// The reference builds every shape BVH on the host, top-down, splitting at the middle of the
// centroid box's largest axis with <= 4 primitives per leaf (make_bvh / build_bvh / split_middle,
// yocto_bvh.cpp:411-571); host/vpt_bvh.cpp reproduces that tree byte for byte and stays the
// default, because only THAT tree gives the reference's visit order (bit-exact ties).
// This is the optional device builder behind the same first-hit gate: an LBVH
//   1. primitive boxes + centroids, centroid bounds (one pass, block reduction + ordered atomics)
//   2. 30-bit Morton code of every centroid, key = code << 32 | primitive (distinct keys)
//   3. cub::DeviceRadixSort on the 64-bit keys
//   4. Karras' hierarchy (HPG 2012): one thread per internal node finds its range and split
//   5. boxes bottom-up (the second thread to arrive at a node carries on)
//   6. subtrees of <= 4 primitives become leaves (the reference's bvh_max_prims), the rest is
//      renumbered so that siblings are adjacent (bvh_node::start / start + 1, yocto_bvh.h:73-79)
// and the result is written in the reference's own node format, so it goes through
// vpt_cuda_upload_scene (wide collapse included) like a host-built tree. A different tree finds
// the same closest distance; which of two primitives at EXACTLY the same distance is reported
// may differ (the reference keeps the later-tested one) - tests record those rays.
#pragma once

#include <cuda_runtime.h>

#include <cub/cub.cuh>

#include "../../include/vpt_cuda.h"
#include "vmath.cuh"

namespace vpt {

constexpr int lbvh_block     = 256;
constexpr int lbvh_leaf_size = 4;  // bvh_max_prims, yocto_bvh.cpp:444

struct lbvh_box {
  float lo[3], hi[3];
};

// order-preserving map float -> uint for atomicMin / atomicMax
__device__ __forceinline__ unsigned lbvh_ordered(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float lbvh_unordered(unsigned u) {
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// 1. boxes, centroids, centroid bounds. bounds[0..2] = min (ordered), [3..5] = max
__global__ void __launch_bounds__(lbvh_block)
    k_lbvh_boxes(const float* __restrict__ positions, const int* __restrict__ elems, int verts_per,
        int n, lbvh_box* __restrict__ boxes, float* __restrict__ centroids, unsigned* bounds) {
  __shared__ unsigned smin[3], smax[3];
  if (threadIdx.x < 3) smin[threadIdx.x] = 0xffffffffu, smax[threadIdx.x] = 0;
  __syncthreads();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    lbvh_box b;
    for (int k = 0; k < 3; k++) b.lo[k] = flt_max, b.hi[k] = -flt_max;
    for (int v = 0; v < verts_per; v++) {
      auto p = positions + 3 * (size_t)elems[(size_t)verts_per * i + v];
      for (int k = 0; k < 3; k++) b.lo[k] = fminf(b.lo[k], p[k]), b.hi[k] = fmaxf(b.hi[k], p[k]);
    }
    boxes[i] = b;
    for (int k = 0; k < 3; k++) {
      float c                       = (b.lo[k] + b.hi[k]) / 2;  // center(bbox), yocto_bvh.cpp:456
      centroids[3 * (size_t)i + k] = c;
      atomicMin(&smin[k], lbvh_ordered(c));
      atomicMax(&smax[k], lbvh_ordered(c));
    }
  }
  __syncthreads();
  if (threadIdx.x < 3) {
    atomicMin(&bounds[threadIdx.x], smin[threadIdx.x]);
    atomicMax(&bounds[3 + threadIdx.x], smax[threadIdx.x]);
  }
}

__device__ __forceinline__ unsigned lbvh_spread(unsigned v) {  // 10 bits -> every third bit
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

// 2. keys: x in bits 29, 26, ...; y in 28, 25, ...; z in 27, 24, ...
__global__ void __launch_bounds__(lbvh_block)
    k_lbvh_keys(const float* __restrict__ centroids, const unsigned* __restrict__ bounds, int n,
        unsigned long long* __restrict__ keys, int* __restrict__ ids) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned code = 0;
  for (int k = 0; k < 3; k++) {
    float lo = lbvh_unordered(bounds[k]), hi = lbvh_unordered(bounds[3 + k]);
    float ext = hi - lo;
    float u   = ext > 0 ? (centroids[3 * (size_t)i + k] - lo) / ext : 0.0f;
    auto  q   = (unsigned)fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
    code |= lbvh_spread(q) << (2 - k);
  }
  keys[i] = ((unsigned long long)code << 32) | (unsigned)i;
  ids[i]  = i;
}

// child reference inside the builder: index, top bit set for a leaf (sorted position)
constexpr int lbvh_leaf_bit = 1 << 30;

__device__ __forceinline__ int lbvh_delta(const unsigned long long* keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  return __clzll(keys[i] ^ keys[j]);  // keys are distinct: never 64
}

// 4. Karras 2012, figure 4: range and split of internal node i of n - 1
__global__ void __launch_bounds__(lbvh_block)
    k_lbvh_hierarchy(const unsigned long long* __restrict__ keys, int n, int2* __restrict__ children,
        int2* __restrict__ ranges, int* __restrict__ parent_internal, int* __restrict__ parent_leaf,
        signed char* __restrict__ axis) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  int d     = lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1) > 0 ? 1 : -1;
  int dmin  = lbvh_delta(keys, n, i, i - d);
  int lmax  = 2;
  while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2)
    if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  int j     = i + l * d;
  int dnode = lbvh_delta(keys, n, i, j);
  int s     = 0;
  for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
    if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    if (t == 1) break;
  }
  int gamma = i + s * d + (d < 0 ? -1 : 0);
  int first = i < j ? i : j, last = i < j ? j : i;
  int left  = first == gamma ? (gamma | lbvh_leaf_bit) : gamma;
  int right = last == gamma + 1 ? ((gamma + 1) | lbvh_leaf_bit) : gamma + 1;
  children[i] = make_int2(left, right);
  ranges[i]   = make_int2(first, last);
  if (left & lbvh_leaf_bit) parent_leaf[gamma] = i; else parent_internal[gamma] = i;
  if (right & lbvh_leaf_bit) parent_leaf[gamma + 1] = i; else parent_internal[gamma + 1] = i;
  // split axis = axis of the highest Morton bit in which the two sides differ (x: bits 29, 26, ..)
  int bit = 63 - __clzll(keys[gamma] ^ keys[gamma + 1]);  // 32 .. 61 for differing codes
  axis[i] = bit >= 32 ? (signed char)((61 - bit) % 3) : (signed char)0;
}

// a box another block may have written a moment ago: read around L1 (a neighbouring box of the
// same line may already sit there from before the write)
__device__ __forceinline__ lbvh_box lbvh_load_box(const lbvh_box* p) {
  lbvh_box b;
  auto     f = reinterpret_cast<const float*>(p);
  for (int k = 0; k < 3; k++) b.lo[k] = __ldcg(f + k), b.hi[k] = __ldcg(f + 3 + k);
  return b;
}

// 5. boxes bottom-up
__global__ void __launch_bounds__(lbvh_block)
    k_lbvh_fit(const lbvh_box* __restrict__ prim_boxes, const int* __restrict__ sorted_ids, int n,
        const int2* __restrict__ children, const int* __restrict__ parent_internal,
        const int* __restrict__ parent_leaf, lbvh_box* __restrict__ node_boxes,
        lbvh_box* __restrict__ leaf_boxes, int* __restrict__ arrived) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  leaf_boxes[i] = prim_boxes[sorted_ids[i]];
  if (n == 1) return;
  __threadfence();
  int node = parent_leaf[i];
  while (node >= 0) {
    if (atomicAdd(&arrived[node], 1) == 0) return;  // the sibling's thread fits this node
    __threadfence();
    auto c = children[node];
    auto a = lbvh_load_box((c.x & lbvh_leaf_bit) ? leaf_boxes + (c.x & ~lbvh_leaf_bit) : node_boxes + c.x);
    auto b = lbvh_load_box((c.y & lbvh_leaf_bit) ? leaf_boxes + (c.y & ~lbvh_leaf_bit) : node_boxes + c.y);
    lbvh_box m;
    for (int k = 0; k < 3; k++) m.lo[k] = fminf(a.lo[k], b.lo[k]), m.hi[k] = fmaxf(a.hi[k], b.hi[k]);
    node_boxes[node] = m;
    __threadfence();
    node = node == 0 ? -1 : parent_internal[node];
  }
}

// 6a. an internal node stays internal iff it holds more than lbvh_leaf_size primitives (then
// so do all its ancestors: the flag alone says whether the node is in the final tree)
__global__ void __launch_bounds__(lbvh_block)
    k_lbvh_alive(const int2* __restrict__ ranges, int n, int* __restrict__ alive) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  alive[i] = ranges[i].y - ranges[i].x + 1 > lbvh_leaf_size ? 1 : 0;
}

// 6b. emit: alive node j with rank r_j puts its children at 1 + 2 r_j and 2 + 2 r_j
__global__ void __launch_bounds__(lbvh_block)
    k_lbvh_emit(const int2* __restrict__ children, const int2* __restrict__ ranges,
        const signed char* __restrict__ axis, const int* __restrict__ alive,
        const int* __restrict__ rank, const lbvh_box* __restrict__ node_boxes,
        const lbvh_box* __restrict__ leaf_boxes, int n, vpt_bvh_node* __restrict__ out) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n - 1 || !alive[j]) return;
  auto put = [&](int at, int child) {
    vpt_bvh_node node;
    bool leaf = (child & lbvh_leaf_bit) != 0;
    int  idx  = child & ~lbvh_leaf_bit;
    auto box  = leaf ? leaf_boxes[idx] : node_boxes[idx];
    for (int k = 0; k < 3; k++) node.bbox_min[k] = box.lo[k], node.bbox_max[k] = box.hi[k];
    if (!leaf && alive[idx]) {
      node.start = 1 + 2 * rank[idx], node.num = 2, node.axis = axis[idx], node.internal = 1;
    } else {
      node.start    = leaf ? idx : ranges[idx].x;
      node.num      = (int16_t)(leaf ? 1 : ranges[idx].y - ranges[idx].x + 1);
      node.axis     = 0;
      node.internal = 0;
    }
    out[at] = node;
  };
  auto c = children[j];
  put(1 + 2 * rank[j], c.x);
  put(2 + 2 * rank[j], c.y);
  if (j == 0) {  // the root's own record
    vpt_bvh_node node;
    for (int k = 0; k < 3; k++)
      node.bbox_min[k] = node_boxes[0].lo[k], node.bbox_max[k] = node_boxes[0].hi[k];
    node.start = 1, node.num = 2, node.axis = axis[0], node.internal = 1;
    out[0] = node;
  }
}

}  // namespace vpt
