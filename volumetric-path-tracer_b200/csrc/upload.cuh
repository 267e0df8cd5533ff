// upload.cuh — host side of vpt_cuda_upload_scene: validates a vpt_scene_desc
// and flattens it into the 16-byte-aligned HBM layout of dscene.cuh.
// Host-only code (runs once per scene); nothing here is on the per-path route.
#pragma once

#if defined(VPT_HOST_POOL)
// tests/hostsim only: keep the flattened arrays in host memory
#include <cstdlib>
typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2 };
inline const char* cudaGetErrorString(cudaError_t) { return "host pool error"; }
#else
#include <cuda_runtime.h>
#endif

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/vpt_cuda.h"
#include "bvh.cuh"
#include "dscene.cuh"
#include "sdf.cuh"

namespace vpt {

struct device_pool {
  std::vector<void*> blocks;
  size_t             bytes = 0;
  template <typename T>
  cudaError_t upload(const std::vector<T>& host, const T** out) {
    *out = nullptr;
    // keep every pool non-null so kernels may form (never dereference) pointers
    auto n = host.size() * sizeof(T);
#if defined(VPT_HOST_POOL)
    void* d = aligned_alloc(256, ((n ? n : 16) + 255) / 256 * 256);
    if (!d) return cudaErrorMemoryAllocation;
    blocks.push_back(d);
    bytes += n;
    if (n) memcpy(d, host.data(), n);
#else
    void* d = nullptr;
    auto  e = cudaMalloc(&d, n ? n : 16);
    if (e != cudaSuccess) return e;
    blocks.push_back(d);
    bytes += n;
    if (n) {
      e = cudaMemcpy(d, host.data(), n, cudaMemcpyHostToDevice);
      if (e != cudaSuccess) return e;
    }
#endif
    *out = (const T*)d;
    return cudaSuccess;
  }
  void release() {
#if defined(VPT_HOST_POOL)
    for (auto b : blocks) free(b);
#else
    for (auto b : blocks) cudaFree(b);
#endif
    blocks.clear();
    bytes = 0;
  }
};

inline f4 pack(f3 v, float w) { return {v.x, v.y, v.z, w}; }
inline f4 pack(f3 v, int w) { return {v.x, v.y, v.z, i2f(w)}; }
inline frame3 to_frame(const float* f) {
  return {{f[0], f[1], f[2]}, {f[3], f[4], f[5]}, {f[6], f[7], f[8]}, {f[9], f[10], f[11]}};
}

// depth of the deepest leaf (root = 1); -1 on malformed child links
inline int bvh_depth(const vpt_bvh_desc& b) {
  if (b.num_nodes <= 0) return 0;
  std::vector<std::pair<int, int>> stack = {{0, 1}};
  int                              deepest = 0;
  size_t                           visited = 0;
  while (!stack.empty()) {
    auto [id, depth] = stack.back();
    stack.pop_back();
    if (id < 0 || id >= b.num_nodes || ++visited > (size_t)b.num_nodes) return -1;
    deepest = depth > deepest ? depth : deepest;
    auto& n = b.nodes[id];
    if (n.internal) {
      stack.push_back({n.start, depth + 1});
      stack.push_back({n.start + 1, depth + 1});
    }
  }
  return deepest;
}

struct flat_scene {
  DScene      d{};
  device_pool pool;
  int         stack_depth = 1;  // traversal stack entries a thread can need
  long long   num_triangles = 0;  // triangle primitives over all shapes (picks the extend build, wavefront.cuh)
  int         num_cameras = 0;
  std::vector<float> camera_aspect;
};

// Returns VPT_OK or a negative status and fills `error`.
inline int flatten_scene(const vpt_scene_desc& in, flat_scene& out, std::string& error) {
  auto fail = [&](int code, const std::string& msg) {
    error = msg;
    return code;
  };
  if (in.num_cameras < 0 || in.num_instances < 0 || in.num_shapes < 0 || in.num_materials < 0 ||
      in.num_textures < 0 || in.num_environments < 0 || in.num_volumes < 0 ||
      in.num_vol_instances < 0 || in.num_sdfs < 0 || in.num_lights < 0)
    return fail(VPT_ERR_INVALID, "negative element count in scene description");
  if (in.num_shapes > 0 && !in.shape_bvhs)
    return fail(VPT_ERR_INVALID, "shape_bvhs is null");

  std::vector<f4>  nodes, wnodes, prims, inst_trav, inst_frame, vcol, materials, cameras, environments;
  std::vector<f4>  vol_instances, volumes, sdfs, sdf_bounds, texf;
  std::vector<f4>  scene_leaves, inst_root;
  std::vector<i4>  shapes, elems, textures, lights;
  std::vector<float>         vpos, vnrm, vtex, cdf, voxels, srgb_lut(256);
  std::vector<unsigned char> texb;

  // Appends one BVH in the child-pair layout of dscene.cuh. `leaf_base` / `leaf_stride`
  // rebase a leaf's `start` (f4 index of its first primitive record, or its first
  // scene_leaves slot). Returns the root's {box, ref} through root0 / root1.
  // VPT_NO_WIDE=1: upload without the wide nodes, every ray walks the binary records (A/B runs)
  const char* no_wide_env = getenv("VPT_NO_WIDE");
  const bool  wide_on     = !(no_wide_env && atoi(no_wide_env) != 0);
  auto push_bvh = [&](const vpt_bvh_desc& bvh, int leaf_base, int leaf_stride, f4& root0,
                      f4& root1, int& wocc, std::string& why) -> bool {
    int              pair_base = (int)(nodes.size() / 4);
    std::vector<int> pair_of(bvh.num_nodes, -1);
    int              npairs = 0;
    for (int n = 0; n < bvh.num_nodes; n++)
      if (bvh.nodes[n].internal) pair_of[n] = npairs++;
    auto ref_of = [&](int n, int& ref) -> bool {
      auto& node = bvh.nodes[n];
      if (node.internal) {
        ref = pair_base + pair_of[n];
        return true;
      }
      long long start = (long long)leaf_base + (long long)node.start * leaf_stride;
      if (node.num < 1 || node.num > 7 || start >= leaf_start_limit) {
        why = "bvh leaf with " + std::to_string(node.num) + " primitives at record " +
              std::to_string(start) + " does not fit the device leaf reference";
        return false;
      }
      ref = make_ref(((int)start << 3) | (int)node.num, ref_leaf);
      return true;
    };
    auto box0 = [](const vpt_bvh_node& n) {
      return f4{n.bbox_min[0], n.bbox_min[1], n.bbox_min[2], n.bbox_max[0]};
    };
    for (int n = 0; n < bvh.num_nodes; n++) {
      auto& node = bvh.nodes[n];
      if (!node.internal) continue;
      if (node.start < 1 || node.start + 1 >= bvh.num_nodes) {
        why = "bvh child link out of range";
        return false;
      }
      auto &c0 = bvh.nodes[node.start], &c1 = bvh.nodes[node.start + 1];
      int  r0 = 0, r1 = 0;
      if (!ref_of(node.start, r0) || !ref_of(node.start + 1, r1)) return false;
      nodes.push_back(box0(c0));
      nodes.push_back({c0.bbox_max[1], c0.bbox_max[2], i2f(r0), i2f((int)node.axis & 3)});
      nodes.push_back(box0(c1));
      nodes.push_back({c1.bbox_max[1], c1.bbox_max[2], i2f(r1), i2f(0)});
    }
    int rr = 0;
    if (!ref_of(0, rr)) return false;
    root0 = box0(bvh.nodes[0]);
    root1 = {bvh.nodes[0].bbox_max[1], bvh.nodes[0].bbox_max[2], i2f(rr), i2f(wide_none)};
    // ---- the same tree collapsed into 4-slot wide nodes (dscene.cuh, bvh.cuh) ----
    // A wide node starts from a binary node's two children and, while a slot is free,
    // replaces the internal slot with the largest box by that node's two children, in place,
    // so slots stay in the tree's left-to-right order. The order table records, for every
    // sign octant of 1/d, where the reference's walk (near child first by the sign on each
    // node's split axis, yocto_bvh.cpp:743-749) reaches each slot.
    wocc = 0;
    if (!wide_on) return true;
    for (int n = 0; n < bvh.num_nodes; n++)
      for (int k = 0; k < 3; k++)
        if (!(bvh.nodes[n].bbox_min[k] == bvh.nodes[n].bbox_min[k]) ||
            !(bvh.nodes[n].bbox_max[k] == bvh.nodes[n].bbox_max[k]))
          return true;  // a NaN box: the ternary min / max of the binary walk decide
    if (!bvh.nodes[0].internal) {
      root1.w = i2f(rr);  // a single leaf is its own wide root
      return true;
    }
    int              wide_base = (int)(wnodes.size() / 8);
    std::vector<int> wide_of(bvh.num_nodes, -1), queue, nslots_of;
    auto wide_ref = [&](int n) {
      if (wide_of[n] < 0) {
        wide_of[n] = wide_base + (int)queue.size();
        queue.push_back(n);
      }
      return wide_of[n];
    };
    struct lnode {
      int node, l, r;  // l < 0: a slot
    };
    root1.w = i2f(wide_ref(0));
    for (size_t qi = 0; qi < queue.size(); qi++) {
      int  n    = queue[qi];
      auto tree = std::vector<lnode>{{n, 1, 2}, {bvh.nodes[n].start, -1, -1},
          {bvh.nodes[n].start + 1, -1, -1}};
      int nslots = 2;
      while (nslots < 4) {
        int    best = -1;
        double best_area = -1;
        for (int i = 0; i < (int)tree.size(); i++) {
          auto& b = bvh.nodes[tree[i].node];
          if (tree[i].l >= 0 || !b.internal) continue;
          double dx = (double)b.bbox_max[0] - b.bbox_min[0], dy = (double)b.bbox_max[1] - b.bbox_min[1],
                 dz = (double)b.bbox_max[2] - b.bbox_min[2];
          double area = dx * dy + dy * dz + dz * dx;
          if (!(area >= 0)) area = 0;
          if (area > best_area) best = i, best_area = area;
        }
        if (best < 0) break;
        int start    = bvh.nodes[tree[best].node].start;
        tree[best].l = (int)tree.size(), tree[best].r = (int)tree.size() + 1;
        tree.push_back({start, -1, -1});
        tree.push_back({start + 1, -1, -1});
        nslots++;
      }
      // slots in left-to-right order; visit sequence per octant
      std::vector<int> slot_of(tree.size(), -1), slots;
      struct walker {
        const std::vector<lnode>& tree;
        const vpt_bvh_desc&       bvh;
        void run(int i, int oct, std::vector<int>& seq) const {
          if (tree[i].l < 0) {
            seq.push_back(i);
            return;
          }
          bool flip = oct >= 0 && ((oct >> ((int)bvh.nodes[tree[i].node].axis & 3)) & 1);
          run(flip ? tree[i].r : tree[i].l, oct, seq);
          run(flip ? tree[i].l : tree[i].r, oct, seq);
        }
      } walk{tree, bvh};
      walk.run(0, -1, slots);
      for (int k = 0; k < (int)slots.size(); k++) slot_of[slots[k]] = k;
      f4  rec[8];
      int refs[4] = {make_ref(0, ref_empty), make_ref(0, ref_empty), make_ref(0, ref_empty),
          make_ref(0, ref_empty)};
      float box[6][4] = {};
      for (int k = 0; k < (int)slots.size(); k++) {
        int   node = tree[slots[k]].node;
        auto& b    = bvh.nodes[node];
        for (int c = 0; c < 3; c++) box[c][k] = b.bbox_min[c], box[3 + c][k] = b.bbox_max[c];
        if (b.internal) {
          if (b.start < 1 || b.start + 1 >= bvh.num_nodes) {
            why = "bvh child link out of range";
            return false;
          }
          refs[k] = wide_ref(node);
        } else if (!ref_of(node, refs[k])) {
          return false;
        }
      }
      for (int c = 0; c < 6; c++) rec[c] = {box[c][0], box[c][1], box[c][2], box[c][3]};
      rec[6] = {i2f(refs[0]), i2f(refs[1]), i2f(refs[2]), i2f(refs[3])};
      unsigned order[4] = {0, 0, 0, 0};
      for (int oct = 0; oct < 8; oct++) {
        std::vector<int> seq;
        walk.run(0, oct, seq);
        unsigned bits = 0;
        for (int rank = 0; rank < (int)seq.size(); rank++) bits |= (1u << rank) << (4 * slot_of[seq[rank]]);
        order[oct >> 1] |= bits << ((oct & 1) * 16);
      }
      rec[7] = {i2f((int)order[0]), i2f((int)order[1]), i2f((int)order[2]), i2f((int)order[3])};
      for (auto& r : rec) wnodes.push_back(r);
      nslots_of.push_back(nslots);
    }
    // stack entries a walk of this wide tree can hold (children are queued after parents)
    std::vector<int> occ(queue.size(), 0);
    for (int qi = (int)queue.size() - 1; qi >= 0; qi--) {
      int deepest = 0;
      auto rf = wnodes[8 * (size_t)(wide_base + qi) + 6];
      int  r4[4] = {f2i(rf.x), f2i(rf.y), f2i(rf.z), f2i(rf.w)};
      for (int k = 0; k < 4; k++)
        if (r4[k] >= 0) deepest = occ[r4[k] - wide_base] > deepest ? occ[r4[k] - wide_base] : deepest;
      occ[qi] = nslots_of[qi] - 1 + deepest > nslots_of[qi] ? nslots_of[qi] - 1 + deepest : nslots_of[qi];
    }
    wocc = occ[0];
    return true;
  };

  // ---- shapes: vertex pools, element indices, leaf-ordered primitive records ----
  std::vector<int> shape_kind(in.num_shapes, kind_empty);
  std::vector<f4>  shape_root0(in.num_shapes, f4{0, 0, 0, 0}), shape_root1(in.num_shapes, f4{0, 0, 0, 0});
  int              max_shape_depth = 0, max_shape_wocc = 0;
  for (int s = 0; s < in.num_shapes; s++) {
    auto& sh = in.shapes[s];
    if (sh.num_points > 0 || sh.num_lines > 0)
      return fail(VPT_ERR_UNSUPPORTED,
          "shape " + std::to_string(s) + " has point/line primitives (unsupported on the GPU path)");
    // same precedence as the reference: triangles, then quads
    int kind  = sh.num_triangles > 0 ? kind_triangles : (sh.num_quads > 0 ? kind_quads : kind_empty);
    int nelem = kind == kind_triangles ? sh.num_triangles : (kind == kind_quads ? sh.num_quads : 0);
    shape_kind[s] = kind;
    int pos_off = (int)(vpos.size() / 3);
    vpos.insert(vpos.end(), sh.positions, sh.positions + 3 * (size_t)sh.num_positions);
    int nrm_off = -1, tex_off = -1, col_off = -1;
    if (sh.num_normals > 0) {
      nrm_off = (int)(vnrm.size() / 3);
      vnrm.insert(vnrm.end(), sh.normals, sh.normals + 3 * (size_t)sh.num_normals);
    }
    if (sh.num_texcoords > 0) {
      tex_off = (int)(vtex.size() / 2);
      vtex.insert(vtex.end(), sh.texcoords, sh.texcoords + 2 * (size_t)sh.num_texcoords);
    }
    if (sh.num_colors > 0) {
      col_off = (int)vcol.size();
      for (int i = 0; i < sh.num_colors; i++)
        vcol.push_back({sh.colors[4 * i], sh.colors[4 * i + 1], sh.colors[4 * i + 2],
            sh.colors[4 * i + 3]});
    }
    int elem_off = (int)elems.size();
    for (int e = 0; e < nelem; e++) {
      i4 q = kind == kind_triangles
                 ? i4{sh.triangles[3 * e], sh.triangles[3 * e + 1], sh.triangles[3 * e + 2], -1}
                 : i4{sh.quads[4 * e], sh.quads[4 * e + 1], sh.quads[4 * e + 2], sh.quads[4 * e + 3]};
      int nv = kind == kind_triangles ? 3 : 4;
      int v[4] = {q.x, q.y, q.z, q.w};
      for (int k = 0; k < nv; k++)
        if (v[k] < 0 || v[k] >= sh.num_positions)
          return fail(VPT_ERR_INVALID, "vertex index out of range in shape " + std::to_string(s));
      elems.push_back(q);
    }
    shapes.push_back({kind, pos_off, nrm_off, tex_off});
    shapes.push_back({col_off, elem_off, nelem, 0});

    if (kind == kind_empty) continue;
    auto& bvh = in.shape_bvhs[s];
    if (bvh.num_nodes <= 0 || bvh.num_primitives != nelem)
      return fail(VPT_ERR_INVALID, "shape bvh " + std::to_string(s) + " does not match its shape");
    int depth = bvh_depth(bvh);
    if (depth < 0) return fail(VPT_ERR_INVALID, "malformed shape bvh " + std::to_string(s));
    max_shape_depth = depth > max_shape_depth ? depth : max_shape_depth;
    int stride    = kind == kind_triangles ? 3 : 4;
    int prim_base = (int)prims.size();
    if (kind == kind_triangles) out.num_triangles += bvh.num_primitives;
    for (int k = 0; k < bvh.num_primitives; k++) {
      int e = bvh.primitives[k];
      if (e < 0 || e >= nelem)
        return fail(VPT_ERR_INVALID, "bvh primitive out of range in shape " + std::to_string(s));
      auto q   = elems[elem_off + e];
      int  v[4] = {q.x, q.y, q.z, q.w};
      for (int c = 0; c < stride; c++) {
        auto p = sh.positions + 3 * (size_t)v[c];
        prims.push_back({p[0], p[1], p[2], i2f(e)});
      }
    }
    for (int n = 0; n < bvh.num_nodes; n++) {
      auto& node = bvh.nodes[n];
      if (!node.internal && (node.start < 0 || node.start + node.num > bvh.num_primitives))
        return fail(VPT_ERR_INVALID, "bvh leaf out of range in shape " + std::to_string(s));
    }
    auto why  = std::string{};
    int  wocc = 0;
    if (!push_bvh(bvh, prim_base, stride, shape_root0[s], shape_root1[s], wocc, why))
      return fail(VPT_ERR_UNSUPPORTED, why + " (shape " + std::to_string(s) + ")");
    max_shape_wocc = wocc > max_shape_wocc ? wocc : max_shape_wocc;
  }

  // ---- instances + scene bvh ----------------------------------------------------
  for (int i = 0; i < in.num_instances; i++) {
    auto& inst = in.instances[i];
    if (inst.shape < 0 || inst.shape >= in.num_shapes || inst.material < 0 ||
        inst.material >= in.num_materials)
      return fail(VPT_ERR_INVALID, "instance " + std::to_string(i) + " references a missing shape/material");
    auto frame = to_frame(inst.frame);
    auto inv   = inverse_frame(frame, true);  // yocto_bvh.cpp:854-855
    inst_trav.push_back(pack(inv.x, shape_kind[inst.shape] == kind_empty ? -1 : 0));
    inst_trav.push_back(pack(inv.y, shape_kind[inst.shape]));
    inst_trav.push_back(pack(inv.z, inst.shape));
    inst_trav.push_back(pack(inv.o, inst.material));
    inst_root.push_back(shape_root0[inst.shape]);
    inst_root.push_back(shape_root1[inst.shape]);
    inst_frame.push_back(pack(frame.x, inst.shape));
    inst_frame.push_back(pack(frame.y, inst.material));
    inst_frame.push_back(pack(frame.z, 0));
    inst_frame.push_back(pack(frame.o, 0));
  }
  // world box of an instance for culling (bvh.cuh): the 8 transformed corners of the shape's
  // root box, enlarged by 1e-4 * cond(frame) of (extent + |coordinates|); valid for ray origins
  // up to 100x that scale. Anything non-finite or empty disables culling for the instance.
  auto cull_record = [&](int i, f4& b0, f4& b1) {
    auto& inst = in.instances[i];
    b0 = {0, 0, 0, 0}, b1 = {0, 0, i2f(i), -1.0f};
    if (shape_kind[inst.shape] == kind_empty) return;
    auto   r0 = shape_root0[inst.shape], r1 = shape_root1[inst.shape];
    double lo[3] = {r0.x, r0.y, r0.z}, hi[3] = {r0.w, r1.x, r1.y};
    double wmin[3] = {1e300, 1e300, 1e300}, wmax[3] = {-1e300, -1e300, -1e300};
    for (int c = 0; c < 8; c++) {
      double p[3] = {c & 1 ? hi[0] : lo[0], c & 2 ? hi[1] : lo[1], c & 4 ? hi[2] : lo[2]};
      for (int k = 0; k < 3; k++) {
        double w = (double)inst.frame[0 + k] * p[0] + (double)inst.frame[3 + k] * p[1] +
                   (double)inst.frame[6 + k] * p[2] + (double)inst.frame[9 + k];
        if (!(w >= -1e30 && w <= 1e30)) return;  // NaN / overflow: never cull
        wmin[k] = w < wmin[k] ? w : wmin[k];
        wmax[k] = w > wmax[k] ? w : wmax[k];
      }
    }
    double scale = 0;
    for (int k = 0; k < 3; k++) {
      double m = std::fabs(wmin[k]) > std::fabs(wmax[k]) ? std::fabs(wmin[k]) : std::fabs(wmax[k]);
      double s = (wmax[k] - wmin[k]) + m;
      scale    = s > scale ? s : scale;
    }
    if (!(scale > 0)) return;
    // The reference tests the ray in object space, where its rounding error grows with the
    // condition number of the frame's linear part (eps * cond * |o|): the margin scales with
    // cond (Frobenius bound, 1 for a rigid frame), and a frame beyond 1e3 is never culled.
    double m[9], inv[9], f2 = 0, i2 = 0;
    for (int k = 0; k < 9; k++) m[k] = inst.frame[k], f2 += m[k] * m[k];
    double det = m[0] * (m[4] * m[8] - m[5] * m[7]) - m[3] * (m[1] * m[8] - m[2] * m[7]) +
                 m[6] * (m[1] * m[5] - m[2] * m[4]);
    if (!(std::fabs(det) > 0) || !(f2 < 1e300)) return;
    inv[0] = m[4] * m[8] - m[5] * m[7], inv[1] = m[2] * m[7] - m[1] * m[8], inv[2] = m[1] * m[5] - m[2] * m[4];
    inv[3] = m[5] * m[6] - m[3] * m[8], inv[4] = m[0] * m[8] - m[2] * m[6], inv[5] = m[2] * m[3] - m[0] * m[5];
    inv[6] = m[3] * m[7] - m[4] * m[6], inv[7] = m[1] * m[6] - m[0] * m[7], inv[8] = m[0] * m[4] - m[1] * m[3];
    for (int k = 0; k < 9; k++) i2 += (inv[k] / det) * (inv[k] / det);
    double cond = std::sqrt(f2 * i2) / 3;
    if (!(cond < 1e3)) return;
    double margin = 1e-4 * scale * (cond > 1 ? cond : 1);
    b0 = {(float)(wmin[0] - margin), (float)(wmin[1] - margin), (float)(wmin[2] - margin),
        (float)(wmax[0] + margin)};
    b1 = {(float)(wmax[1] + margin), (float)(wmax[2] + margin), i2f(i), (float)(100 * scale)};
  };
  int scene_root = 0, scene_wroot = wide_none, scene_depth = 0, scene_leaf = 0, scene_wocc = 0;
  f4  scene_box0{0, 0, 0, 0}, scene_box1{0, 0, 0, 0};
  if (in.num_instances > 0) {
    auto& bvh = in.scene_bvh;
    if (bvh.num_nodes <= 0 || bvh.num_primitives != in.num_instances)
      return fail(VPT_ERR_INVALID, "scene bvh does not match the instance list");
    scene_depth = bvh_depth(bvh);
    if (scene_depth < 0) return fail(VPT_ERR_INVALID, "malformed scene bvh");
    for (int k = 0; k < bvh.num_primitives; k++) {
      if (bvh.primitives[k] < 0 || bvh.primitives[k] >= in.num_instances)
        return fail(VPT_ERR_INVALID, "scene bvh primitive out of range");
      f4 b0, b1;
      cull_record(bvh.primitives[k], b0, b1);
      scene_leaves.push_back(b0), scene_leaves.push_back(b1);
    }
    for (int n = 0; n < bvh.num_nodes; n++) {
      auto& node = bvh.nodes[n];
      if (!node.internal) {
        if (node.start < 0 || node.start + node.num > bvh.num_primitives)
          return fail(VPT_ERR_INVALID, "scene bvh leaf out of range");
        scene_leaf = node.num > scene_leaf ? node.num : scene_leaf;
      }
    }
    auto why = std::string{};
    f4   root1;
    if (!push_bvh(bvh, 0, 1, scene_box0, root1, scene_wocc, why))
      return fail(VPT_ERR_UNSUPPORTED, why + " (scene bvh)");
    scene_box1  = {root1.x, root1.y, 0, 0};
    scene_root  = f2i(root1.z);
    scene_wroot = f2i(root1.w);
  }
  // one stack for both levels (bvh.cuh): a sibling per scene level, the queued
  // instances of one scene leaf, the sentinel, then a sibling per shape level
  out.stack_depth = scene_depth + 1 + scene_leaf + 1 + max_shape_depth + 1;
  // the wide walk pushes up to four slots per visit
  int wide_depth = (scene_wocc > 1 ? scene_wocc : 1) + scene_leaf + 1 +
                   (max_shape_wocc > 1 ? max_shape_wocc : 1) + 1;
  out.stack_depth = wide_depth > out.stack_depth ? wide_depth : out.stack_depth;
  if (out.stack_depth > vpt_stack_depth)
    return fail(VPT_ERR_UNSUPPORTED, "bvh deeper than the device traversal stack (" +
                                         std::to_string(out.stack_depth) + " > " +
                                         std::to_string(vpt_stack_depth) + ")");

  // ---- materials, textures ---------------------------------------------------------
  auto tex_ok = [&](int handle) { return handle >= -1 && handle < in.num_textures; };
  for (int m = 0; m < in.num_materials; m++) {
    auto& mt = in.materials[m];
    // SDF scenes of the reference carry texture handles past the texture list
    // (tests/06_gridsdf: color_tex 1, 2 with one texture); eval_material(scene, int) ignores
    // textures, so those are legal there and only instanced (mesh) materials are checked below
    materials.push_back({i2f(mt.type), mt.emission[0], mt.emission[1], mt.emission[2]});
    materials.push_back({mt.color[0], mt.color[1], mt.color[2], mt.roughness});
    materials.push_back({mt.metallic, mt.ior, mt.scattering[0], mt.scattering[1]});
    materials.push_back({mt.scattering[2], mt.scanisotropy, mt.trdepth, mt.opacity});
    materials.push_back({i2f(mt.emission_tex), i2f(mt.color_tex), i2f(mt.roughness_tex),
        i2f(mt.scattering_tex)});
    materials.push_back({i2f(mt.normal_tex), 0, 0, 0});
  }
  for (int i = 0; i < in.num_instances; i++) {
    auto& mt = in.materials[in.instances[i].material];
    if (!tex_ok(mt.emission_tex) || !tex_ok(mt.color_tex) || !tex_ok(mt.roughness_tex) ||
        !tex_ok(mt.scattering_tex) || !tex_ok(mt.normal_tex))
      return fail(VPT_ERR_INVALID, "material " + std::to_string(in.instances[i].material) +
                                       " references a missing texture");
  }
  for (int t = 0; t < in.num_textures; t++) {
    auto&  tx = in.textures[t];
    if (tx.width < 0 || tx.height < 0 || ((tx.pixelsf || tx.pixelsb) && (tx.width == 0 || tx.height == 0)))
      return fail(VPT_ERR_INVALID, "texture " + std::to_string(t) + " has a bad size");
    size_t n  = (size_t)tx.width * tx.height;
    if (tx.pixelsf) {
      textures.push_back({tx.width, tx.height, (tx.linear ? 1 : 0) | 2, (int)texf.size()});
      for (size_t i = 0; i < n; i++)
        texf.push_back({tx.pixelsf[4 * i], tx.pixelsf[4 * i + 1], tx.pixelsf[4 * i + 2],
            tx.pixelsf[4 * i + 3]});
    } else if (tx.pixelsb) {
      textures.push_back({tx.width, tx.height, (tx.linear ? 1 : 0), (int)(texb.size() / 4)});
      texb.insert(texb.end(), tx.pixelsb, tx.pixelsb + 4 * n);
    } else {
      textures.push_back({0, 0, 0, 0});
    }
  }
  // srgb_to_rgb(byte / 255.0f) with the host libm (yocto_color.h:212-226)
  for (int b = 0; b < 256; b++) {
    float srgb  = b / 255.0f;
    srgb_lut[b] = ((double)srgb <= 0.04045) ? srgb / 12.92f
                                            : powf((srgb + 0.055f) / (1.0f + 0.055f), 2.4f);
  }

  // ---- cameras, environments, lights -----------------------------------------------
  for (int c = 0; c < in.num_cameras; c++) {
    auto& cm    = in.cameras[c];
    auto  frame = to_frame(cm.frame);
    cameras.push_back(pack(frame.x, cm.orthographic ? 1 : 0));
    cameras.push_back(pack(frame.y, cm.lens));
    cameras.push_back(pack(frame.z, cm.film));
    cameras.push_back(pack(frame.o, cm.aspect));
    cameras.push_back({cm.focus, cm.aperture, 0, 0});
    out.camera_aspect.push_back(cm.aspect);
  }
  out.num_cameras = in.num_cameras;
  for (int e = 0; e < in.num_environments; e++) {
    auto& ev    = in.environments[e];
    auto  frame = to_frame(ev.frame);
    auto  inv   = inverse_frame(frame, false);  // inverse(environment.frame): rigid
    if (!tex_ok(ev.emission_tex))
      return fail(VPT_ERR_INVALID, "environment texture out of range");
    environments.push_back(pack(frame.x, 0)), environments.push_back(pack(frame.y, 0));
    environments.push_back(pack(frame.z, 0)), environments.push_back(pack(frame.o, 0));
    environments.push_back(pack(inv.x, 0)), environments.push_back(pack(inv.y, 0));
    environments.push_back(pack(inv.z, 0)), environments.push_back(pack(inv.o, 0));
    environments.push_back({ev.emission[0], ev.emission[1], ev.emission[2], i2f(ev.emission_tex)});
  }
  for (int l = 0; l < in.num_lights; l++) {
    auto& lt = in.lights[l];
    if (lt.instance >= in.num_instances || lt.environment >= in.num_environments ||
        lt.sdf >= in.num_sdfs)
      return fail(VPT_ERR_INVALID, "light " + std::to_string(l) + " references a missing object");
    lights.push_back({lt.instance, lt.environment, lt.sdf, (int)cdf.size()});
    // .y: the running sums never decrease (lights.cuh sample_discrete may then probe in any order)
    int sorted = lt.num_elements > 0 && lt.elements_cdf ? 1 : 0;
    for (int k = 1; sorted && k < lt.num_elements; k++)
      if (!(lt.elements_cdf[k] >= lt.elements_cdf[k - 1])) sorted = 0;
    lights.push_back({lt.num_elements, sorted, 0, 0});
    // a null pointer reserves the range: vpt_cuda_upload_scene_device_lights fills it on the device
    if (lt.num_elements > 0 && lt.elements_cdf)
      cdf.insert(cdf.end(), lt.elements_cdf, lt.elements_cdf + lt.num_elements);
    else if (lt.num_elements > 0)
      cdf.resize(cdf.size() + lt.num_elements, 0.0f);
  }

  // ---- sdf grids and functions -------------------------------------------------------
  // Culling record of one distance field (dscene.cuh `sdf_bounds`, sdf.cuh): world -> local is
  // q = M p + o (the FORWARD frame, yocto_sdfs.cpp:13, :20); the field is an exact distance to
  // a shape inside the local box [lo, hi]. W = bounding box of M^-1 ([lo, hi] - o), enlarged;
  // s = smallest singular value of M (Jacobi on M^T M, double), shrunk. VPT_NO_SDF_CULL=1: s = 0.
  const char* no_sdf_cull_env = getenv("VPT_NO_SDF_CULL");
  const bool  sdf_cull_on     = !(no_sdf_cull_env && atoi(no_sdf_cull_env) != 0);
  auto push_sdf_bounds = [&](const float* fr, const double lo[3], const double hi[3], bool ok,
                             bool grid) {
    f4 b0{0, 0, 0, 0}, b1{0, 0, 0, grid ? 1.0f : 0.0f};
    auto never = [&]() { sdf_bounds.push_back(b0), sdf_bounds.push_back(b1); };
    if (!ok || !sdf_cull_on) return never();
    double m[3][3];  // m[r][c]: column c of the frame is fr[3c .. 3c+2]
    for (int c = 0; c < 3; c++)
      for (int r = 0; r < 3; r++) m[r][c] = fr[3 * c + r];
    double det = m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) -
                 m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
                 m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
    if (!(std::fabs(det) > 1e-30) || !(std::fabs(det) < 1e30)) return never();
    double inv[3][3];
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) {
        int r1 = (c + 1) % 3, r2 = (c + 2) % 3, c1 = (r + 1) % 3, c2 = (r + 2) % 3;
        inv[r][c] = (m[r1][c1] * m[r2][c2] - m[r1][c2] * m[r2][c1]) / det;  // adjugate / det
      }
    // eigenvalues of A = M^T M by cyclic Jacobi rotations
    double a[3][3];
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) a[i][j] = m[0][i] * m[0][j] + m[1][i] * m[1][j] + m[2][i] * m[2][j];
    for (int sweep = 0; sweep < 32; sweep++)
      for (int pi = 0; pi < 2; pi++)
        for (int qi = pi + 1; qi < 3; qi++) {
          if (std::fabs(a[pi][qi]) < 1e-300) continue;
          double th = (a[qi][qi] - a[pi][pi]) / (2 * a[pi][qi]);
          double tt = (th >= 0 ? 1 : -1) / (std::fabs(th) + std::sqrt(th * th + 1));
          double cs = 1 / std::sqrt(tt * tt + 1), sn = tt * cs;
          for (int k = 0; k < 3; k++) {
            double akp = a[k][pi], akq = a[k][qi];
            a[k][pi] = cs * akp - sn * akq, a[k][qi] = sn * akp + cs * akq;
          }
          for (int k = 0; k < 3; k++) {
            double apk = a[pi][k], aqk = a[qi][k];
            a[pi][k] = cs * apk - sn * aqk, a[qi][k] = sn * apk + cs * aqk;
          }
        }
    double lmin = std::fmin(a[0][0], std::fmin(a[1][1], a[2][2]));
    if (!(lmin > 0)) return never();
    double smin = std::sqrt(lmin) * (1 - 1e-3);
    double wmin[3] = {1e300, 1e300, 1e300}, wmax[3] = {-1e300, -1e300, -1e300};
    for (int c = 0; c < 8; c++) {
      double q[3] = {(c & 1 ? hi[0] : lo[0]) - fr[9], (c & 2 ? hi[1] : lo[1]) - fr[10],
          (c & 4 ? hi[2] : lo[2]) - fr[11]};
      for (int k = 0; k < 3; k++) {
        double w = inv[k][0] * q[0] + inv[k][1] * q[1] + inv[k][2] * q[2];
        if (!(w >= -1e30 && w <= 1e30)) return never();
        wmin[k] = std::fmin(wmin[k], w), wmax[k] = std::fmax(wmax[k], w);
      }
    }
    double scale = 0;
    for (int k = 0; k < 3; k++)
      scale = std::fmax(scale, (wmax[k] - wmin[k]) + std::fmax(std::fabs(wmin[k]), std::fabs(wmax[k])));
    double margin = 1e-4 * scale + 1e-30;
    b0 = {(float)(wmin[0] - margin), (float)(wmin[1] - margin), (float)(wmin[2] - margin),
        (float)(smin * smin)};
    b1 = {(float)(wmax[0] + margin), (float)(wmax[1] + margin), (float)(wmax[2] + margin),
        grid ? 1.0f : 0.0f};
    sdf_bounds.push_back(b0), sdf_bounds.push_back(b1);
  };
  for (int v = 0; v < in.num_volumes; v++) {
    auto&  vol = in.volumes[v];
    if (vol.whd[0] <= 0 || vol.whd[1] <= 0 || vol.whd[2] <= 0 || !vol.voxels)
      return fail(VPT_ERR_INVALID, "sdf grid " + std::to_string(v) + " is empty");
    size_t n   = (size_t)vol.whd[0] * vol.whd[1] * vol.whd[2];
    volumes.push_back({i2f(vol.whd[0]), i2f(vol.whd[1]), i2f(vol.whd[2]), vol.res});
    volumes.push_back({i2f((int)voxels.size()), 0, 0, 0});
    voxels.insert(voxels.end(), vol.voxels, vol.voxels + n);
  }
  for (int v = 0; v < in.num_vol_instances; v++) {
    auto& vi = in.vol_instances[v];
    if (vi.volume < 0 || vi.volume >= in.num_volumes || vi.material < 0 ||
        vi.material >= in.num_materials)
      return fail(VPT_ERR_INVALID, "sdf grid instance references a missing volume/material");
    auto frame = to_frame(vi.frame);
    vol_instances.push_back(pack(frame.x, vi.scalef));
    vol_instances.push_back(pack(frame.y, vi.volume));
    vol_instances.push_back(pack(frame.z, vi.material));
    vol_instances.push_back(pack(frame.o, 0));
    auto& gv   = in.volumes[vi.volume];
    auto  size = proxy_box_size(frame.o, gv.whd[0], gv.whd[1], gv.whd[2], gv.res, vi.scalef);
    vol_instances.push_back(pack(size * 0.5f, 0));
    vol_instances.push_back(pack(size, 0));
    // the proxy box of eval_sdf (yocto_sdfs.cpp:36-40): [o, o + res whd scalef] in LOCAL space
    auto&  vol = in.volumes[vi.volume];
    double lo[3], hi[3];
    bool   ok = true;
    for (int k = 0; k < 3; k++) {
      double size = (double)vol.res * vol.whd[k] * vi.scalef;
      lo[k] = 0, hi[k] = size * (1 + 1e-5);
      ok    = ok && size > 0 && size < 1e30;
    }
    push_sdf_bounds(vi.frame, lo, hi, ok, true);
  }
  for (int s = 0; s < in.num_sdfs; s++) {
    auto& sd = in.sdfs[s];
    if (sd.type < 0 || sd.type > VPT_SDF_TORUS || sd.material < 0 || sd.material >= in.num_materials)
      return fail(VPT_ERR_INVALID, "sdf " + std::to_string(s) + " has a bad type/material");
    auto frame = to_frame(sd.frame);
    auto inv   = inverse_frame(frame, false);  // yocto_pathtrace.cpp:330
    sdfs.push_back(pack(frame.x, sd.type)), sdfs.push_back(pack(frame.y, sd.material));
    sdfs.push_back(pack(frame.z, 0)), sdfs.push_back(pack(frame.o, 0));
    sdfs.push_back(pack(inv.x, 0)), sdfs.push_back(pack(inv.y, 0));
    sdfs.push_back(pack(inv.z, 0)), sdfs.push_back(pack(inv.o, 0));
    sdfs.push_back({sd.params[0], sd.params[1], sd.params[2], sd.params[3]});
    sdfs.push_back({sd.whd[0], sd.whd[1], sd.whd[2], 0});
    // local bounding box of the shape each typed closure measures the distance to
    // (yocto_sceneio.cpp:3690-3726, yocto_sdfs.h:43-80); a plane has none
    double lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    bool   ok = true;
    auto   sym = [&](double x, double y, double z) {
      ok = ok && x >= 0 && y >= 0 && z >= 0;
      lo[0] = -x, lo[1] = -y, lo[2] = -z, hi[0] = x, hi[1] = y, hi[2] = z;
    };
    switch (sd.type) {
      case VPT_SDF_BBOX: sym(sd.params[1], sd.params[2], sd.params[3]); ok = ok && sd.params[0] >= 0; break;
      case VPT_SDF_BOX:
        for (int k = 0; k < 3; k++) hi[k] = sd.whd[k], ok = ok && sd.whd[k] >= 0;
        break;
      case VPT_SDF_CAPPED_CONE: {
        double r = std::fmax(sd.params[1], sd.params[2]);
        sym(r, sd.params[0], r);
        ok = ok && sd.params[1] >= 0 && sd.params[2] >= 0;
      } break;
      case VPT_SDF_SPHERE: sym(sd.params[0], sd.params[0], sd.params[0]); break;
      case VPT_SDF_TORUS:
        sym((double)sd.params[0] + sd.params[1], sd.params[1], (double)sd.params[0] + sd.params[1]);
        ok = ok && sd.params[0] >= 0;
        break;
      default: ok = false; break;  // plane: unbounded
    }
    for (int k = 0; k < 3; k++) ok = ok && std::isfinite(lo[k]) && std::isfinite(hi[k]);
    push_sdf_bounds(sd.frame, lo, hi, ok, false);
  }

  // offsets into the pools are stored as int
  const size_t int_max = 0x7fffffff;
  if (texf.size() > int_max || texb.size() / 4 > int_max || voxels.size() > int_max ||
      cdf.size() > int_max || vpos.size() / 3 > int_max || elems.size() > int_max ||
      nodes.size() > int_max)
    return fail(VPT_ERR_UNSUPPORTED, "scene pools exceed 2^31 elements");

  // ---- to the device -----------------------------------------------------------------
  auto& d = out.d;
  auto& p = out.pool;
  cudaError_t e = cudaSuccess;
#define VPT_UP(vec, field) \
  if (e == cudaSuccess) e = p.upload(vec, &d.field)
  VPT_UP(nodes, nodes);
  VPT_UP(wnodes, wnodes);
  VPT_UP(prims, prims);
  VPT_UP(scene_leaves, scene_leaves);
  VPT_UP(inst_trav, inst_trav);
  VPT_UP(inst_root, inst_root);
  VPT_UP(inst_frame, inst_frame);
  VPT_UP(shapes, shapes);
  VPT_UP(vpos, vpos);
  VPT_UP(vnrm, vnrm);
  VPT_UP(vtex, vtex);
  VPT_UP(vcol, vcol);
  VPT_UP(elems, elems);
  VPT_UP(materials, materials);
  VPT_UP(textures, textures);
  VPT_UP(texb, texb);
  VPT_UP(texf, texf);
  VPT_UP(srgb_lut, srgb_lut);
  VPT_UP(cameras, cameras);
  VPT_UP(environments, environments);
  VPT_UP(lights, lights);
  VPT_UP(cdf, cdf);
  VPT_UP(vol_instances, vol_instances);
  VPT_UP(volumes, volumes);
  VPT_UP(voxels, voxels);
  VPT_UP(sdfs, sdfs);
  VPT_UP(sdf_bounds, sdf_bounds);
#undef VPT_UP
  if (e != cudaSuccess) {
    p.release();
    return fail(e == cudaErrorMemoryAllocation ? VPT_ERR_OUT_OF_MEMORY : VPT_ERR_CUDA,
        std::string("scene upload failed: ") + cudaGetErrorString(e));
  }
  d.scene_root        = scene_root;
  d.scene_wroot       = scene_wroot;
  d.scene_box0 = scene_box0, d.scene_box1 = scene_box1;
  d.num_instances     = in.num_instances;
  d.num_environments  = in.num_environments;
  d.num_lights        = in.num_lights;
  d.num_vol_instances = in.num_vol_instances;
  d.num_sdfs          = in.num_sdfs;
  return VPT_OK;
}

}  // namespace vpt
