// dscene.cuh — the flattened scene as it lies in HBM, and the load helpers.
//
// Everything a path touches is a 16-byte-aligned record read with 128-bit
// loads through the read-only path:
//   nodes      4 x f4 (64 bytes) per INTERNAL BVH node = the boxes of its two
//              children: {c0.min.xyz, c0.max.x} {c0.max.yz, ref0, axis}
//              {c1.min.xyz, c1.max.x} {c1.max.yz, ref1, 0}. A ref >= 0 is the
//              record of an internal child; a leaf child has no record, its ref
//              is ~((first primitive record (f4 index) << 3 | count) << 2)
//              (scene BVH: first slot of scene_leaves). `axis` is the split axis
//              of the node the record belongs to. All shape BVHs and the scene
//              BVH are concatenated. One 64-byte fetch per visited node.
//   wnodes     8 x f4 (128 bytes) per WIDE node: up to four descendants of one binary
//              node (its two children, the widest of them expanded again while slots
//              are free), as SoA so that one 16-byte load holds one box coordinate of
//              all four slots: {min.x[4]} {min.y[4]} {min.z[4]} {max.x[4]} {max.y[4]}
//              {max.z[4]} {ref[4]} {order[8 x 16 bit]}. Slot refs are wide-node indices
//              or the same leaf refs as above (unused slot: ref_empty). `order` holds,
//              for each of the 8 sign octants of 1/d, the one-hot rank (4 bits per
//              slot) in which the reference's depth-first walk reaches the slots
//              (bvh.cuh). Rays whose 1/d is not finite walk `nodes` instead.
//   prims      primitives gathered in BVH leaf order, vertex positions inline:
//              quad = 4 x f4 {p_k.xyz, payload}, triangle = 3 x f4;
//              payload of the first f4 = element index.
//   scene_leaves 2 x f4 per scene-BVH primitive slot: the instance's world box
//              enlarged for culling {min.xyz, max.x} {max.yz, instance id,
//              largest |ray origin| culling is valid for (-1: never cull)}
//   inst_trav  4 x f4 per instance: columns of inverse(frame, non_rigid) in xyz;
//              w = {0 or -1 (no geometry), shape kind, shape id, material id}
//   inst_root  2 x f4 per instance: the shape BVH's root box and refs
//              {min.xyz, max.x} {max.yz, ref in nodes, ref in wnodes}
//   inst_frame 4 x f4 per instance: forward frame columns.
// The reference layout this replaces: bvh_data / shape_data / instance_data
// (yocto_bvh.h:73-92, yocto_shape.h:74, yocto_scene.h:142-148).
#pragma once

#include "vmath.cuh"

namespace vpt {

enum shape_kind : int { kind_empty = 0, kind_triangles = 1, kind_quads = 2 };

struct DScene {
  // acceleration
  const f4*  nodes;
  const f4*  wnodes;
  const f4*  prims;
  const f4*  scene_leaves;
  const f4*  inst_trav;
  const f4*  inst_root;
  const f4*  inst_frame;
  f4         scene_box0, scene_box1;  // root box of the scene BVH {min.xyz, max.x} {max.yz, 0, 0}
  int        scene_root;              // ref of the scene BVH root (valid if num_instances > 0)
  int        scene_wroot;             // the same root in `wnodes` (wide_none: walk `nodes`)
  int        num_instances;
  // shapes for shading: 2 x i4 per shape {kind, pos, nrm, tex} {col, elem, 0, 0}
  const i4*    shapes;
  const float* vpos;   // 3 floats per vertex
  const float* vnrm;   // 3 floats per vertex
  const float* vtex;   // 2 floats per vertex
  const f4*    vcol;   // 4 floats per vertex
  const i4*    elems;  // vertex indices per element (triangles: w = -1)
  // materials: 6 x f4 per material
  const f4* materials;
  // textures: i4 {w, h, flags(bit0 linear, bit1 float), offset} ; texel pools
  const i4*            textures;
  const unsigned char* texb;  // 4 bytes per texel
  const f4*            texf;
  const float*         srgb_lut;  // 256 entries: srgb_to_rgb(b / 255.0f), host glibc
  // cameras: 5 x f4 per camera
  const f4* cameras;
  // environments: 9 x f4 per env: frame(4) inverse rigid frame(4) {emission, tex}
  const f4* environments;
  int       num_environments;
  // lights: 2 x i4 per light {instance, environment, sdf, cdf_off} {cdf_n, cdf sorted, 0, 0}
  const i4*    lights;
  const float* cdf;
  int          num_lights;
  // sdf grids / functions
  // 6 x f4: frame cols; w = {scalef, volume, material, 0}; then the proxy box of eval_sdf
  // (yocto_sdfs.cpp:36-40), constant per instance: {size / 2, 0} {size, 0}
  const f4*    vol_instances;
  const f4*    volumes;        // 2 x f4: {w, h, d (ints), res} {offset (int), 0, 0, 0}
  const float* voxels;
  const f4*    sdfs;  // 10 x f4: frame(4; w = type, material), inverse rigid (4), params, whd
  // 2 x f4 per distance field (grid instances, then functions): a world-space box holding the
  // pre-image of the field's local bounding box {min.xyz, s^2} {max.xyz, 1 if grid}; s bounds
  // the frame's smallest singular value from below (0: never skip this field), see sdf.cuh
  const f4*    sdf_bounds;
  int          num_vol_instances;
  int          num_sdfs;
};

struct DParams {
  int camera;
  int shader;
  int samples;   // total target (samples == 1 -> preview branch)
  int bounces;
  int noimplicit_mis;
  int spheretrace_maxiter;
  int width, height;
};

// ---- loads -----------------------------------------------------------------
VPT_FN f4 ld4(const f4* p) {
#if defined(__CUDA_ARCH__)
  auto v = __ldg(reinterpret_cast<const float4*>(p));
  return {v.x, v.y, v.z, v.w};
#else
  return *p;
#endif
}
VPT_FN i4 ld4(const i4* p) {
#if defined(__CUDA_ARCH__)
  auto v = __ldg(reinterpret_cast<const int4*>(p));
  return {v.x, v.y, v.z, v.w};
#else
  return *p;
#endif
}
VPT_FN float ld1(const float* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
VPT_FN int ld1(const int* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
VPT_FN f3 ld3(const float* p) { return {ld1(p), ld1(p + 1), ld1(p + 2)}; }
VPT_FN f2 ld2(const float* p) { return {ld1(p), ld1(p + 1)}; }

VPT_FN frame3 ld_frame(const f4* p) {
  auto a = ld4(p), b = ld4(p + 1), c = ld4(p + 2), d = ld4(p + 3);
  return {xyz(a), xyz(b), xyz(c), xyz(d)};
}

// ---- ray / hit -------------------------------------------------------------
struct ray3 {
  f3    o, d;
  float tmin, tmax;
};
VPT_FN ray3 make_ray(f3 o, f3 d) { return {o, d, 1e-4f, flt_max}; }  // yocto_geometry.h:118-133
VPT_FN f3 ray_point(const ray3& r, float t) { return r.o + r.d * t; }

struct hit_t {
  int   instance, element;
  f2    uv;
  float distance;
  bool  hit;
};

}  // namespace vpt
