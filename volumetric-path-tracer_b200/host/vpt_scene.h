// vpt_scene.h — host-side scene model of the B200 path tracer.
//
// Plain structs with the same fields, names and defaults as the reference's
// scene model (libs/yocto/yocto_scene.h:84-249, yocto_shape.h:74,
// yocto_bvh.h:73-92) restricted to what the render path reads, so a
// yocto::scene_data converts field by field (host/yocto_adapter.h). Scenes are
// loaded from ".vptscene" archives (vpt_archive.h); asset loading (JSON, PLY,
// OBJ, HDR ...) stays with the reference's loaders and is out of scope here.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/vpt_cuda.h"
#include "../csrc/vmath.cuh"

namespace vpt {

using vec2f   = f2;
using vec3f   = f3;
using vec4f   = f4;
using frame3f = frame3;
struct vec3i { int x, y, z; };
struct vec4i { int x, y, z, w; };
struct vec4b { uint8_t x, y, z, w; };

inline const frame3f identity3x4f = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 0, 0}};

struct camera_data {
  frame3f frame        = identity3x4f;
  bool    orthographic = false;
  float   lens         = 0.050f;
  float   film         = 0.036f;
  float   aspect       = 1.500f;
  float   focus        = 10000;
  float   aperture     = 0;
};

struct texture_data {
  int                width  = 0;
  int                height = 0;
  bool               linear = false;
  std::vector<vec4f> pixelsf = {};
  std::vector<vec4b> pixelsb = {};
};

enum struct material_type {
  matte, glossy, reflective, transparent, refractive, subsurface, volumetric, gltfpbr
};

struct material_data {
  material_type type         = material_type::matte;
  vec3f         emission     = {0, 0, 0};
  vec3f         color        = {0, 0, 0};
  float         roughness    = 0;
  float         metallic     = 0;
  float         ior          = 1.5f;
  vec3f         scattering   = {0, 0, 0};
  float         scanisotropy = 0;
  float         trdepth      = 0.01f;
  float         opacity      = 1;
  int emission_tex = invalidid, color_tex = invalidid, roughness_tex = invalidid,
      scattering_tex = invalidid, normal_tex = invalidid;
};

struct instance_data {
  frame3f frame    = identity3x4f;
  int     shape    = invalidid;
  int     material = invalidid;
};

struct environment_data {
  frame3f frame        = identity3x4f;
  vec3f   emission     = {0, 0, 0};
  int     emission_tex = invalidid;
};

struct shape_data {
  std::vector<int>   points    = {};  // unsupported on the GPU path (kept to report it)
  std::vector<int>   lines     = {};  // 2 ints per line; unsupported
  std::vector<vec3i> triangles = {};
  std::vector<vec4i> quads     = {};
  std::vector<vec3f> positions = {};
  std::vector<vec3f> normals   = {};
  std::vector<vec2f> texcoords = {};
  std::vector<vec4f> colors    = {};
};

// voxel-grid sdf (yocto_scene.h:203-219), x fastest
struct volume_data {
  vec3i              whd = {0, 0, 0};
  std::vector<float> vol = {};
  float              res = 0;
};
struct volume_instance {
  int     volume   = invalidid;
  int     material = invalidid;
  float   scalef   = 1;
  frame3f frame    = identity3x4f;
};

// analytic sdf with typed parameters; the reference stores a std::function
// closure built from these (yocto_sceneio.cpp:3685-3728)
enum struct sdf_type { bbox, box, capped_cone, plane, sphere, torus };
struct sdf_data {
  int      material  = invalidid;
  frame3f  frame     = identity3x4f;
  vec3f    whd       = {0, 0, 0};
  sdf_type type      = sdf_type::sphere;
  float    params[4] = {0, 0, 0, 0};
};

// subdivision surface (yocto_scene.h:161-183): face-varying control cage of shape `shape`
struct subdiv_data {
  std::vector<vec4i> quadspos = {}, quadsnorm = {}, quadstexcoord = {};
  std::vector<vec3f> positions = {}, normals = {};
  std::vector<vec2f> texcoords = {};
  int   subdivisions = 0;
  bool  catmullclark = true, smooth = true;
  float displacement     = 0;
  int   displacement_tex = invalidid;
  int   shape            = invalidid;
};

struct scene_data {
  std::vector<camera_data>      cameras       = {};
  std::vector<instance_data>    instances     = {};
  std::vector<environment_data> environments  = {};
  std::vector<shape_data>       shapes        = {};
  std::vector<texture_data>     textures      = {};
  std::vector<material_data>    materials     = {};
  std::vector<volume_data>      volumes       = {};
  std::vector<volume_instance>  vol_instances = {};
  std::vector<sdf_data>         sdfs          = {};
  std::vector<subdiv_data>      subdivs       = {};  // pending: see tesselate_surfaces
};

// bvh (yocto_bvh.h:73-92); the node is the C ABI's vpt_bvh_node (32 bytes)
using bvh_node = vpt_bvh_node;
struct bvh_data {
  std::vector<bvh_node> nodes      = {};
  std::vector<int>      primitives = {};
  std::vector<bvh_data> shapes     = {};
};
using bvh_scene = bvh_data;
using bvh_shape = bvh_data;

// Loads a ".vptscene" archive written by tools/vpt_export or by the
// Python fixture generators (volumetric-path-tracer_b200/archive.py).
bool load_scene(const std::string& filename, scene_data& scene, std::string& error);
// the `subdivs` arrays of an archive (tools/vpt_export --keep-subdivs / --dump-subdivs) only
bool load_subdivs(const std::string& filename, scene_data& scene, std::string& error);

}  // namespace vpt
