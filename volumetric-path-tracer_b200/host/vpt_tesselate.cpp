// vpt_tesselate.cpp — tesselate_surfaces (yocto_pathtrace.h:128, yocto_pathtrace.cpp:1230-1280)
// with the numeric stages on the device (include/vpt_cuda.h):
//   Catmull-Clark levels for positions and texture coordinates  -> vpt_cuda_catmullclark
//   smooth vertex normals                                        -> vpt_cuda_vertex_normals
//   texture displacement along the normals                       -> vpt_cuda_displace
// The host keeps the index bookkeeping between them: split_facevarying (one vertex per distinct
// {position, normal, texcoord} index triple, numbered in first-appearance order,
// yocto_shape.cpp:2597-2650) and quads_to_triangles (:2565-2573). The resulting shapes equal the
// reference's tessellation bit for bit (tests/test_gpu_tesselate.py).
#include <cstring>
#include <stdexcept>
#include <unordered_map>

#include "vpt_pathtrace.h"

namespace vpt {

namespace {

void check(int status, const char* what) {
  if (status != VPT_OK) throw std::runtime_error(std::string(what) + ": " + vpt_cuda_last_error());
}

template <typename V>
void subdivide(std::vector<vec4i>& quads, std::vector<V>& verts, int dim, bool lock_boundary, int levels) {
  if (quads.empty() || levels <= 0) return;
  int32_t* out_quads = nullptr;
  float*   out_verts = nullptr;
  int32_t  nq = 0, nv = 0;
  check(vpt_cuda_catmullclark((const int32_t*)quads.data(), (int)quads.size(), (const float*)verts.data(),
            (int)verts.size(), dim, lock_boundary ? 1 : 0, levels, &out_quads, &nq, &out_verts, &nv),
      "vpt_cuda_catmullclark");
  quads.resize(nq), verts.resize(nv);
  if (nq) memcpy(quads.data(), out_quads, sizeof(vec4i) * (size_t)nq);
  if (nv) memcpy(verts.data(), out_verts, sizeof(V) * (size_t)nv);
  vpt_cuda_free_host(out_quads), vpt_cuda_free_host(out_verts);
}

std::vector<vec3f> vertex_normals(const void* elems, size_t nelems, int verts_per,
    const std::vector<vec3f>& positions) {
  auto normals = std::vector<vec3f>(positions.size());
  check(vpt_cuda_vertex_normals((const int32_t*)elems, (int)nelems, verts_per, (const float*)positions.data(),
            (int)positions.size(), (float*)normals.data()),
      "vpt_cuda_vertex_normals");
  return normals;
}

struct triple_hash {
  size_t operator()(const vec3i& v) const {
    auto h = std::hash<int>{};
    auto s = (size_t)0;
    for (auto x : {v.x, v.y, v.z}) s ^= h(x) + 0x9e3779b9 + (s << 6) + (s >> 2);
    return s;
  }
};
struct triple_eq {
  bool operator()(const vec3i& a, const vec3i& b) const { return a.x == b.x && a.y == b.y && a.z == b.z; }
};

}  // namespace

void tesselate_surface(shape_data& shape, const subdiv_data& subdiv_, const scene_data& scene) {
  auto subdiv = subdiv_;
  if (subdiv.subdivisions != 0) {
    subdivide(subdiv.quadspos, subdiv.positions, 3, false, subdiv.subdivisions);
    subdivide(subdiv.quadstexcoord, subdiv.texcoords, 2, true, subdiv.subdivisions);
    if (subdiv.smooth) {
      subdiv.normals   = vertex_normals(subdiv.quadspos.data(), subdiv.quadspos.size(), 4, subdiv.positions);
      subdiv.quadsnorm = subdiv.quadspos;
    } else {
      subdiv.normals = {}, subdiv.quadsnorm = {};
    }
  }
  // split_facevarying: vertices numbered in the order their index triple first appears
  auto vert_map = std::unordered_map<vec3i, int, triple_hash, triple_eq>{};
  auto triples  = std::vector<vec3i>{};
  shape.quads.assign(subdiv.quadspos.size(), vec4i{0, 0, 0, 0});
  for (size_t fid = 0; fid < subdiv.quadspos.size(); fid++) {
    for (int c = 0; c < 4; c++) {
      auto v  = vec3i{(&subdiv.quadspos[fid].x)[c],
          !subdiv.quadsnorm.empty() ? (&subdiv.quadsnorm[fid].x)[c] : -1,
          !subdiv.quadstexcoord.empty() ? (&subdiv.quadstexcoord[fid].x)[c] : -1};
      auto it = vert_map.find(v);
      if (it == vert_map.end()) {
        it = vert_map.insert({v, (int)triples.size()}).first;
        triples.push_back(v);
      }
      (&shape.quads[fid].x)[c] = it->second;
    }
  }
  shape.positions.clear(), shape.normals.clear(), shape.texcoords.clear();
  if (!subdiv.positions.empty())
    for (auto& v : triples) shape.positions.push_back(subdiv.positions[v.x]);
  if (!subdiv.normals.empty())
    for (auto& v : triples) shape.normals.push_back(subdiv.normals[v.y]);
  if (!subdiv.texcoords.empty())
    for (auto& v : triples) shape.texcoords.push_back(subdiv.texcoords[v.z]);
  // quads_to_triangles
  shape.triangles.clear();
  for (auto& q : shape.quads) {
    shape.triangles.push_back({q.x, q.y, q.w});
    if (q.z != q.w) shape.triangles.push_back({q.z, q.w, q.y});
  }
  shape.quads = {}, shape.points = {}, shape.lines = {};

  if (subdiv.displacement != 0 && subdiv.displacement_tex >= 0 && !shape.triangles.empty()) {
    if (subdiv.displacement_tex >= (int)scene.textures.size())
      throw std::runtime_error("subdiv displacement texture out of range");
    if (shape.texcoords.size() != shape.positions.size())
      throw std::runtime_error("displaced subdiv without texture coordinates");
    if (shape.normals.empty())
      shape.normals = vertex_normals(shape.triangles.data(), shape.triangles.size(), 3, shape.positions);
    auto& t    = scene.textures[subdiv.displacement_tex];
    auto  desc = vpt_texture_desc{t.width, t.height, t.linear ? 1 : 0,
        t.pixelsf.empty() ? nullptr : (const float*)t.pixelsf.data(),
        t.pixelsb.empty() ? nullptr : (const uint8_t*)t.pixelsb.data()};
    check(vpt_cuda_displace((float*)shape.positions.data(), (const float*)shape.normals.data(),
              (const float*)shape.texcoords.data(), (int)shape.positions.size(), &desc, subdiv.displacement),
        "vpt_cuda_displace");
    if (subdiv.smooth)
      shape.normals = vertex_normals(shape.triangles.data(), shape.triangles.size(), 3, shape.positions);
    else
      shape.normals = {};
  }
}

void tesselate_surfaces(scene_data& scene) {
  for (auto& subdiv : scene.subdivs) {
    if (subdiv.shape < 0 || subdiv.shape >= (int)scene.shapes.size())
      throw std::runtime_error("subdiv references a missing shape");
    tesselate_surface(scene.shapes[subdiv.shape], subdiv, scene);
  }
  scene.subdivs.clear();  // done: a second call is a no-op
}

}  // namespace vpt
