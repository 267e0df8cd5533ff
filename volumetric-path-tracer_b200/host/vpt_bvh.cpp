// vpt_bvh.cpp — host BVH build: middle split on the largest centroid axis,
// at most 4 primitives per leaf, children stored adjacently, built with an
// explicit stack. It produces the same tree as the reference's build_bvh
// (libs/yocto/yocto_bvh.cpp:411-611) — same split rule, same std::partition,
// same node order — so leaf order and therefore equal-distance tie-breaking
// match; the tests compare node arrays byte for byte.
#include <algorithm>
#include <stdexcept>
#include <thread>

#include "vpt_pathtrace.h"

namespace vpt {

namespace {

struct bbox3f {
  vec3f min = {flt_max, flt_max, flt_max};
  vec3f max = {-flt_max, -flt_max, -flt_max};
};
bbox3f merge(const bbox3f& a, const vec3f& b) { return {vmin(a.min, b), vmax(a.max, b)}; }
bbox3f merge(const bbox3f& a, const bbox3f& b) { return {vmin(a.min, b.min), vmax(a.max, b.max)}; }
vec3f  center(const bbox3f& a) { return (a.min + a.max) / 2; }

constexpr int bvh_max_prims = 4;

// split position and axis; falls back to the median when the split is empty
std::pair<int, int> split_middle(std::vector<int>& primitives, const std::vector<vec3f>& centers,
    int start, int end) {
  auto cbbox = bbox3f{};
  for (auto i = start; i < end; i++) cbbox = merge(cbbox, centers[primitives[i]]);
  auto csize = cbbox.max - cbbox.min;
  if (is_zero(csize)) return {(start + end) / 2, 0};
  auto axis = 0;
  if (csize.x >= csize.y && csize.x >= csize.z) axis = 0;
  if (csize.y >= csize.x && csize.y >= csize.z) axis = 1;
  if (csize.z >= csize.x && csize.z >= csize.y) axis = 2;
  auto split  = comp(center(cbbox), axis);
  auto middle = (int)(std::partition(primitives.data() + start, primitives.data() + end,
                          [axis, split, &centers](int primitive) {
                            return comp(centers[primitive], axis) < split;
                          }) -
                      primitives.data());
  if (middle == start || middle == end) return {(start + end) / 2, axis};
  return {middle, axis};
}

void build_nodes(bvh_data& bvh, const std::vector<bbox3f>& bboxes) {
  bvh.nodes.clear();
  bvh.nodes.reserve(bboxes.size() * 2);
  bvh.primitives.resize(bboxes.size());
  for (auto idx = 0; idx < (int)bboxes.size(); idx++) bvh.primitives[idx] = idx;
  auto centers = std::vector<vec3f>(bboxes.size());
  for (auto idx = 0; idx < (int)bboxes.size(); idx++) centers[idx] = center(bboxes[idx]);

  struct job { int node, start, end; };
  auto stack = std::vector<job>{{0, 0, (int)bboxes.size()}};
  bvh.nodes.emplace_back();
  while (!stack.empty()) {
    auto [nodeid, start, end] = stack.back();
    stack.pop_back();
    auto bbox = bbox3f{};
    for (auto i = start; i < end; i++) bbox = merge(bbox, bboxes[bvh.primitives[i]]);
    auto node = bvh_node{};
    node.bbox_min[0] = bbox.min.x, node.bbox_min[1] = bbox.min.y, node.bbox_min[2] = bbox.min.z;
    node.bbox_max[0] = bbox.max.x, node.bbox_max[1] = bbox.max.y, node.bbox_max[2] = bbox.max.z;
    if (end - start > bvh_max_prims) {
      auto [mid, axis] = split_middle(bvh.primitives, centers, start, end);
      node.internal    = 1;
      node.axis        = (int8_t)axis;
      node.num         = 2;
      node.start       = (int)bvh.nodes.size();
      bvh.nodes.emplace_back();
      bvh.nodes.emplace_back();
      stack.push_back({node.start + 0, start, mid});
      stack.push_back({node.start + 1, mid, end});
    } else {
      node.internal = 0;
      node.axis     = 0;
      node.num      = (int16_t)(end - start);
      node.start    = start;
    }
    bvh.nodes[nodeid] = node;
  }
  bvh.nodes.shrink_to_fit();
}

bbox3f transform_bbox(const frame3f& a, const bbox3f& b) {
  vec3f corners[8] = {{b.min.x, b.min.y, b.min.z}, {b.min.x, b.min.y, b.max.z},
      {b.min.x, b.max.y, b.min.z}, {b.min.x, b.max.y, b.max.z}, {b.max.x, b.min.y, b.min.z},
      {b.max.x, b.min.y, b.max.z}, {b.max.x, b.max.y, b.min.z}, {b.max.x, b.max.y, b.max.z}};
  auto  xformed    = bbox3f{};
  for (auto& corner : corners) xformed = merge(xformed, transform_point(a, corner));
  return xformed;
}

}  // namespace

bvh_shape make_bvh(const shape_data& shape) {
  if (!shape.points.empty() || !shape.lines.empty())
    throw std::runtime_error("point and line primitives are not supported by the GPU path");
  auto bboxes = std::vector<bbox3f>{};
  if (!shape.triangles.empty()) {
    bboxes.resize(shape.triangles.size());
    for (size_t idx = 0; idx < shape.triangles.size(); idx++) {
      auto& t  = shape.triangles[idx];
      auto &p0 = shape.positions[t.x], &p1 = shape.positions[t.y], &p2 = shape.positions[t.z];
      bboxes[idx] = {vmin(p0, vmin(p1, p2)), vmax(p0, vmax(p1, p2))};
    }
  } else if (!shape.quads.empty()) {
    bboxes.resize(shape.quads.size());
    for (size_t idx = 0; idx < shape.quads.size(); idx++) {
      auto& q  = shape.quads[idx];
      auto &p0 = shape.positions[q.x], &p1 = shape.positions[q.y], &p2 = shape.positions[q.z],
           &p3 = shape.positions[q.w];
      bboxes[idx] = {vmin(p0, vmin(p1, vmin(p2, p3))), vmax(p0, vmax(p1, vmax(p2, p3)))};
    }
  }
  auto bvh = bvh_data{};
  build_nodes(bvh, bboxes);
  return bvh;
}

// make_bvh(shape) on the device: LBVH in the reference's node format (vpt_cuda_build_shape_bvh)
bvh_shape make_bvh_device(const shape_data& shape, vpt_bvh_build_stats* stats) {
  auto desc = vpt_shape_desc{};
  desc.positions = (const float*)shape.positions.data(), desc.num_positions = (int)shape.positions.size();
  desc.triangles = (const int*)shape.triangles.data(), desc.num_triangles = (int)shape.triangles.size();
  desc.quads = (const int*)shape.quads.data(), desc.num_quads = (int)shape.quads.size();
  desc.num_points = (int)shape.points.size(), desc.num_lines = (int)shape.lines.size() / 2;
  auto n   = desc.num_triangles > 0 ? desc.num_triangles : desc.num_quads;
  auto bvh = bvh_data{};
  if (stats) *stats = vpt_bvh_build_stats{};
  if (n <= 0) return bvh;
  bvh.nodes.resize(2 * (size_t)n + 1);
  bvh.primitives.resize(n);
  int count = 0;
  if (vpt_cuda_build_shape_bvh(&desc, bvh.nodes.data(), bvh.nodes.size(), bvh.primitives.data(),
          bvh.primitives.size(), &count, stats) != VPT_OK)
    throw std::runtime_error(std::string("vpt_cuda_build_shape_bvh: ") + vpt_cuda_last_error());
  bvh.nodes.resize(count);
  return bvh;
}

bvh_scene make_bvh(const scene_data& scene, const pathtrace_params& params) {
  auto bvh = bvh_data{};
  bvh.shapes.resize(scene.shapes.size());
  if (params.device_bvh) {
    for (size_t idx = 0; idx < scene.shapes.size(); idx++)
      bvh.shapes[idx] = make_bvh_device(scene.shapes[idx]);
  } else {
  // one shape per task, like the reference's parallel_for over shapes
  auto nthreads = params.noparallel ? 1u : std::max(1u, std::thread::hardware_concurrency());
  nthreads      = (unsigned)std::min<size_t>(nthreads, std::max<size_t>(1, scene.shapes.size()));
  auto error    = std::string{};
  auto worker   = [&](unsigned tid) {
    try {
      for (auto idx = (size_t)tid; idx < scene.shapes.size(); idx += nthreads)
        bvh.shapes[idx] = make_bvh(scene.shapes[idx]);
    } catch (std::exception& e) {
      error = e.what();
    }
  };
  if (nthreads <= 1) {
    worker(0);
  } else {
    auto threads = std::vector<std::thread>{};
    for (auto t = 0u; t < nthreads; t++) threads.emplace_back(worker, t);
    for (auto& t : threads) t.join();
  }
  if (!error.empty()) throw std::runtime_error(error);
  }

  auto bboxes = std::vector<bbox3f>(scene.instances.size());
  for (size_t idx = 0; idx < bboxes.size(); idx++) {
    auto& instance = scene.instances[idx];
    auto& sbvh     = bvh.shapes.at(instance.shape);
    if (sbvh.nodes.empty()) continue;
    auto& root = sbvh.nodes[0];
    auto  box  = bbox3f{{root.bbox_min[0], root.bbox_min[1], root.bbox_min[2]},
        {root.bbox_max[0], root.bbox_max[1], root.bbox_max[2]}};
    bboxes[idx] = transform_bbox(instance.frame, box);
  }
  build_nodes(bvh, bboxes);
  return bvh;
}

}  // namespace vpt
