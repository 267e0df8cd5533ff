// warpsim — TEST-ONLY design aid. Replays the extend queue of one wavefront iteration on a
// model of the persistent extend kernel's warps (32 lanes, refill threshold, the real trav_t
// from csrc/bvh.cuh compiled for the host) and counts, per SIMT schedule, how many warp
// instructions the walk costs and how many lanes are live in each of them. The per-operation
// instruction costs are estimates read off an ncu source table; only ratios between schedules
// mean anything. Nothing in the product builds, links or calls this.
//
// usage: warpsim scene.vptscene [--shader S] [--resolution R] [--bounces B] [--warps W]
#define VPT_HOST_POOL 1
#define VPT_WORK_COUNTERS 1
#define VPT_RAY_SINK 1
#include <algorithm>
#include <cstring>
#include <random>

#include "../../volumetric-path-tracer_b200/csrc/shaders.cuh"
#include "../../volumetric-path-tracer_b200/csrc/tiling.cuh"
#include "../../volumetric-path-tracer_b200/csrc/upload.cuh"
#include "../../volumetric-path-tracer_b200/host/vpt_archive.h"
#include "../../volumetric-path-tracer_b200/host/vpt_pathtrace.h"

using namespace vpt;

struct lane_t {
  trav_t t;
  ray3   ray;
  bool   has = false;
  int    op = 0, arg = 0;
  alignas(8) int mem[vpt_stack_depth * vpt_stack_words];
  stack_t st() { return stack_t{mem, 1}; }
};

struct cost_t {
  double inst = 0, lanes = 0;  // warp instructions, sum of live lanes over them
  void   add(double c, int n) { inst += c, lanes += c * n; }
};
struct tally_t {
  cost_t node, leaf, inst, pop, sleaf, round, refill, retire;
  double total() const { return node.inst + leaf.inst + inst.inst + pop.inst + sleaf.inst + round.inst + refill.inst + retire.inst; }
  double lanes() const { return node.lanes + leaf.lanes + inst.lanes + pop.lanes + sleaf.lanes + round.lanes + refill.lanes + retire.lanes; }
};

// estimated warp instructions per execution (ncu r02c source table, 03_volume)
constexpr double c_node = 200, c_tri = 70, c_quad = 150, c_inst = 150, c_pop = 22, c_sleaf = 45,
                 c_round = 35, c_refill = 90, c_retire = 70;

struct warp_t {
  lane_t lane[32];
  bool   more = true;
};

struct sim_t {
  const DScene&            sc;
  const std::vector<ray3>& queue;
  size_t                   head = 0;
  tally_t                  tl;
  int                      refill_at = 8;
  sim_t(const DScene& s, const std::vector<ray3>& q) : sc(s), queue(q) {}

  void refill(warp_t& w) {
    int idle = 0;
    for (auto& l : w.lane) idle += !l.has;
    if (!w.more || idle < refill_at) return;
    tl.refill.add(c_refill, idle);
    for (auto& l : w.lane) {
      if (l.has) continue;
      if (head >= queue.size()) { w.more = false; break; }
      l.ray = queue[head++];
      l.t.begin(sc, -1, l.ray, l.st());
      l.has = true, l.op = 0;
    }
    if (head >= queue.size()) w.more = false;
  }
  int leaf_cost(lane_t& l, int& iters) {  // one lane's leaf loop
    iters = l.arg & 7;
    return l.t.kind == kind_quads ? 1 : 0;
  }
  void run_leaves(warp_t& w, unsigned mask) {
    // lanes loop over their own primitive counts; quads and triangles are separate branches
    int maxq = 0, maxt = 0, nq[5] = {0}, nt[5] = {0};
    for (int i = 0; i < 32; i++)
      if (mask >> i & 1) {
        auto& l = w.lane[i];
        int   n = l.arg & 7;
        if (l.t.kind == kind_quads) { maxq = std::max(maxq, n); for (int k = 0; k < n && k < 5; k++) nq[k]++; }
        else { maxt = std::max(maxt, n); for (int k = 0; k < n && k < 5; k++) nt[k]++; }
        l.t.test_leaf(sc, l.arg);
        l.op = 0;
      }
    for (int k = 0; k < maxq && k < 5; k++) tl.leaf.add(c_quad, nq[k]);
    for (int k = 0; k < maxt && k < 5; k++) tl.leaf.add(c_tri, nt[k]);
  }
  void run_nodes(warp_t& w, unsigned mask) {
    tl.node.add(c_node, __builtin_popcount(mask));
    for (int i = 0; i < 32; i++)
      if (mask >> i & 1) w.lane[i].t.visit_node(sc, w.lane[i].arg, w.lane[i].st()), w.lane[i].op = 0;
  }
  void run_instances(warp_t& w, unsigned mask) {
    tl.inst.add(c_inst, __builtin_popcount(mask));
    for (int i = 0; i < 32; i++)
      if (mask >> i & 1)
        w.lane[i].t.enter_instance(sc, w.lane[i].ray, w.lane[i].arg, w.lane[i].st()), w.lane[i].op = 0;
  }
  // one pop per lane of `mask` (fetch with one pop): cost c_pop + scene-leaf loops
  void pop_once(warp_t& w, unsigned mask) {
    tl.pop.add(c_pop, __builtin_popcount(mask));
    int sl_iters = 0, sl_lanes = 0;
    for (int i = 0; i < 32; i++)
      if (mask >> i & 1) {
        auto& l  = w.lane[i];
        auto  c0 = (long long)g_work.cull_tests;
        int   cur0 = l.t.cur;
        // peek: a scene leaf costs its instance loop
        auto en = l.st().get(l.t.cur - 1);
        bool sleaf = en.ref < 0 && ((~en.ref) & 3) == ref_leaf && l.t.kind < 0 && en.t0 <= l.t.tmax * box_slack;
        l.op = l.t.fetch(sc, l.ray, l.st(), 1, l.arg);
        if (sleaf) sl_iters = std::max(sl_iters, (int)((~en.ref) >> 2 & 7)), sl_lanes++;
        (void)c0, (void)cur0;
      }
    if (sl_iters) tl.sleaf.add(c_sleaf * sl_iters, sl_lanes);
  }
  unsigned ballot(warp_t& w, int op) {
    unsigned m = 0;
    for (int i = 0; i < 32; i++)
      if (w.lane[i].has && w.lane[i].op == op) m |= 1u << i;
    return m;
  }
  unsigned can_pop(warp_t& w) {
    unsigned m = 0;
    for (int i = 0; i < 32; i++)
      if (w.lane[i].has && w.lane[i].op == 0 && w.lane[i].t.cur != 0) m |= 1u << i;
    return m;
  }
  bool retire(warp_t& w) {
    int n = 0, live = 0;
    for (auto& l : w.lane) {
      if (l.has && l.op == 0 && l.t.cur == 0) l.has = false, n++;
      live += l.has;
    }
    if (n) tl.retire.add(c_retire, n);
    return live > 0 || w.more;
  }

  // schedule A: the shipped kernel (<= pops pops per round with the node visits inside the loop,
  // then instance entry, then leaf test)
  bool step_bounded(warp_t& w, int pops) {
    refill(w);
    int live = 0;
    for (auto& l : w.lane) live += l.has;
    tl.round.add(c_round, live);
    for (int k = 0; k < pops; k++) {
      unsigned m = can_pop(w);
      if (!m) break;
      pop_once(w, m);
      unsigned nodes = ballot(w, trav_t::op_node);
      if (nodes) run_nodes(w, nodes);
    }
    unsigned in = ballot(w, trav_t::op_instance);
    if (in) run_instances(w, in);
    unsigned lf = ballot(w, trav_t::op_leaf);
    if (lf) run_leaves(w, lf);
    return retire(w);
  }
  // schedule B: election - fetch up to `pops` entries, then run the operation most lanes hold
  // (and any other operation that at least `also` lanes hold)
  bool step_elect(warp_t& w, int pops, int also) {
    refill(w);
    int live = 0;
    for (auto& l : w.lane) live += l.has;
    tl.round.add(c_round, live);
    for (int k = 0; k < pops; k++) {
      unsigned m = can_pop(w);
      if (!m) break;
      pop_once(w, m);
    }
    unsigned nd = ballot(w, trav_t::op_node), in = ballot(w, trav_t::op_instance), lf = ballot(w, trav_t::op_leaf);
    int cn = __builtin_popcount(nd), ci = __builtin_popcount(in), cl = __builtin_popcount(lf);
    int best = std::max(cn, std::max(ci, cl));
    if (best > 0) {
      if (cn == best || cn >= also) { if (cn) run_nodes(w, nd); }
      if ((ci == best && cn != best) || ci >= also) { if (ci) run_instances(w, in); }
      if ((cl == best && cn != best && ci != best) || cl >= also) { if (cl) run_leaves(w, lf); }
    }
    return retire(w);
  }
};

// schedule C: a pool of R rays per warp whose state lives in shared memory; every operation
// runs on up to 32 rays that hold it, compacted to the lanes (c_state extra instructions per
// execution to load / store the ray state)
constexpr double c_state = 25;
struct pool_t {
  std::vector<lane_t> ray;
  bool                more = true;
};
struct poolsim_t {
  const DScene&            sc;
  const std::vector<ray3>& queue;
  size_t                   head = 0;
  tally_t                  tl;
  int                      min_lanes = 24;
  poolsim_t(const DScene& s, const std::vector<ray3>& q) : sc(s), queue(q) {}
  bool step(pool_t& p) {
    int R = (int)p.ray.size();
    // refill
    int idle = 0;
    for (auto& l : p.ray) idle += !l.has;
    if (p.more && idle >= 8) {
      tl.refill.add(c_refill * ((idle + 31) / 32), std::min(idle, 32));
      for (auto& l : p.ray) {
        if (l.has) continue;
        if (head >= queue.size()) { p.more = false; break; }
        l.ray = queue[head++];
        l.t.begin(sc, -1, l.ray, l.st());
        l.has = true, l.op = 0;
      }
    }
    // classify: pop for every ray without an operation (stale entries / sentinels inline)
    std::vector<int> ids[6];
    for (int i = 0; i < R; i++) {
      auto& l = p.ray[i];
      if (l.has && l.op == 0 && l.t.cur != 0) ids[0].push_back(i);
    }
    for (size_t k = 0; k < ids[0].size(); k += 32) {
      int n = (int)std::min<size_t>(32, ids[0].size() - k);
      tl.pop.add(c_pop + c_state, n);
      for (int j = 0; j < n; j++) {
        auto& l  = p.ray[ids[0][k + j]];
        auto  en = l.st().get(l.t.cur - 1);
        bool  sleaf = en.ref < 0 && ((~en.ref) & 3) == ref_leaf && l.t.kind < 0 && en.t0 <= l.t.tmax * box_slack;
        if (sleaf) l.op = 4;
        else l.op = l.t.fetch(sc, l.ray, l.st(), 1, l.arg);
      }
    }
    for (int op = 1; op <= 4; op++) ids[op].clear();
    int live = 0;
    for (int i = 0; i < R; i++)
      if (p.ray[i].has) {
        live++;
        if (p.ray[i].op) ids[p.ray[i].op].push_back(i);
      }
    tl.round.add(c_round, std::min(live, 32));
    int best = 0;
    for (int op = 1; op <= 4; op++)
      if (ids[op].size() > ids[best == 0 ? op : best].size() || best == 0) best = ids[op].size() > (best ? ids[best].size() : 0) ? op : best;
    if (best == 0) best = 1;
    for (int op = 1; op <= 4; op++) {
      size_t n = ids[op].size();
      if (n == 0) continue;
      // full chunks always; the remainder only if it is the best operation or has enough lanes
      size_t full = n / 32 * 32;
      size_t run  = full;
      if (n - full >= (size_t)min_lanes || (op == best && full == 0)) run = n;
      for (size_t k = 0; k < run; k += 32) {
        int m = (int)std::min<size_t>(32, run - k);
        if (op == trav_t::op_node) {
          tl.node.add(c_node + c_state, m);
          for (int j = 0; j < m; j++) { auto& l = p.ray[ids[op][k + j]]; l.t.visit_node(sc, l.arg, l.st()); l.op = 0; }
        } else if (op == trav_t::op_instance) {
          tl.inst.add(c_inst + c_state, m);
          for (int j = 0; j < m; j++) { auto& l = p.ray[ids[op][k + j]]; l.t.enter_instance(sc, l.ray, l.arg, l.st()); l.op = 0; }
        } else if (op == trav_t::op_leaf) {
          int maxq = 0, maxt = 0, nq[8] = {0}, nt[8] = {0};
          for (int j = 0; j < m; j++) {
            auto& l = p.ray[ids[op][k + j]];
            int   c = l.arg & 7;
            if (l.t.kind == kind_quads) { maxq = std::max(maxq, c); for (int q = 0; q < c; q++) nq[q]++; }
            else { maxt = std::max(maxt, c); for (int q = 0; q < c; q++) nt[q]++; }
            l.t.test_leaf(sc, l.arg);
            l.op = 0;
          }
          tl.leaf.add(c_state, m);
          for (int q = 0; q < maxq; q++) tl.leaf.add(c_quad, nq[q]);
          for (int q = 0; q < maxt; q++) tl.leaf.add(c_tri, nt[q]);
        } else {
          int iters = 0;
          for (int j = 0; j < m; j++) {
            auto& l  = p.ray[ids[op][k + j]];
            auto  en = l.st().get(l.t.cur - 1);
            iters    = std::max(iters, (int)((~en.ref) >> 2 & 7));
            l.op     = l.t.fetch(sc, l.ray, l.st(), 1, l.arg);
          }
          tl.sleaf.add(c_sleaf * iters + c_state, m);
        }
      }
    }
    int n = 0;
    live  = 0;
    for (auto& l : p.ray) {
      if (l.has && l.op == 0 && l.t.cur == 0) l.has = false, n++;
      live += l.has;
    }
    if (n) tl.retire.add(c_retire * ((n + 31) / 32), std::min(n, 32));
    return live > 0 || p.more;
  }
};

// schedule D: K rays per LANE (state in the lane's own shared-memory column: no compaction, no
// bank conflicts). Each iteration pops for the rays without an operation, elects the operation
// most lanes hold for at least one of their rays (plus any other that `also` lanes hold) and
// runs it for one ray per lane.
struct colsim_t {
  const DScene&            sc;
  const std::vector<ray3>& queue;
  size_t                   head = 0;
  tally_t                  tl;
  int                      K = 3, also = 24, refill_at = 8;
  colsim_t(const DScene& s, const std::vector<ray3>& q) : sc(s), queue(q) {}
  bool step(pool_t& p) {
    int idle = 0;
    for (auto& l : p.ray) idle += !l.has;
    if (p.more && idle >= refill_at) {
      // refill: per column index k, the lanes whose ray k is idle
      for (int k = 0; k < K; k++) {
        int n = 0;
        for (int i = 0; i < 32; i++) n += !p.ray[k * 32 + i].has;
        if (n) tl.refill.add(c_refill, n);
      }
      for (auto& l : p.ray) {
        if (l.has) continue;
        if (head >= queue.size()) { p.more = false; break; }
        l.ray = queue[head++];
        l.t.begin(sc, -1, l.ray, l.st());
        l.has = true, l.op = 0;
      }
    }
    for (int k = 0; k < K; k++) {
      int n = 0;
      for (int i = 0; i < 32; i++) {
        auto& l = p.ray[k * 32 + i];
        if (l.has && l.op == 0 && l.t.cur != 0) {
          n++;
          auto en    = l.st().get(l.t.cur - 1);
          bool sleaf = en.ref < 0 && ((~en.ref) & 3) == ref_leaf && l.t.kind < 0 && en.t0 <= l.t.tmax * box_slack;
          if (sleaf) l.op = 4;
          else l.op = l.t.fetch(sc, l.ray, l.st(), 1, l.arg);
        }
      }
      if (n) tl.pop.add(c_pop + c_state, n);
    }
    tl.round.add(c_round, 32);
    // per operation: the lanes holding it, and the ray each would run
    int pickr[5][32], cnt[5] = {0};
    for (int op = 1; op <= 4; op++)
      for (int i = 0; i < 32; i++) {
        pickr[op][i] = -1;
        for (int k = 0; k < K; k++)
          if (p.ray[k * 32 + i].has && p.ray[k * 32 + i].op == op) { pickr[op][i] = k * 32 + i; break; }
        cnt[op] += pickr[op][i] >= 0;
      }
    int best = 1;
    for (int op = 2; op <= 4; op++)
      if (cnt[op] > cnt[best]) best = op;
    for (int op = 1; op <= 4; op++) {
      if (cnt[op] == 0 || !(op == best || cnt[op] >= also)) continue;
      int m = cnt[op];
      if (op == trav_t::op_node) {
        tl.node.add(c_node + c_state, m);
        for (int i = 0; i < 32; i++) if (pickr[op][i] >= 0) { auto& l = p.ray[pickr[op][i]]; l.t.visit_node(sc, l.arg, l.st()); l.op = 0; }
      } else if (op == trav_t::op_instance) {
        tl.inst.add(c_inst + c_state, m);
        for (int i = 0; i < 32; i++) if (pickr[op][i] >= 0) { auto& l = p.ray[pickr[op][i]]; l.t.enter_instance(sc, l.ray, l.arg, l.st()); l.op = 0; }
      } else if (op == trav_t::op_leaf) {
        int maxq = 0, maxt = 0, nq[8] = {0}, nt[8] = {0};
        for (int i = 0; i < 32; i++) if (pickr[op][i] >= 0) {
          auto& l = p.ray[pickr[op][i]];
          int   c = l.arg & 7;
          if (l.t.kind == kind_quads) { maxq = std::max(maxq, c); for (int q = 0; q < c; q++) nq[q]++; }
          else { maxt = std::max(maxt, c); for (int q = 0; q < c; q++) nt[q]++; }
          l.t.test_leaf(sc, l.arg);
          l.op = 0;
        }
        tl.leaf.add(c_state, m);
        for (int q = 0; q < maxq; q++) tl.leaf.add(c_quad, nq[q]);
        for (int q = 0; q < maxt; q++) tl.leaf.add(c_tri, nt[q]);
      } else {
        int iters = 0;
        for (int i = 0; i < 32; i++) if (pickr[op][i] >= 0) {
          auto& l  = p.ray[pickr[op][i]];
          auto  en = l.st().get(l.t.cur - 1);
          iters    = std::max(iters, (int)((~en.ref) >> 2 & 7));
          l.op     = l.t.fetch(sc, l.ray, l.st(), 1, l.arg);
        }
        tl.sleaf.add(c_sleaf * iters + c_state, m);
      }
    }
    int live = 0;
    for (int k = 0; k < K; k++) {
      int n = 0;
      for (int i = 0; i < 32; i++) {
        auto& l = p.ray[k * 32 + i];
        if (l.has && l.op == 0 && l.t.cur == 0) l.has = false, n++;
        live += l.has;
      }
      if (n) tl.retire.add(c_retire, n);
    }
    return live > 0 || p.more;
  }
};

int main(int argc, const char** argv) {
  if (argc < 2) return 2;
  auto params = pathtrace_params{};
  params.shader = (pathtrace_shader_type)1, params.resolution = 640, params.bounces = 64, params.samples = 1;
  int nwarps = 148 * 6 * 4;
  for (int i = 2; i < argc; i++) {
    auto a    = std::string{argv[i]};
    auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "--shader") {
      auto name = next();
      for (size_t k = 0; k < pathtrace_shader_names.size(); k++)
        if (pathtrace_shader_names[k] == name) params.shader = (pathtrace_shader_type)k;
    } else if (a == "--resolution") params.resolution = std::stoi(next());
    else if (a == "--bounces") params.bounces = std::stoi(next());
    else if (a == "--warps") nwarps = std::stoi(next());
  }
  auto error = std::string{};
  auto scene = scene_data{};
  if (!load_scene(argv[1], scene, error)) return fprintf(stderr, "%s\n", error.c_str()), 1;
  auto bvh     = make_bvh(scene, params);
  auto lights  = make_lights(scene, params);
  auto storage = scene_desc_storage{};
  make_scene_desc(storage, scene, bvh, lights);
  auto flat = flat_scene{};
  if (flatten_scene(storage.desc, flat, error) != VPT_OK) return fprintf(stderr, "%s\n", error.c_str()), 1;
  auto& sc    = flat.d;
  auto  state = make_state(scene, params);
  auto  prm   = DParams{params.camera, (int)params.shader, params.samples, params.bounces, 0,
      params.spheretrace_maxiter, state.width, state.height};
  // the extend queue of a steady-state iteration: slots in tile order, each at a random bounce
  // of its pixel's path
  auto queue = std::vector<ray3>{};
  auto sink  = std::vector<ray3>{};
  auto pick  = std::mt19937{7};
  alignas(8) int stack_mem[vpt_stack_depth * vpt_stack_words];
  auto st = stack_t{stack_mem, 1};
  int  slots = tile_slot_count(state.width, state.height);
  for (int s = 0; s < slots; s++) {
    int idx = tile_pixel_index(s, state.width, state.height);
    if (idx < 0) continue;
    auto rng = rng_t{state.rngs[idx].state, state.rngs[idx].inc};
    sink.clear();
    vpt_ray_sink = &sink;
    auto ray = sample_camera(sc, prm, idx, rng);
    switch ((int)params.shader) {
      case 0: shade_path<0>(sc, prm, ray, rng, st); break;
      case 1: shade_path<1>(sc, prm, ray, rng, st); break;
      default: shade_path<2>(sc, prm, ray, rng, st); break;
    }
    vpt_ray_sink = nullptr;
    if (!sink.empty()) queue.push_back(sink[pick() % sink.size()]);
  }
  printf("warpsim: %zu rays, %d warps (%.1f rays per warp)\n", queue.size(), nwarps, queue.size() / (double)nwarps);
  auto report = [&](const char* name, sim_t& sim) {
    auto& t = sim.tl;
    auto  f = [&](const char* n, cost_t& c) {
      printf("    %-7s %5.1f%% of instructions at %4.1f lanes\n", n, 100 * c.inst / t.total(), c.inst ? c.lanes / c.inst : 0);
    };
    printf("%s: %.1f warp instructions per ray, %.2f lanes per instruction\n", name, t.total() / queue.size(),
        t.lanes() / t.total());
    f("node", t.node), f("leaf", t.leaf), f("inst", t.inst), f("pop", t.pop), f("sleaf", t.sleaf), f("round", t.round),
        f("refill", t.refill), f("retire", t.retire);
  };
  auto run = [&](const char* name, auto step, int refill_at) {
    auto sim   = sim_t{sc, queue};
    sim.refill_at = refill_at;
    auto warps = std::vector<warp_t>(nwarps);
    auto alive = std::vector<char>(nwarps, 1);
    int  left  = nwarps;
    while (left > 0)
      for (int i = 0; i < nwarps; i++)
        if (alive[i] && !step(sim, warps[i])) alive[i] = 0, left--;
    report(name, sim);
  };
  for (int K : {1, 2, 3, 4})
    for (int also : {20, 33}) {
      auto sim  = colsim_t{sc, queue};
      sim.K = K, sim.also = also, sim.refill_at = 8;
      int  np    = std::max(1, 148 * 768 / (32 * K));
      auto pools = std::vector<pool_t>(np);
      for (auto& p : pools) p.ray.resize(32 * K);
      auto alive = std::vector<char>(np, 1);
      int  left  = np;
      while (left > 0)
        for (int i = 0; i < np; i++)
          if (alive[i] && !sim.step(pools[i])) alive[i] = 0, left--;
      auto& t = sim.tl;
      auto  f = [&](const char* n, cost_t& c) {
        printf("    %-7s %5.1f%% of instructions at %4.1f lanes\n", n, 100 * c.inst / t.total(), c.inst ? c.lanes / c.inst : 0);
      };
      printf("%d rays per lane, also >= %d: %.1f warp instructions per ray, %.2f lanes per instruction\n", K, also,
          t.total() / queue.size(), t.lanes() / t.total());
      f("node", t.node), f("leaf", t.leaf), f("inst", t.inst), f("pop", t.pop), f("sleaf", t.sleaf), f("round", t.round),
          f("refill", t.refill), f("retire", t.retire);
    }
  for (int R : {96})
    for (int minl : {24}) {
      auto sim      = poolsim_t{sc, queue};
      sim.min_lanes = minl;
      int  np       = nwarps * 32 / R * (R > 64 ? 2 : 1) / (R > 64 ? 2 : 1);
      np            = std::max(1, 148 * 640 / R);  // ~640 rays per SM
      auto pools    = std::vector<pool_t>(np);
      for (auto& p : pools) p.ray.resize(R);
      auto alive = std::vector<char>(np, 1);
      int  left  = np;
      while (left > 0)
        for (int i = 0; i < np; i++)
          if (alive[i] && !sim.step(pools[i])) alive[i] = 0, left--;
      char name[64];
      snprintf(name, sizeof name, "pool of %d rays, remainder >= %d", R, minl);
      auto& t = sim.tl;
      auto  f = [&](const char* n, cost_t& c) {
        printf("    %-7s %5.1f%% of instructions at %4.1f lanes\n", n, 100 * c.inst / t.total(), c.inst ? c.lanes / c.inst : 0);
      };
      printf("%s: %.1f warp instructions per ray, %.2f lanes per instruction\n", name, t.total() / queue.size(),
          t.lanes() / t.total());
      f("node", t.node), f("leaf", t.leaf), f("inst", t.inst), f("pop", t.pop), f("sleaf", t.sleaf), f("round", t.round),
          f("refill", t.refill), f("retire", t.retire);
    }
  run("bounded 4 pops (shipped)", [](sim_t& s, warp_t& w) { return s.step_bounded(w, 4); }, 8);
  run("bounded 2 pops", [](sim_t& s, warp_t& w) { return s.step_bounded(w, 2); }, 8);
  run("bounded 1 pop", [](sim_t& s, warp_t& w) { return s.step_bounded(w, 1); }, 8);
  for (int pops : {3})
    for (int also : {16}) {
      char name[64];
      snprintf(name, sizeof name, "elect pops %d also %d", pops, also);
      run(name, [=](sim_t& s, warp_t& w) { return s.step_elect(w, pops, also); }, 8);
    }
  run("elect pops 2 also 12 refill 4", [](sim_t& s, warp_t& w) { return s.step_elect(w, 2, 12); }, 4);
  return 0;
}
