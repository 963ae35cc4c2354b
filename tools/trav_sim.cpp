// tools/trav_sim.cpp — CPU model of the persistent-warp traversal scheduler (csrc/traverse.cuh trace_persistent).
//
// Development tool, not part of the product: it replays the warp-round schedule of the CUDA traversal kernels on the host,
// over the REAL compressed wide BVH that flatten_scene() builds, for a batch of rays, and reports what a design variant would
// do to lane utilisation and to the number of warp-instructions issued per ray — so that scheduling designs can be compared
// without GPU time.  Stage costs (SASS instructions per stage) are taken from the ncu source view of k_extend
// (profiles/r01_ncu_src_extend_v8.txt) and can be overridden.
//
// Build: see tools/trav_sim.py (g++ with csrc/flatten.cpp + csrc/bvh_build.cpp; the GPU BVH builder is stubbed out).
#include <cstdio>
#include <cstring>
#include <cmath>
#include <cfloat>
#include <vector>
#include <string>
#include <algorithm>
#include "../include/spt.h"
#include "../simple-path-tracer_b200/csrc/flatten.h"

namespace spt {
bool build_cwbvh_device(const std::vector<Box3>&, int, std::vector<CwNode>&, std::vector<int>&, Box3*, std::string* err, int) {
  if (err) *err = "trav_sim: no GPU builder";
  return false;
}
} // namespace spt

namespace {

struct F3 { float x, y, z; };
inline F3 sub(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(F3 a, F3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline F3 cross(F3 a, F3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }

struct RaySetup { F3 o, d, idir; uint32_t octinv; };
inline float clamp_dir(float d) { float a = std::fmax(std::fabs(d), 1e-18f); return d < 0.f || (d == 0.f && std::signbit(d)) ? -a : a; }
inline void ray_setup(RaySetup& r, F3 o, F3 d) {
  r.o = o; r.d = d;
  r.idir = {1.0f / clamp_dir(d.x), 1.0f / clamp_dir(d.y), 1.0f / clamp_dir(d.z)};
  const uint32_t oct = ((uint32_t)std::signbit(d.x) << 2) | ((uint32_t)std::signbit(d.y) << 1) | (uint32_t)std::signbit(d.z);
  r.octinv = 7u - oct;
}

struct Group { uint32_t x, y; };

inline uint32_t bfind(uint32_t v) { return 31u - (uint32_t)__builtin_clz(v); }

// node_children() of traverse.cuh on the host
void node_children(const spt::CwNode& n, const RaySetup& r, float tmin, float tmax, Group& ng, Group& tg) {
  float s[3], a[3], b[3];
  const float o[3] = {r.o.x, r.o.y, r.o.z}, id[3] = {r.idir.x, r.idir.y, r.idir.z};
  for (int k = 0; k < 3; k++) { s[k] = std::ldexp(1.0f, n.e[k]); a[k] = s[k] * id[k]; b[k] = (n.p[k] - o[k]) * id[k]; }
  uint32_t hitmask = 0;
  for (int j = 0; j < 8; j++) {
    float t0[3], t1[3];
    for (int k = 0; k < 3; k++) {
      const bool neg = id[k] < 0.f;
      const float qn = (float)(neg ? n.qhi[k][j] : n.qlo[k][j]), qf = (float)(neg ? n.qlo[k][j] : n.qhi[k][j]);
      t0[k] = std::fma(qn, a[k], b[k]); t1[k] = std::fma(qf, a[k], b[k]);
    }
    const float cmin = std::fmax(std::fmax(t0[0], t0[1]), std::fmax(t0[2], tmin));
    const float cmax = std::fmin(std::fmin(t1[0], t1[1]), std::fmin(t1[2], tmax)) * 1.000002f;
    if (cmin <= cmax) {
      const uint32_t m = n.meta[j];
      const bool inner = (m & (m << 1) & 0x10u) != 0; // bits 3 and 4 both set: 24 + slot
      const uint32_t bit = inner ? ((m ^ r.octinv) & 0x1fu) : (m & 0x1fu);
      hitmask |= (m >> 5) << bit;
    }
  }
  ng.x = n.child_base; tg.x = n.prim_base;
  ng.y = (hitmask & 0xff000000u) | n.imask;
  tg.y = hitmask & 0x00ffffffu;
}

bool tri_test(const float4* tris, uint32_t idx, F3 o, F3 d, float& t) {
  const float4 a = tris[3 * (size_t)idx], b = tris[3 * (size_t)idx + 1], c = tris[3 * (size_t)idx + 2];
  const F3 p0{a.x, a.y, a.z}, e1{b.x, b.y, b.z}, e2{c.x, c.y, c.z};
  const F3 q = cross(d, e2);
  float det = dot(e1, q);
  if (det == 0.0f) return false;
  det = 1.0f / det;
  const F3 s = sub(o, p0);
  const float v = dot(s, q) * det;
  if (!(v >= 0.0f)) return false;
  const F3 rr = cross(s, e1);
  const float w = dot(d, rr) * det, u = (1.0f - v) - w;
  if (!(w >= 0.0f && u >= 0.0f)) return false;
  t = dot(e2, rr) * det;
  return true;
}

struct Lane {
  bool has = false;
  float tmin = 0, tmax = 0; // tmax = running hit.t
  RaySetup r, world;
  Group ng{0, 0}, tg{0, 0};
  std::vector<Group> stack;
  int blas_base = 0, cur_inst = -1;
  bool hit = false;
};

struct Sim {
  const spt::HostScene& hs;
  bool any;
  uint64_t n_nodes = 0, n_tris = 0, n_insts = 0;
  explicit Sim(const spt::HostScene& h, bool a) : hs(h), any(a) {}

  bool has_node(const Lane& L) const { return L.ng.y > 0x00ffffffu; }
  bool has_leaf(const Lane& L) const { return L.tg.y != 0u; }
  void begin(Lane& L, const SptRay& ray) {
    L.has = true; L.tmin = ray.t_min; L.tmax = ray.t_max; L.hit = false;
    ray_setup(L.r, {ray.origin[0], ray.origin[1], ray.origin[2]}, {ray.direction[0], ray.direction[1], ray.direction[2]});
    L.world = L.r;
    L.ng = {0u, 0x80000000u}; L.tg = {0u, 0u}; L.stack.clear(); L.blas_base = 0; L.cur_inst = -1;
  }
  void node(Lane& L) {
    const uint32_t bit = bfind(L.ng.y), base = L.ng.x, imask = L.ng.y;
    L.ng.y &= ~(1u << bit);
    if (L.tg.y != 0u) L.stack.push_back(L.tg); // eager mode only: leftover triangles wait on the stack
    if (L.ng.y > 0x00ffffffu) L.stack.push_back(L.ng);
    const uint32_t slot = (bit - 24u) ^ L.r.octinv;
    const uint32_t rel = (uint32_t)__builtin_popcount(imask & ~(0xffffffffu << slot));
    n_nodes++;
    node_children(hs.nodes[base + rel], L.r, L.tmin, L.tmax, L.ng, L.tg);
  }
  // one triangle test for lane L's ray; returns true when ANY found an occluder
  bool tri_one(Lane& L, uint32_t pidx) {
    n_tris++;
    float t;
    if (tri_test(hs.tris.data(), pidx, L.r.o, L.r.d, t) && t > L.tmin && t < L.tmax) {
      if (any) { L.hit = true; return true; }
      L.tmax = t; L.hit = true;
    }
    return false;
  }
  bool tri(Lane& L) {
    const uint32_t k = bfind(L.tg.y);
    L.tg.y &= ~(1u << k);
    return tri_one(L, L.tg.x + k);
  }
  bool inst(Lane& L) {
    const uint32_t k = bfind(L.tg.y);
    L.tg.y &= ~(1u << k);
    const int iid = hs.tlas_inst[L.tg.x + k];
    const DInstance& I = hs.inst[iid];
    n_insts++;
    const F3 o = L.r.o, d = L.r.d;
    const F3 o2{I.inv0.x * o.x + I.inv0.y * o.y + I.inv0.z * o.z + I.inv0.w, I.inv1.x * o.x + I.inv1.y * o.y + I.inv1.z * o.z + I.inv1.w,
                I.inv2.x * o.x + I.inv2.y * o.y + I.inv2.z * o.z + I.inv2.w};
    const F3 d2{I.inv0.x * d.x + I.inv0.y * d.y + I.inv0.z * d.z, I.inv1.x * d.x + I.inv1.y * d.y + I.inv1.z * d.z, I.inv2.x * d.x + I.inv2.y * d.y + I.inv2.z * d.z};
    if (I.kind == 0) { // sphere
      n_tris++;
      const float4 cr = hs.spheres[I.root];
      const F3 oc = sub(o2, {cr.x, cr.y, cr.z});
      const float a = dot(d2, d2), b = dot(d2, oc), c = dot(oc, oc) - cr.w * cr.w;
      float delta = b * b - a * c;
      if (delta >= 0.f) {
        delta = std::sqrt(delta);
        const float mn = (-b - delta) / a, mx = (-b + delta) / a;
        if (any) { if (mn < L.tmax && mx > L.tmin) { L.hit = true; return true; } }
        else { const float t = mn < L.tmin ? mx : mn; if (L.tmin < t && t < L.tmax) { L.tmax = t; L.hit = true; } }
      }
      return false;
    }
    if (L.tg.y != 0u) L.stack.push_back(L.tg);
    if (L.ng.y > 0x00ffffffu) L.stack.push_back(L.ng);
    L.blas_base = (int)L.stack.size(); L.cur_inst = iid;
    const bool small = (I.kind & 2) != 0;
    ray_setup(L.r, o2, d2);
    L.ng = {small ? 0u : (uint32_t)I.root, small ? 0u : 0x80000000u};
    L.tg = {small ? (uint32_t)I.root : 0u, small ? (1u << (I.kind >> 8)) - 1u : 0u};
    return false;
  }
  bool pop(Lane& L) {
    if (L.cur_inst >= 0 && (int)L.stack.size() == L.blas_base) { L.cur_inst = -1; L.r = L.world; }
    if (L.stack.empty()) return true;
    const Group e = L.stack.back(); L.stack.pop_back();
    if (e.y > 0x00ffffffu) { L.ng = e; L.tg = {0, 0}; } else { L.tg = e; L.ng = {0, 0}; }
    return false;
  }
};

} // namespace

extern "C" {

struct SimCfg {
  int32_t any;          // 1: any-hit (connect), 0: closest hit (extend)
  int32_t n_warps;      // warps simulated side by side (share the fetch cursor)
  int32_t chunk;        // rays a warp reserves at a time
  int32_t min_idle;     // refill threshold
  int32_t tri_per_round;
  int32_t tri_dist;     // 1: the pending triangle tests of the whole warp are dealt to all 32 lanes (one or more passes per round)
  int32_t node_twice;   // 1: a second node stage at the end of the round for lanes that have a node pending and no leaf
  int32_t inst_dist;    // EAGER node step: 1 = a lane with a node group pending takes its node step even when a triangle group is still pending
                        // (the leftover triangles go on the stack and are tested after the new subtree)
  int32_t slots;        // ray slots per warp (32 = one ray per lane in registers; 64 / 96: ray state in shared memory, every stage
                        // compacts the slots that need it onto the 32 lanes, up to max_pass passes of 32 per round)
  int32_t max_pass;     // passes per stage and round in the slot design
  int32_t pop_end;      // 1: a second pop stage at the END of the round (a ray that has nothing left finishes in the round of its last test)
  int32_t order;        // 0: node, inst, tri   1: inst, node, tri   2: inst, node, inst, tri
  int32_t gate_leaf;    // one-ray-per-lane design: the instance / triangle stages run only when at least this many lanes want them, or the
                        // stage was skipped in the previous gate_wait rounds
  int32_t gate_wait;
  int32_t gate;         // slot design: a pass of a stage only runs when at least this many slots need it (or nothing else can run)
  float c_round, c_pop, c_node, c_inst, c_tri, c_tri_dist; // warp-instructions per executed stage
  float c_tri_setup;    // tri_dist 1: once per round that has triangle work
  float c_slot;         // slot design: extra instructions per stage pass (compaction + shared-memory load / store of the ray state)
};
struct SimOut {
  double rounds_per_ray, lanes_node, lanes_inst, lanes_tri, lanes_pop, busy_frac, issue_per_ray, lane_util;
  double nodes_per_ray, tris_per_ray, insts_per_ray, hit_rate, tri_passes_per_round;
};

void* sim_scene_create(const SptSceneDesc* d) {
  setenv("SPT_BVH_BUILD", "host", 1);
  auto* hs = new spt::HostScene();
  const std::string err = spt::flatten_scene(*d, *hs);
  if (!err.empty()) { fprintf(stderr, "trav_sim: %s\n", err.c_str()); delete hs; return nullptr; }
  return hs;
}
void sim_scene_destroy(void* h) { delete (spt::HostScene*)h; }

// Closest hit (any = 0) or any hit (any = 1) of every ray through the compressed wide BVH, one ray at a time, in the stage order of the CUDA
// traversal: a CPU check of the host BVH builder and of the conservative 8-bit box quantisation (tests/test_host_bvh.py compares with the oracle).
int sim_trace(void* scene, const SptRay* rays, int64_t n, int32_t any, float* out_t, uint8_t* out_hit) {
  const spt::HostScene& hs = *(const spt::HostScene*)scene;
  Sim S(hs, any != 0);
  for (int64_t i = 0; i < n; i++) {
    Lane L;
    out_t[i] = rays[i].t_max; out_hit[i] = 0;
    if (hs.nodes.empty() || hs.inst.empty()) continue;
    S.begin(L, rays[i]);
    for (;;) {
      bool fin = false;
      if (S.has_node(L) && !S.has_leaf(L)) S.node(L);
      if (S.has_leaf(L) && L.cur_inst < 0) fin = S.inst(L);
      while (!fin && S.has_leaf(L) && L.cur_inst >= 0) fin = S.tri(L);
      if (!fin && !S.has_node(L) && !S.has_leaf(L)) fin = S.pop(L);
      if (fin) break;
    }
    out_t[i] = L.tmax; out_hit[i] = L.hit ? 1 : 0;
  }
  return 0;
}

int sim_run(void* scene, const SptRay* rays, int64_t n, const SimCfg* cfg, SimOut* out) {
  const spt::HostScene& hs = *(const spt::HostScene*)scene;
  Sim S(hs, cfg->any != 0);
  const int NW = std::max(1, cfg->n_warps);
  const int NS = std::max(32, cfg->slots);
  struct Warp { std::vector<Lane> lane; int64_t w_next = 0, w_end = 0; bool exhausted = false, done = false; int skipped_inst = 0, skipped_tri = 0; };
  std::vector<Warp> warps(NW);
  for (auto& W : warps) W.lane.resize(NS);
  int64_t cursor = 0;
  double issue = 0, useful = 0; // warp-instructions issued; sum over stages of cost * active/32
  uint64_t rounds = 0, st_node = 0, st_inst = 0, st_tri = 0, st_pop = 0, a_node = 0, a_inst = 0, a_tri = 0, a_pop = 0, busy = 0, hits = 0, tri_passes = 0;
  int alive = NW;
  const bool slot_design = NS > 32;
  const int max_pass = std::max(1, cfg->max_pass);
  std::vector<char> fin(NS);
  while (alive > 0) {
    for (Warp& W : warps) {
      if (W.done) continue;
      // ---- refill
      int idle = 0;
      for (auto& L : W.lane) idle += !L.has;
      if (idle == NS || idle >= cfg->min_idle) {
        if (W.w_next >= W.w_end && !W.exhausted) {
          const int64_t base = cursor; cursor += cfg->chunk;
          if (base >= n) W.exhausted = true; else { W.w_next = base; W.w_end = std::min<int64_t>(base + cfg->chunk, n); }
        }
        const int64_t take = std::min<int64_t>(idle, W.w_end - W.w_next);
        if (take == 0 && idle == NS) { W.done = true; alive--; continue; }
        int64_t k = 0;
        for (auto& L : W.lane) if (!L.has && k < take) { S.begin(L, rays[W.w_next + k]); k++; }
        W.w_next += take;
      }
      rounds++;
      double cost = cfg->c_round, use = 0;
      std::fill(fin.begin(), fin.end(), 0);
      for (auto& L : W.lane) busy += L.has;
      // a stage: `want(l)` selects the slots, `run(l)` executes one step; up to 32 slots per pass
      bool gated = false; // gating is in force this round only if SOME stage has at least `gate` slots waiting (else nothing would ever run)
      auto stage = [&](auto want, auto run, float c, uint64_t& st, uint64_t& a, int passes, bool gate_me) {
        for (int p = 0; p < passes; p++) {
          int act = 0;
          if (slot_design && gate_me && gated) {
            int cnt = 0; for (int m = 0; m < NS; m++) cnt += W.lane[m].has && !fin[m] && want(W.lane[m]);
            if (cnt < cfg->gate) return;
          }
          for (int l = 0; l < NS && act < 32; l++) if (W.lane[l].has && !fin[l] && want(W.lane[l])) {
            act++; if (run(W.lane[l])) fin[l] = 1;
          }
          if (!act) return;
          const float cc = c + (slot_design ? cfg->c_slot : 0.f);
          cost += cc; use += cc * act / 32.0; st++; a += act;
        }
      };
      auto w_pop = [&](const Lane& L) { return !S.has_node(L) && !S.has_leaf(L); };
      auto w_node = [&](const Lane& L) { return S.has_node(L) && (!S.has_leaf(L) || (cfg->inst_dist == 1 && L.cur_inst >= 0)); };
      auto w_inst = [&](const Lane& L) { return S.has_leaf(L) && L.cur_inst < 0; };
      auto w_tri = [&](const Lane& L) { return S.has_leaf(L) && L.cur_inst >= 0; };
      const int P = slot_design ? max_pass : 1;
      if (slot_design && cfg->gate > 1) {
        int cn = 0, ci = 0, ct = 0;
        for (int m = 0; m < NS; m++) if (W.lane[m].has) { cn += w_node(W.lane[m]) || w_pop(W.lane[m]); ci += w_inst(W.lane[m]); ct += w_tri(W.lane[m]); }
        gated = std::max(cn, std::max(ci, ct)) >= cfg->gate;
      }
      stage(w_pop, [&](Lane& L) { return S.pop(L); }, cfg->c_pop, st_pop, a_pop, slot_design ? 2 : 1, false);
      auto count_want = [&](auto want) { int c = 0; for (int m = 0; m < NS; m++) c += W.lane[m].has && !fin[m] && want(W.lane[m]); return c; };
      bool run_inst = true, run_tri = true;
      if (!slot_design && cfg->gate_leaf > 1) {
        const int ci = count_want(w_inst);
        run_inst = ci >= cfg->gate_leaf || W.skipped_inst >= cfg->gate_wait || ci == 0;
        W.skipped_inst = run_inst ? 0 : W.skipped_inst + 1;
      }
      if (cfg->order >= 1) stage(w_inst, [&](Lane& L) { return S.inst(L); }, cfg->c_inst, st_inst, a_inst, P, true);
      stage(w_node, [&](Lane& L) { S.node(L); return false; }, cfg->c_node, st_node, a_node, P, true);
      if (cfg->order != 1 && run_inst) stage(w_inst, [&](Lane& L) { return S.inst(L); }, cfg->c_inst, st_inst, a_inst, P, true);
      if (!slot_design && cfg->gate_leaf > 1) {
        const int ct = count_want(w_tri);
        run_tri = ct >= cfg->gate_leaf || W.skipped_tri >= cfg->gate_wait || ct == 0;
        W.skipped_tri = run_tri ? 0 : W.skipped_tri + 1;
      }
      if (cfg->tri_dist == 2) {
        // helper passes: every lane with pending triangles tests its first one; lanes WITHOUT triangle work take the second triangle of
        // the i-th lane that has two or more (ballot matching, ray handed over by shuffles).  tri_per_round passes at most.
        for (int p = 0; p < cfg->tri_per_round; p++) {
          int owners = 0, helpers_avail = 0, helped = 0;
          for (int l = 0; l < NS; l++) { Lane& L = W.lane[l]; const bool w = L.has && !fin[l] && w_tri(L); owners += w; helpers_avail += !w; }
          if (!owners) break;
          for (int l = 0; l < NS; l++) { Lane& L = W.lane[l]; if (L.has && !fin[l] && w_tri(L)) {
            if (S.tri(L)) { fin[l] = 1; continue; }
            if (L.tg.y != 0u && helped < helpers_avail) { helped++; if (S.tri(L)) fin[l] = 1; }
          } }
          cost += cfg->c_tri_dist; use += cfg->c_tri_dist * (owners + helped) / 32.0; st_tri++; a_tri += owners + helped; tri_passes++;
        }
      } else if (!cfg->tri_dist) {
        if (run_tri) stage(w_tri, [&](Lane& L) { return S.tri(L); }, cfg->c_tri, st_tri, a_tri, slot_design ? P * 2 : cfg->tri_per_round, true);
      } else {
        std::vector<std::pair<int, uint32_t>> work;
        for (int l = 0; l < NS; l++) { Lane& L = W.lane[l]; if (L.has && !fin[l] && w_tri(L)) {
          uint32_t m = L.tg.y; while (m) { const uint32_t k = bfind(m); m &= ~(1u << k); work.push_back({l, L.tg.x + k}); } L.tg.y = 0; } }
        for (size_t b = 0; b < work.size(); b += 32) {
          int a = 0;
          const size_t e = std::min(work.size(), b + 32);
          for (size_t i = b; i < e; i++) { Lane& L = W.lane[work[i].first]; if (fin[work[i].first]) continue; a++; if (S.tri_one(L, work[i].second)) fin[work[i].first] = 1; }
          if (a) { cost += cfg->c_tri_dist; use += cfg->c_tri_dist * a / 32.0; st_tri++; a_tri += a; tri_passes++; }
        }
        if (!work.empty()) cost += cfg->c_tri_setup; // prefix sum + scatter of the work list, once per round
      }
      if (cfg->node_twice) stage(w_node, [&](Lane& L) { S.node(L); return false; }, cfg->c_node, st_node, a_node, 1, true);
      if (cfg->pop_end) stage(w_pop, [&](Lane& L) { return S.pop(L); }, cfg->c_pop, st_pop, a_pop, 1, false);
      for (int l = 0; l < NS; l++) if (fin[l]) { hits += W.lane[l].hit; W.lane[l].has = false; }
      issue += cost; useful += use;
    }
  }
  out->rounds_per_ray = (double)busy / (double)n; // lane-rounds with a ray, per ray
  out->busy_frac = (double)busy / ((double)NS * rounds);
  out->lanes_node = st_node ? (double)a_node / st_node : 0; out->lanes_inst = st_inst ? (double)a_inst / st_inst : 0;
  out->lanes_tri = st_tri ? (double)a_tri / st_tri : 0; out->lanes_pop = st_pop ? (double)a_pop / st_pop : 0;
  out->issue_per_ray = issue / (double)n; out->lane_util = 32.0 * useful / issue;
  out->nodes_per_ray = (double)S.n_nodes / n; out->tris_per_ray = (double)S.n_tris / n; out->insts_per_ray = (double)S.n_insts / n;
  out->hit_rate = (double)hits / n; out->tri_passes_per_round = (double)tri_passes / rounds;
  return 0;
}

} // extern "C"
