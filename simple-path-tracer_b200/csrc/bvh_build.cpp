// bvh_build.cpp — host-side acceleration-structure build.
//
// Replaces BvhAccel::new (src/primitive/bvh.rs:23-215).  The reference's builder is NOT reproduced (its bucket
// rule and "surface area" are broken — SURVEY.md Q1/Q2 — and e.g. leave bunny_5k as one 5000-triangle leaf);
// closest-hit results do not depend on the tree, only on the leaf tests, so a proper tree is a drop-in.
//   1. binary BVH by 16-bin surface-area heuristic over primitive boxes (leaves <= max_leaf primitives)
//   2. greedy collapse to 8-wide nodes (open the child with the largest area until 8 children)
//   3. compression to 80-byte nodes with 8-bit quantised child boxes and octant-ordered child slots
//      (Ylitie, Karras, Laine: "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017)
#include "bvh_build.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <cfloat>

namespace spt {

namespace {

struct B2 { // binary node
  Box3 box;
  int left = -1, right = -1; // children, or -1 for a leaf
  int first = 0, count = 0;  // primitive range for a leaf
};

inline float area(const Box3& b) {
  float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  if (dx < 0 || dy < 0 || dz < 0) return 0.f;
  return 2.f * (dx * dy + dy * dz + dz * dx);
}
inline void grow(Box3& a, const Box3& b) {
  for (int k = 0; k < 3; k++) { a.lo[k] = std::min(a.lo[k], b.lo[k]); a.hi[k] = std::max(a.hi[k], b.hi[k]); }
}
inline Box3 empty_box() { Box3 b; for (int k = 0; k < 3; k++) { b.lo[k] = FLT_MAX; b.hi[k] = -FLT_MAX; } return b; }

struct Builder {
  const std::vector<Box3>& boxes;
  std::vector<int> order;      // primitive permutation
  std::vector<float> cent;     // 3 per primitive
  std::vector<B2> nodes;
  int max_leaf;

  Builder(const std::vector<Box3>& b, int ml) : boxes(b), max_leaf(ml) {
    order.resize(b.size());
    cent.resize(3 * b.size());
    for (size_t i = 0; i < b.size(); i++) {
      order[i] = (int)i;
      for (int k = 0; k < 3; k++) cent[3 * i + k] = 0.5f * (b[i].lo[k] + b[i].hi[k]);
    }
  }

  int build(int first, int count) {
    int id = (int)nodes.size();
    nodes.emplace_back();
    Box3 nb = empty_box(), cb = empty_box();
    for (int i = first; i < first + count; i++) {
      grow(nb, boxes[order[i]]);
      for (int k = 0; k < 3; k++) { float c = cent[3 * order[i] + k]; cb.lo[k] = std::min(cb.lo[k], c); cb.hi[k] = std::max(cb.hi[k], c); }
    }
    nodes[id].box = nb;
    if (count <= max_leaf) { nodes[id].first = first; nodes[id].count = count; return id; }
    // binned SAH
    const int NB = 16;
    float best_cost = FLT_MAX; int best_axis = -1, best_bin = -1;
    for (int ax = 0; ax < 3; ax++) {
      float ext = cb.hi[ax] - cb.lo[ax];
      if (!(ext > 0.f)) continue;
      Box3 bb[NB]; int bc[NB];
      for (int j = 0; j < NB; j++) { bb[j] = empty_box(); bc[j] = 0; }
      float scale = NB / ext;
      for (int i = first; i < first + count; i++) {
        int p = order[i];
        int j = std::min(NB - 1, std::max(0, (int)((cent[3 * p + ax] - cb.lo[ax]) * scale)));
        grow(bb[j], boxes[p]); bc[j]++;
      }
      float ra[NB]; Box3 acc = empty_box();
      for (int j = NB - 1; j > 0; j--) { grow(acc, bb[j]); ra[j] = area(acc); }
      acc = empty_box(); int lc = 0;
      for (int j = 0; j < NB - 1; j++) {
        grow(acc, bb[j]); lc += bc[j];
        int rc = count - lc;
        if (lc == 0 || rc == 0) continue;
        float cost = area(acc) * lc + ra[j + 1] * rc;
        if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = j; }
      }
    }
    int mid;
    if (best_axis >= 0) {
      float ext = cb.hi[best_axis] - cb.lo[best_axis], scale = NB / ext, lo = cb.lo[best_axis];
      int ax = best_axis, bin = best_bin;
      auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](int p) {
        int j = std::min(NB - 1, std::max(0, (int)((cent[3 * p + ax] - lo) * scale)));
        return j <= bin;
      });
      mid = (int)(it - order.begin());
    } else mid = first; // all centroids coincide
    if (mid == first || mid == first + count) { // fall back to an object-median split
      mid = first + count / 2;
      int ax = 0; float e = -1.f;
      for (int k = 0; k < 3; k++) if (nb.hi[k] - nb.lo[k] > e) { e = nb.hi[k] - nb.lo[k]; ax = k; }
      std::nth_element(order.begin() + first, order.begin() + mid, order.begin() + first + count,
                       [&](int a, int b) { return cent[3 * a + ax] < cent[3 * b + ax]; });
    }
    int l = build(first, mid - first);
    int r = build(mid, first + count - mid);
    nodes[id].left = l; nodes[id].right = r;
    return id;
  }
};

inline float exp2i(int e) { return std::ldexp(1.0f, e); }

} // namespace

void build_cwbvh(const std::vector<Box3>& prim_boxes, int max_leaf, uint32_t node_base, uint32_t prim_base,
                 std::vector<CwNode>& out_nodes, std::vector<int>& out_prim_order, Box3* root_box) {
  out_nodes.clear(); out_prim_order.clear();
  if (root_box) *root_box = empty_box();
  if (prim_boxes.empty()) return;
  max_leaf = std::max(1, std::min(3, max_leaf));
  Builder B(prim_boxes, max_leaf);
  int root = B.build(0, (int)prim_boxes.size());
  if (root_box) *root_box = B.nodes[root].box;

  struct Item { int b2; uint32_t out; };
  std::vector<Item> queue;
  out_nodes.emplace_back();
  queue.push_back({root, 0});
  for (size_t qi = 0; qi < queue.size(); qi++) {
    const Item it = queue[qi];
    const B2& N = B.nodes[it.b2];
    // --- collapse: gather up to 8 children
    int ch[8]; int nc = 0;
    if (N.left < 0) ch[nc++] = it.b2; // a leaf root: one leaf child
    else {
      ch[nc++] = N.left; ch[nc++] = N.right;
      while (nc < 8) {
        int best = -1; float ba = -1.f;
        for (int i = 0; i < nc; i++) if (B.nodes[ch[i]].left >= 0) { float a = area(B.nodes[ch[i]].box); if (a > ba) { ba = a; best = i; } }
        if (best < 0) break;
        int c = ch[best];
        ch[best] = B.nodes[c].left; ch[nc++] = B.nodes[c].right;
      }
    }
    // --- octant-ordered slot assignment (greedy on cost = (centroid_c - centroid_node) . dir_s)
    float pc[3]; for (int k = 0; k < 3; k++) pc[k] = 0.5f * (N.box.lo[k] + N.box.hi[k]);
    float cost[8][8]; int slot_of[8]; int child_in[8];
    for (int s = 0; s < 8; s++) child_in[s] = -1;
    for (int c = 0; c < nc; c++) {
      slot_of[c] = -1;
      const Box3& cb = B.nodes[ch[c]].box;
      for (int s = 0; s < 8; s++) {
        float d = 0.f;
        for (int k = 0; k < 3; k++) { float sg = ((s >> (2 - k)) & 1) ? -1.f : 1.f; d += (0.5f * (cb.lo[k] + cb.hi[k]) - pc[k]) * sg; }
        cost[c][s] = d;
      }
    }
    for (int n = 0; n < nc; n++) {
      float bm = FLT_MAX; int bc = -1, bs = -1;
      for (int c = 0; c < nc; c++) if (slot_of[c] < 0) for (int s = 0; s < 8; s++) if (child_in[s] < 0 && cost[c][s] < bm) { bm = cost[c][s]; bc = c; bs = s; }
      slot_of[bc] = bs; child_in[bs] = bc;
    }
    // --- quantisation frame
    CwNode cw; memset(&cw, 0, sizeof(cw));
    int8_t e[3];
    for (int k = 0; k < 3; k++) {
      float ext = N.box.hi[k] - N.box.lo[k];
      int ek = -127;
      if (ext > 0.f) {
        ek = (int)std::ceil(std::log2((double)ext / 255.0));
        while ((double)exp2i(ek) * 255.0 < (double)ext) ek++;
        ek = std::max(-126, std::min(127, ek));
      }
      e[k] = (int8_t)ek;
      cw.p[k] = N.box.lo[k];
    }
    uint8_t imask = 0;
    uint32_t child_base = (uint32_t)out_nodes.size();
    uint32_t prim_first = (uint32_t)out_prim_order.size();
    uint32_t n_inner = 0, n_prims = 0;
    for (int s = 0; s < 8; s++) {
      int c = child_in[s];
      if (c < 0) { // empty slot: inverted box, no bits
        for (int k = 0; k < 3; k++) { cw.qlo[k][s] = 255; cw.qhi[k][s] = 0; }
        cw.meta[s] = 0;
        continue;
      }
      const B2& C = B.nodes[ch[c]];
      for (int k = 0; k < 3; k++) {
        double sc = e[k] == -127 ? 0.0 : (double)exp2i(e[k]);
        int lo = 0, hi = 0;
        if (sc > 0.0) {
          lo = (int)std::floor(((double)C.box.lo[k] - (double)cw.p[k]) / sc);
          hi = (int)std::ceil(((double)C.box.hi[k] - (double)cw.p[k]) / sc);
          lo = std::max(0, std::min(255, lo)); hi = std::max(0, std::min(255, hi));
          while (lo > 0 && (double)cw.p[k] + lo * sc > (double)C.box.lo[k]) lo--;
          while (hi < 255 && (double)cw.p[k] + hi * sc < (double)C.box.hi[k]) hi++;
        }
        cw.qlo[k][s] = (uint8_t)lo; cw.qhi[k][s] = (uint8_t)hi;
      }
      if (C.left >= 0) {
        imask |= (uint8_t)(1u << s);
        cw.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
        n_inner++;
      } else {
        uint32_t unary = C.count == 1 ? 1u : (C.count == 2 ? 3u : 7u);
        cw.meta[s] = (uint8_t)((unary << 5) | n_prims);
        for (int i = 0; i < C.count; i++) out_prim_order.push_back(B.order[C.first + i]);
        n_prims += (uint32_t)C.count;
      }
    }
    cw.e[0] = e[0]; cw.e[1] = e[1]; cw.e[2] = e[2]; cw.imask = imask;
    cw.child_base = node_base + child_base;
    cw.prim_base = prim_base + prim_first;
    // inner children get consecutive nodes in slot order
    out_nodes.resize(out_nodes.size() + n_inner);
    uint32_t r = 0;
    for (int s = 0; s < 8; s++) {
      int c = child_in[s];
      if (c >= 0 && B.nodes[ch[c]].left >= 0) queue.push_back({ch[c], child_base + r++});
    }
    out_nodes[it.out] = cw;
  }
}

} // namespace spt
