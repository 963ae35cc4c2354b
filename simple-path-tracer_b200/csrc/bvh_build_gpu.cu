// bvh_build_gpu.cu — on-device build of the compressed 8-wide BVH for large meshes (replaces the host
// flatten's SAH build where that dominates start-up; the reference analogue is BvhAccel::new, src/primitive/bvh.rs:23-215).
//
//   1. 63-bit Morton codes of the primitive centroids                                k_morton
//   2. radix sort of (code, primitive)                                               cub::DeviceRadixSort
//   3. binary tree over the sorted primitives, either
//        LBVH: radix tree (Karras, HPG 2012), one thread per node + bottom-up refit  k_radix_tree, k_refit
//        PLOC: rounds of nearest-neighbour search in a Morton window + merge of mutual pairs + compaction
//              (Meister & Bittner, TVCG 2018)                                        k_ploc_nn, k_ploc_merge, k_ploc_compact
//   4. level-by-level collapse to 8-wide nodes (open the largest child until 8), octant-ordered slots, 8-bit quantised
//      child boxes; nodes and primitive ranges are allocated with atomics           k_collapse
//
// The output has exactly the format of the host builder (bvh_build.cpp): same 80-byte nodes, inner children consecutive
// in slot order, the primitives of a node's leaf slots consecutive in leaf order.  Closest-hit results do not depend on
// the tree, so the two builders are interchangeable; the LBVH is built ~100x faster and traverses a little slower
// (numbers in profiles/README.md), which is why flatten() only uses it for big meshes.
#include "bvh_build.h"

#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cfloat>
#include <cstdio>
#include <string>

namespace spt {
namespace {

#define BG_CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { if (err) *err = std::string(#x) + ": " + cudaGetErrorString(e_); cleanup(); return false; } } while (0)

__device__ __forceinline__ uint64_t spread21(uint32_t v) { // 21 bits -> every third bit
  uint64_t x = v & 0x1fffffu;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

__global__ void k_morton(const Box3* __restrict__ boxes, int n, float3 lo, float3 inv, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Box3 b = boxes[i];
  const float cx = (0.5f * (b.lo[0] + b.hi[0]) - lo.x) * inv.x, cy = (0.5f * (b.lo[1] + b.hi[1]) - lo.y) * inv.y, cz = (0.5f * (b.lo[2] + b.hi[2]) - lo.z) * inv.z;
  const uint32_t qx = (uint32_t)fminf(fmaxf(cx * 2097152.0f, 0.0f), 2097151.0f), qy = (uint32_t)fminf(fmaxf(cy * 2097152.0f, 0.0f), 2097151.0f),
                 qz = (uint32_t)fminf(fmaxf(cz * 2097152.0f, 0.0f), 2097151.0f);
  keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
  vals[i] = (uint32_t)i;
}

// length of the common prefix of the keys at sorted positions i and j; equal keys are told apart by their position
__device__ __forceinline__ int delta(const uint64_t* __restrict__ keys, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint64_t a = keys[i], b = keys[j];
  if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
  return __clzll((long long)(a ^ b));
}

// Binary nodes: internal 0..n-2, leaf j = n-1+j.  Root = internal 0.
__global__ void k_radix_tree(const uint64_t* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right, int* __restrict__ parent,
                             int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = delta(keys, n, i, i - d);
  int lmax = 2;
  while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
  int l = 0;
  for (int t = lmax / 2; t >= 1; t /= 2) if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = delta(keys, n, i, j);
  int s = 0;
  for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
    if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    if (t == 1) break;
  }
  const int g = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const int lc = lo == g ? n - 1 + g : g, rc = hi == g + 1 ? n - 1 + g + 1 : g + 1;
  left[i] = lc; right[i] = rc; count[i] = hi - lo + 1;
  parent[lc] = i; parent[rc] = i;
  if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ void box_union(Box3& a, const Box3& b) {
#pragma unroll
  for (int k = 0; k < 3; k++) { a.lo[k] = fminf(a.lo[k], b.lo[k]); a.hi[k] = fmaxf(a.hi[k], b.hi[k]); }
}
__device__ __forceinline__ float box_area(const Box3& b) {
  const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
  if (dx < 0 || dy < 0 || dz < 0) return 0.f;
  return 2.f * (dx * dy + dy * dz + dz * dx);
}

__device__ __forceinline__ Box3 load_box_cg(const Box3* p) {
  Box3 b;
#pragma unroll
  for (int k = 0; k < 3; k++) { b.lo[k] = __ldcg(&p->lo[k]); b.hi[k] = __ldcg(&p->hi[k]); }
  return b;
}

__global__ void k_refit(const Box3* __restrict__ prim_boxes, const uint32_t* __restrict__ vals, int n, const int* __restrict__ left,
                        const int* __restrict__ right, const int* __restrict__ parent, int* __restrict__ arrived, Box3* __restrict__ nbox) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  nbox[n - 1 + j] = prim_boxes[vals[j]];
  __threadfence();
  int node = parent[n - 1 + j];
  while (node >= 0) {
    if (atomicAdd(&arrived[node], 1) == 0) return; // the sibling subtree is not done yet: its last thread will come through here
    __threadfence();
    Box3 b = load_box_cg(nbox + left[node]); // written by another SM: read through L2, not a possibly stale L1 line
    box_union(b, load_box_cg(nbox + right[node]));
    nbox[node] = b;
    __threadfence();
    node = parent[node];
  }
}


// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner, TVCG 2018) ---------------------------------------------
// Clusters stay in Morton order.  Every round each cluster looks SPT_PLOC_RADIUS positions to either side for the neighbour
// whose union with it has the smallest surface area; pairs that chose each other are merged into a new inner node, the rest
// wait; a prefix sum compacts the survivors.  Near-SAH quality at a few times the cost of the radix tree.
#ifndef SPT_PLOC_RADIUS
#define SPT_PLOC_RADIUS 16
#endif

__global__ void k_ploc_init(const Box3* __restrict__ prim_boxes, const uint32_t* __restrict__ vals, int n, int* __restrict__ cid, Box3* __restrict__ cbox,
                            Box3* __restrict__ nbox) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const Box3 b = prim_boxes[vals[j]];
  cid[j] = n - 1 + j; cbox[j] = b; nbox[n - 1 + j] = b;
}

__global__ void k_ploc_nn(const Box3* __restrict__ cbox, int m, int* __restrict__ nn) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const Box3 bi = cbox[i];
  float best = FLT_MAX; int bj = -1;
  const int j0 = max(0, i - SPT_PLOC_RADIUS), j1 = min(m - 1, i + SPT_PLOC_RADIUS);
  for (int j = j0; j <= j1; j++) {
    if (j == i) continue;
    Box3 u = cbox[j];
    box_union(u, bi);
    const float a = box_area(u);
    if (a < best) { best = a; bj = j; } // ties: the lower position wins on both sides of a pair, so mutual choices stay mutual
  }
  nn[i] = bj;
}

__global__ void k_ploc_merge(const int* __restrict__ cid, const Box3* __restrict__ cbox, int m, const int* __restrict__ nn, int* __restrict__ left,
                             int* __restrict__ right, int* __restrict__ count, Box3* __restrict__ nbox, unsigned int* node_counter, int n,
                             int* __restrict__ cid_tmp, Box3* __restrict__ cbox_tmp, int* __restrict__ valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int j = nn[i];
  if (j >= 0 && nn[j] == i) {
    if (i < j) {
      const int id = (int)atomicAdd(node_counter, 1u);
      const int a = cid[i], b = cid[j];
      Box3 u = cbox[i];
      box_union(u, cbox[j]);
      left[id] = a; right[id] = b;
      count[id] = (a >= n - 1 ? 1 : count[a]) + (b >= n - 1 ? 1 : count[b]);
      nbox[id] = u;
      cid_tmp[i] = id; cbox_tmp[i] = u; valid[i] = 1;
    } else valid[i] = 0; // absorbed by its partner
  } else { cid_tmp[i] = cid[i]; cbox_tmp[i] = cbox[i]; valid[i] = 1; }
}

__global__ void k_ploc_compact(const int* __restrict__ cid_tmp, const Box3* __restrict__ cbox_tmp, const int* __restrict__ valid, const int* __restrict__ pos,
                               int m, int* __restrict__ cid, Box3* __restrict__ cbox) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m || !valid[i]) return;
  cid[pos[i]] = cid_tmp[i]; cbox[pos[i]] = cbox_tmp[i];
}

struct CollapseArgs {
  int n, max_leaf;
  const int *left, *right, *count; // count[x] = primitives under inner node x
  const Box3* nbox;
  const uint32_t* vals;
  CwNode* nodes;
  int* order;
  unsigned int *node_counter, *prim_counter, *out_counter;
  const uint2* items_in;
  uint2* items_out;
  int n_in;
};

__device__ __forceinline__ bool is_leaf(const CollapseArgs& A, int x) { return x >= A.n - 1 || A.count[x] <= A.max_leaf; }
__device__ __forceinline__ int leaf_count(const CollapseArgs& A, int x) { return x >= A.n - 1 ? 1 : A.count[x]; }
// the (at most 3) primitives under a leaf-ified subtree, left to right
__device__ __forceinline__ int leaf_prims(const CollapseArgs& A, int x, int out[3]) {
  int stack[4], sp = 0, k = 0;
  stack[sp++] = x;
  while (sp > 0) {
    const int y = stack[--sp];
    if (y >= A.n - 1) { if (k < 3) out[k++] = (int)A.vals[y - (A.n - 1)]; }
    else { stack[sp++] = A.right[y]; stack[sp++] = A.left[y]; }
  }
  return k;
}

__global__ void k_collapse(const CollapseArgs A) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= A.n_in) return;
  const int b2 = (int)A.items_in[w].x;
  const uint32_t out = A.items_in[w].y;
  const Box3 N = A.nbox[b2];
  // ---- gather up to 8 children: open the inner child with the largest box until there are 8
  int ch[8]; int nc = 0;
  if (is_leaf(A, b2)) ch[nc++] = b2;
  else {
    ch[nc++] = A.left[b2]; ch[nc++] = A.right[b2];
    while (nc < 8) {
      int best = -1; float ba = -1.f;
      for (int i = 0; i < nc; i++) if (!is_leaf(A, ch[i])) { const float a = box_area(A.nbox[ch[i]]); if (a > ba) { ba = a; best = i; } }
      if (best < 0) break;
      const int c = ch[best];
      ch[best] = A.left[c]; ch[nc++] = A.right[c];
    }
  }
  // ---- octant-ordered slots: greedy on (child centre - node centre) . direction of slot s
  float pc[3];
  for (int k = 0; k < 3; k++) pc[k] = 0.5f * (N.lo[k] + N.hi[k]);
  float cost[8][8]; int slot_of[8], child_in[8];
  for (int s = 0; s < 8; s++) child_in[s] = -1;
  for (int c = 0; c < nc; c++) {
    slot_of[c] = -1;
    const Box3 cb = A.nbox[ch[c]];
    for (int s = 0; s < 8; s++) {
      float d = 0.f;
      for (int k = 0; k < 3; k++) { const float sg = ((s >> (2 - k)) & 1) ? -1.f : 1.f; d += (0.5f * (cb.lo[k] + cb.hi[k]) - pc[k]) * sg; }
      cost[c][s] = d;
    }
  }
  for (int it = 0; it < nc; it++) {
    float bm = FLT_MAX; int bc = -1, bs = -1;
    for (int c = 0; c < nc; c++) if (slot_of[c] < 0) for (int s = 0; s < 8; s++) if (child_in[s] < 0 && cost[c][s] < bm) { bm = cost[c][s]; bc = c; bs = s; }
    if (bc < 0) { for (int c = 0; c < nc && bc < 0; c++) if (slot_of[c] < 0) bc = c; for (int s = 0; s < 8 && bs < 0; s++) if (child_in[s] < 0) bs = s; } // non-finite costs
    slot_of[bc] = bs; child_in[bs] = bc;
  }
  // ---- quantisation frame
  CwNode cw;
  memset(&cw, 0, sizeof(cw));
  int8_t e[3];
  for (int k = 0; k < 3; k++) {
    const float ext = N.hi[k] - N.lo[k];
    int ek = -127;
    if (ext > 0.f) {
      ek = (int)ceil(log2((double)ext / 255.0));
      while ((double)ldexpf(1.0f, ek) * 255.0 < (double)ext) ek++;
      ek = max(-126, min(127, ek));
    }
    e[k] = (int8_t)ek;
    cw.p[k] = N.lo[k];
  }
  uint32_t n_inner = 0, n_prims = 0;
  for (int c = 0; c < nc; c++) { if (is_leaf(A, ch[c])) n_prims += (uint32_t)leaf_count(A, ch[c]); else n_inner++; }
  const uint32_t child_base = n_inner ? atomicAdd(A.node_counter, n_inner) : 0u;
  const uint32_t prim_first = n_prims ? atomicAdd(A.prim_counter, n_prims) : 0u;
  const uint32_t item_base = n_inner ? atomicAdd(A.out_counter, n_inner) : 0u;
  uint8_t imask = 0;
  uint32_t r = 0, po = 0;
  for (int s = 0; s < 8; s++) {
    const int c = child_in[s];
    if (c < 0) { // empty slot: inverted box, no bits
      for (int k = 0; k < 3; k++) { cw.qlo[k][s] = 255; cw.qhi[k][s] = 0; }
      cw.meta[s] = 0;
      continue;
    }
    const Box3 C = A.nbox[ch[c]];
    for (int k = 0; k < 3; k++) {
      const double sc = e[k] == -127 ? 0.0 : (double)ldexpf(1.0f, e[k]);
      int lo = 0, hi = 0;
      if (sc > 0.0) {
        lo = (int)floor(((double)C.lo[k] - (double)cw.p[k]) / sc);
        hi = (int)ceil(((double)C.hi[k] - (double)cw.p[k]) / sc);
        lo = max(0, min(255, lo)); hi = max(0, min(255, hi));
        while (lo > 0 && (double)cw.p[k] + lo * sc > (double)C.lo[k]) lo--;
        while (hi < 255 && (double)cw.p[k] + hi * sc < (double)C.hi[k]) hi++;
      }
      cw.qlo[k][s] = (uint8_t)lo; cw.qhi[k][s] = (uint8_t)hi;
    }
    if (!is_leaf(A, ch[c])) {
      imask |= (uint8_t)(1u << s);
      cw.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
      A.items_out[item_base + r] = make_uint2((uint32_t)ch[c], child_base + r); // inner children: consecutive nodes in slot order
      r++;
    } else {
      int prims[3];
      const int cnt = leaf_prims(A, ch[c], prims);
      const uint32_t unary = cnt == 1 ? 1u : (cnt == 2 ? 3u : 7u);
      cw.meta[s] = (uint8_t)((unary << 5) | po);
      for (int i = 0; i < cnt; i++) A.order[prim_first + po + i] = prims[i];
      po += (uint32_t)cnt;
    }
  }
  cw.e[0] = e[0]; cw.e[1] = e[1]; cw.e[2] = e[2]; cw.imask = imask;
  cw.child_base = child_base;
  cw.prim_base = prim_first;
  A.nodes[out] = cw;
}

} // namespace

bool build_cwbvh_device(const std::vector<Box3>& prim_boxes, int max_leaf, std::vector<CwNode>& out_nodes, std::vector<int>& out_prim_order,
                        Box3* root_box, std::string* err, int algorithm) {
  out_nodes.clear(); out_prim_order.clear();
  const int n = (int)prim_boxes.size();
  if (n < 2) { if (err) *err = "build_cwbvh_device: needs at least 2 primitives"; return false; }
  max_leaf = std::max(1, std::min(3, max_leaf));
  // centroid bounds (host, one pass)
  float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  for (const Box3& b : prim_boxes) for (int k = 0; k < 3; k++) { const float c = 0.5f * (b.lo[k] + b.hi[k]); lo[k] = std::min(lo[k], c); hi[k] = std::max(hi[k], c); }
  float inv[3];
  for (int k = 0; k < 3; k++) inv[k] = hi[k] > lo[k] ? 1.0f / (hi[k] - lo[k]) : 0.0f;

  Box3 *d_boxes = nullptr, *d_nbox = nullptr;
  uint64_t *d_keys = nullptr, *d_keys2 = nullptr;
  uint32_t *d_vals = nullptr, *d_vals2 = nullptr;
  int *d_left = nullptr, *d_right = nullptr, *d_parent = nullptr, *d_count = nullptr, *d_arrived = nullptr, *d_order = nullptr;
  int *d_cid[2] = {nullptr, nullptr}, *d_nn = nullptr, *d_valid = nullptr, *d_pos = nullptr;
  Box3 *d_cbox[2] = {nullptr, nullptr};
  CwNode* d_nodes = nullptr;
  uint2 *d_items[2] = {nullptr, nullptr};
  unsigned int* d_ctr = nullptr; // node, prim, out
  void* d_tmp = nullptr;
  auto cleanup = [&]() {
    cudaFree(d_boxes); cudaFree(d_nbox); cudaFree(d_keys); cudaFree(d_keys2); cudaFree(d_vals); cudaFree(d_vals2); cudaFree(d_left); cudaFree(d_right);
    cudaFree(d_parent); cudaFree(d_count); cudaFree(d_arrived); cudaFree(d_cid[0]); cudaFree(d_cid[1]); cudaFree(d_nn); cudaFree(d_valid); cudaFree(d_pos); cudaFree(d_cbox[0]); cudaFree(d_cbox[1]); cudaFree(d_order); cudaFree(d_nodes); cudaFree(d_items[0]); cudaFree(d_items[1]);
    cudaFree(d_ctr); cudaFree(d_tmp);
  };
  const size_t N = (size_t)n;
  BG_CU(cudaMalloc(&d_boxes, N * sizeof(Box3)));
  BG_CU(cudaMalloc(&d_nbox, (2 * N - 1) * sizeof(Box3)));
  BG_CU(cudaMalloc(&d_keys, N * 8)); BG_CU(cudaMalloc(&d_keys2, N * 8));
  BG_CU(cudaMalloc(&d_vals, N * 4)); BG_CU(cudaMalloc(&d_vals2, N * 4));
  BG_CU(cudaMalloc(&d_left, N * 4)); BG_CU(cudaMalloc(&d_right, N * 4)); BG_CU(cudaMalloc(&d_count, N * 4));
  BG_CU(cudaMalloc(&d_parent, (2 * N - 1) * 4)); BG_CU(cudaMalloc(&d_arrived, N * 4)); BG_CU(cudaMalloc(&d_order, N * 4));
  BG_CU(cudaMalloc(&d_nodes, N * sizeof(CwNode))); // a wide node consumes at least one binary inner node: < n of them
  BG_CU(cudaMalloc(&d_items[0], N * sizeof(uint2))); BG_CU(cudaMalloc(&d_items[1], N * sizeof(uint2)));
  BG_CU(cudaMalloc(&d_ctr, 3 * sizeof(unsigned int)));
  BG_CU(cudaMemcpy(d_boxes, prim_boxes.data(), N * sizeof(Box3), cudaMemcpyHostToDevice));
  const int TB = 256, GB = (n + TB - 1) / TB;
  k_morton<<<GB, TB>>>(d_boxes, n, make_float3(lo[0], lo[1], lo[2]), make_float3(inv[0], inv[1], inv[2]), d_keys, d_vals);
  size_t tmp_bytes = 0;
  BG_CU(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, d_keys, d_keys2, d_vals, d_vals2, n, 0, 63));
  BG_CU(cudaMalloc(&d_tmp, tmp_bytes));
  BG_CU(cub::DeviceRadixSort::SortPairs(d_tmp, tmp_bytes, d_keys, d_keys2, d_vals, d_vals2, n, 0, 63));
  int root = 0;
  if (algorithm == 0) { // LBVH: binary radix tree + bottom-up refit
    k_radix_tree<<<GB, TB>>>(d_keys2, n, d_left, d_right, d_parent, d_count);
    BG_CU(cudaMemset(d_arrived, 0, N * 4));
    k_refit<<<GB, TB>>>(d_boxes, d_vals2, n, d_left, d_right, d_parent, d_arrived, d_nbox);
    BG_CU(cudaGetLastError());
    BG_CU(cudaDeviceSynchronize());
  } else { // PLOC
    for (int k = 0; k < 2; k++) { BG_CU(cudaMalloc(&d_cid[k], N * 4)); BG_CU(cudaMalloc(&d_cbox[k], N * sizeof(Box3))); }
    BG_CU(cudaMalloc(&d_nn, N * 4)); BG_CU(cudaMalloc(&d_valid, N * 4)); BG_CU(cudaMalloc(&d_pos, N * 4));
    size_t scan_bytes = 0;
    BG_CU(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_valid, d_pos, n));
    if (scan_bytes > tmp_bytes) { cudaFree(d_tmp); d_tmp = nullptr; BG_CU(cudaMalloc(&d_tmp, scan_bytes)); tmp_bytes = scan_bytes; }
    BG_CU(cudaMemset(d_ctr, 0, sizeof(unsigned int))); // inner-node ids are handed out from 0
    k_ploc_init<<<GB, TB>>>(d_boxes, d_vals2, n, d_cid[0], d_cbox[0], d_nbox);
    int m = n, rounds = 0;
    while (m > 1) {
      const int gb = (m + TB - 1) / TB;
      k_ploc_nn<<<gb, TB>>>(d_cbox[0], m, d_nn);
      k_ploc_merge<<<gb, TB>>>(d_cid[0], d_cbox[0], m, d_nn, d_left, d_right, d_count, d_nbox, d_ctr, n, d_cid[1], d_cbox[1], d_valid);
      BG_CU(cub::DeviceScan::ExclusiveSum(d_tmp, scan_bytes, d_valid, d_pos, m));
      k_ploc_compact<<<gb, TB>>>(d_cid[1], d_cbox[1], d_valid, d_pos, m, d_cid[0], d_cbox[0]);
      int last_pos = 0, last_valid = 0;
      BG_CU(cudaMemcpy(&last_pos, d_pos + (m - 1), 4, cudaMemcpyDeviceToHost));
      BG_CU(cudaMemcpy(&last_valid, d_valid + (m - 1), 4, cudaMemcpyDeviceToHost));
      const int m_new = last_pos + last_valid;
      if (m_new >= m || ++rounds > 100000) { if (err) *err = "build_cwbvh_device: PLOC made no progress"; cleanup(); return false; }
      m = m_new;
    }
    BG_CU(cudaMemcpy(&root, d_cid[0], 4, cudaMemcpyDeviceToHost));
  }
  // ---- collapse, one launch per level of the wide tree
  unsigned int ctr0[3] = {1u, 0u, 0u}; // node 0 is the root
  BG_CU(cudaMemcpy(d_ctr, ctr0, sizeof(ctr0), cudaMemcpyHostToDevice));
  const uint2 root_item = make_uint2((uint32_t)root, 0u);
  BG_CU(cudaMemcpy(d_items[0], &root_item, sizeof(uint2), cudaMemcpyHostToDevice));
  CollapseArgs A;
  A.n = n; A.max_leaf = max_leaf; A.left = d_left; A.right = d_right; A.count = d_count; A.nbox = d_nbox; A.vals = d_vals2;
  A.nodes = d_nodes; A.order = d_order; A.node_counter = d_ctr; A.prim_counter = d_ctr + 1; A.out_counter = d_ctr + 2;
  int n_in = 1, cur = 0, levels = 0;
  while (n_in > 0) {
    BG_CU(cudaMemset(d_ctr + 2, 0, sizeof(unsigned int)));
    A.items_in = d_items[cur]; A.items_out = d_items[cur ^ 1]; A.n_in = n_in;
    k_collapse<<<(n_in + 127) / 128, 128>>>(A);
    BG_CU(cudaGetLastError());
    unsigned int n_out = 0;
    BG_CU(cudaMemcpy(&n_out, d_ctr + 2, sizeof(unsigned int), cudaMemcpyDeviceToHost));
    n_in = (int)n_out; cur ^= 1;
    if (++levels > 4096) { if (err) *err = "build_cwbvh_device: runaway collapse"; cleanup(); return false; }
  }
  unsigned int ctr[3];
  BG_CU(cudaMemcpy(ctr, d_ctr, sizeof(ctr), cudaMemcpyDeviceToHost));
  if ((int)ctr[1] != n) { if (err) *err = "build_cwbvh_device: primitive count mismatch"; cleanup(); return false; }
  out_nodes.resize(ctr[0]); out_prim_order.resize(N);
  BG_CU(cudaMemcpy(out_nodes.data(), d_nodes, (size_t)ctr[0] * sizeof(CwNode), cudaMemcpyDeviceToHost));
  BG_CU(cudaMemcpy(out_prim_order.data(), d_order, N * 4, cudaMemcpyDeviceToHost));
  if (root_box) BG_CU(cudaMemcpy(root_box, d_nbox + root, sizeof(Box3), cudaMemcpyDeviceToHost));
  cleanup();
  return true;
}

} // namespace spt
