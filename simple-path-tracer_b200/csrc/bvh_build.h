// bvh_build.h — host-side compressed wide BVH builder (see bvh_build.cpp).
#pragma once
#include <stdint.h>
#include <string>
#include <vector>

namespace spt {

struct Box3 { float lo[3], hi[3]; };

struct CwNode { // 80 bytes, read on the device as 5 x float4 (layout documented in scene.h)
  float p[3];
  int8_t e[3];
  uint8_t imask;
  uint32_t child_base, prim_base;
  uint8_t meta[8];
  uint8_t qlo[3][8];
  uint8_t qhi[3][8];
};
static_assert(sizeof(CwNode) == 80, "CwNode must be 80 bytes");

// Builds a CWBVH over `prim_boxes`.  Node indices stored in the nodes are offset by node_base, primitive
// indices by prim_base (so several structures can share one device array).  out_prim_order[i] = index into
// prim_boxes of the i-th primitive in leaf order.  max_leaf in 1..3 (primitives per leaf slot).
void build_cwbvh(const std::vector<Box3>& prim_boxes, int max_leaf, uint32_t node_base, uint32_t prim_base,
                 std::vector<CwNode>& out_nodes, std::vector<int>& out_prim_order, Box3* root_box);

// The same structure built on the current CUDA device (bvh_build_gpu.cu: Morton sort -> radix tree -> refit -> wide
// collapse); node and primitive bases are 0.  Needs >= 2 primitives.  Returns false with *err set on a CUDA failure.
bool build_cwbvh_device(const std::vector<Box3>& prim_boxes, int max_leaf, std::vector<CwNode>& out_nodes, std::vector<int>& out_prim_order,
                        Box3* root_box, std::string* err, int algorithm = 1 /* 0 = LBVH (radix tree), 1 = PLOC */);

} // namespace spt
