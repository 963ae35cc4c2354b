// flatten.h — host-side flattening of an SptSceneDesc into the device arrays of scene.h.
#pragma once
#ifndef SPT_DEVICE_BUILD_MIN_TRIS
#define SPT_DEVICE_BUILD_MIN_TRIS (1 << 18) // meshes at least this big get their BLAS built on the GPU
#endif
#include <string>
#include <vector>
#include "../../include/spt.h"
#include "scene.h"
#include "bvh_build.h"

namespace spt {

struct HostScene {
  std::vector<CwNode> nodes;        // TLAS first, then BLASes
  std::vector<float4> tris;         // 3 per triangle, BLAS leaf order
  std::vector<float4> shade_tris;   // 9 per triangle, same order
  std::vector<uint8_t> inst_mat;    // per instance: material tag
  uint32_t mat_mask = 0;
  bool has_medium = false;
  std::vector<int32_t> tlas_inst;
  std::vector<DInstance> inst;
  std::vector<DInstanceShade> inst_shade;
  std::vector<float4> spheres;
  std::vector<DMesh> meshes;
  std::vector<float4> verts;
  std::vector<uint32_t> indices;
  std::vector<DSurface> surfaces;
  std::vector<DMaterial> materials;
  std::vector<DTexture> textures;
  std::vector<DMip> mips;
  std::vector<uchar4> texels;
  std::vector<DMedium> mediums;
  std::vector<DLight> lights;
  std::vector<float> light_prob;
  std::vector<float2> light_alias;
  std::vector<int32_t> inst_light;
  std::vector<float4> env_tex;
  std::vector<float2> env_alias;
  std::vector<float> ss_cdf;
  std::vector<DCamera> cameras;
  float env_scale[3] = {1, 1, 1};
  int env_w = 0, env_h = 0;
  int light_sampler = 0, env_light_index = -1;
  bool has_image_tex = false;
  bool lean = false;                // every texture is a plain scalar and no surface has a normal / emissive map: the SPT_LEAN shade kernels apply
  int64_t tlas_nodes = 0, blas_nodes = 0;
  double seconds_bvh = 0.0, seconds_create = 0.0; // build statistics (spt_scene_info)
  int64_t blas_on_device = 0;
};

// Returns an empty string on success, else the error message.
std::string flatten_scene(const SptSceneDesc& d, HostScene& out);

} // namespace spt
