// scene.h — device-resident scene layout (what SptSceneDesc is flattened into) shared by the host flattener
// and the kernels.  All arrays are plain global memory; hot traversal data is read with 16-byte loads.
//
//  nodes   : compressed 8-wide BVH nodes, 80 B = 5 x float4 each, TLAS first (root = node 0), then one BLAS per
//            mesh.  Layout per node (after Ylitie/Karras/Laine 2017):
//              n0 = (p.x, p.y, p.z, [ex ey ez imask])         p = quantisation origin, e = per-axis exponent
//              n1 = (child_base, prim_base, meta[0..3], meta[4..7])
//              n2 = (qlo_x[0..7], qlo_y[0..7])  n3 = (qlo_z[0..7], qhi_x[0..7])  n4 = (qhi_y[0..7], qhi_z[0..7])
//  tris    : 48 B = 3 x float4 per triangle in BLAS leaf order: (p0, tri_id) (e1 = p1-p0, -) (e2 = p2-p0, -)  — the reference's
//            Triangle (triangle.rs:23-27) reduced to what intersect_ray (:124-147) reads.
//  shade_tris: 144 B = 9 x float4 per triangle in the SAME leaf order: the vertex attributes Triangle::intersect
//            interpolates (triangle.rs:188-213), pre-gathered so shading a hit is one indexed fetch instead of the
//            instance -> mesh -> indices -> vertices chain:
//              (n0, uv0.x) (n1, uv0.y) (n2, uv1.x) (tan0, uv1.y) (tan1, uv2.x) (tan2, uv2.y)
//              (bitan0, prim_pdf) (bitan1, -) (bitan2, -)      prim_pdf = 1/(max(area,1e-3) * n_tris)  (Q6)
//  tlas_inst: TLAS leaf order -> instance id.
//  inst    : 64 B per instance: rows of trans_inv (Affine3A, instance.rs:13) + kind/blas_root/mesh/surface.
//            kind = 0 sphere (root = sphere index) | 1 mesh (root = BLAS root node) |
//                   2 + (n << 8) small mesh of n <= SPT_SMALL_MESH triangles (root = first triangle): the leaf
//                   triangles are tested right after the instance transform, without a BLAS node step.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

#define SPT_SMALL_MESH 4
struct DInstance {      // hot (traversal)
  float4 inv0, inv1, inv2; // row i of trans_inv.matrix3 in xyz, trans_inv.translation[i] in w
  int32_t kind;            // SPT_PRIM_*
  int32_t root;            // mesh: BLAS root node index; sphere: index into spheres
  int32_t mesh;            // mesh id or -1
  int32_t surface;
};
struct DInstanceShade { // cold (once per shaded hit): forward affine and inverse-transpose (transform.rs:7-24)
  float4 fwd0, fwd1, fwd2; // row i of trans.matrix3 in xyz, translation[i] in w
  float4 it0, it1, it2;    // row i of trans_it
};
struct DMesh {
  int32_t vert_base; // first vertex in `verts` (4 float4 per vertex)
  int32_t idx_base;  // first triangle in `indices` (uint3 per triangle, file order, mesh-local vertex ids)
  int32_t n_tris;
  int32_t blas_root;
};
struct DSurface { int32_t material, normal_map, emissive_map, double_sided, inside_medium; float emissive[3]; };
struct DMaterial { int32_t tag, tex[4], rough_chan, metal_chan; float ior; };
struct DTexture {
  int32_t tag, child0, child1, mode, wrap;
  float value[3], tiling[3], offset[3];
  int32_t mip_base, n_mips; // into `mips`
};
struct DMip { int32_t w, h; uint32_t texel_base; int32_t pad; }; // texels: uchar4 RGBA8
struct DMedium { float sigma_t[3], sigma_s[3], g, pad; };
struct DLight {
  int32_t tag, instance;
  float position[3], direction[3], strength[3];
  float cos_inner, cos_outer;
};
struct DCamera { float eye[3], forward[3], right[3], up[3], half_cot_half_fov; };

struct DScene { // passed by value to kernels (fits the 4 KB parameter space comfortably)
  const float4* nodes;
  const float4* tris;
  const float4* shade_tris;
  const int32_t* tlas_inst;
  const DInstance* inst;
  const DInstanceShade* inst_shade;
  const float4* spheres; // center.xyz, radius
  const DMesh* meshes;
  const float4* verts;   // per vertex: (pos, uv.x) (normal, uv.y) (tangent, 0) (bitangent, 0)
  const uint32_t* indices;
  const DSurface* surfaces;
  const DMaterial* materials;
  const DTexture* textures;
  const DMip* mips;
  const uchar4* texels;
  const DMedium* mediums;
  const DLight* lights;
  const float* light_prob;   // power_is: alias props per light (power_is.rs:24-34)
  const float2* light_alias; // power_is: (u, k as int bits) per light (alias_table.rs)
  const int32_t* inst_light; // instance id -> light index or -1 (power_is.rs:15)
  const uint8_t* inst_mat;   // instance id -> material tag of its surface (shade-queue class)
  const float4* env_tex;     // (r,g,b, alias prop) per texel
  const float2* env_alias;   // (u, k as int bits) per texel
  const float* ss_cdf;       // 512 x (x, y) pairs of SS_CDF_TABLE (substrate.rs:187-196)
  DCamera camera;            // the camera of the current render
  float env_scale[3];
  int32_t env_w, env_h;
  int32_t n_lights, light_sampler, env_light_index;
  int32_t n_instances, tlas_empty;
  int32_t has_image_tex;
  uint32_t mat_mask;         // bit t set: some instance uses a material with tag t
  int32_t has_medium;        // some surface has inside_medium
  float num_light_inv;
};
