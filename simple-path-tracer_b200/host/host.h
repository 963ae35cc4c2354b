// host.h — the compiled host side above the C ABI of include/spt.h.
//
// The reference is one Rust binary (src/main.rs) whose host code — JSON scene loader, mesh reader, renderer factory,
// image writer — sits around the hot path.  There is no Rust toolchain in this image, so the host is written in C++ and
// mirrors the reference's interfaces name for name:
//
//   load_scene / load_renderer   src/loader/mod.rs:8-32, src/loader/json.rs:19-242, src/core/loader.rs:12-470
//   create_renderer, RendererT::render, OutputConfig     src/renderer/mod.rs:9-38
//   Scene::get_camera            src/core/scene.rs:29-41
//   Film::save_image (8-bit)     src/core/film.rs:94-99
//
// Everything here is host logic; all rendering goes through libspt_cuda.so (spt_scene_create / spt_render).
#pragma once
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../../include/spt.h"

namespace spt_host {

struct HostError : std::runtime_error { // ≙ anyhow::Error of the reference's loaders
  explicit HostError(const std::string& m) : std::runtime_error(m) {}
};

struct Mesh { // TriMesh, src/primitive/triangle.rs:10-27
  std::vector<float> positions, normals, texcoords, tangents, bitangents;
  std::vector<uint32_t> indices;
};
Mesh load_mesh(const std::string& path); // .obj or .npz, tangents computed
void calc_tangents(Mesh& m);

struct Image8 { int width = 0, height = 0; std::vector<uint8_t> rgba; };
Image8 load_png(const std::string& path);                                              // 8-bit, non-interlaced
void save_png(const std::string& path, int width, int height, const uint8_t* rgb);     // RGB8
std::vector<float> load_exr_rgb(const std::string& path, int& width, int& height);     // scanline EXR, none/rle/zips/zip

// The flat, index-linked scene that becomes an SptSceneDesc (owner of every array the descriptor points into).
struct FlatScene {
  std::vector<Mesh> meshes;
  std::vector<SptSphereDesc> spheres;
  std::vector<SptInstanceDesc> instances;
  std::vector<SptSurfaceDesc> surfaces;
  std::vector<SptMaterialDesc> materials;
  std::vector<SptTextureDesc> textures;
  std::vector<std::vector<uint8_t>> texture_pixels; // parallel to textures (empty unless SPT_TEX_IMAGE)
  std::vector<SptMediumDesc> mediums;
  std::vector<SptLightDesc> lights;
  std::vector<SptCameraDesc> cameras;
  bool has_env = false;
  SptEnvDesc env{};
  std::vector<float> env_rgb;
  int32_t env_light_index = -1;
  int32_t light_sampler = SPT_LS_UNIFORM;
  int32_t aggregate = SPT_AGG_BVH;
  std::map<std::string, int32_t> texture_names, material_names, medium_names, surface_names, instance_names, light_names, camera_names;
  std::map<std::string, std::pair<int32_t, int32_t>> primitive_names; // (kind, index); kind -1 = skipped (out of scope)
  std::vector<std::string> warnings;

  int32_t camera_index(const std::string* name) const; // Scene::get_camera
  const SptSceneDesc& desc();                          // pointers stay valid while *this is alive and unchanged

 private:
  std::vector<SptMeshDesc> mesh_descs_;
  SptSceneDesc desc_{};
};

FlatScene load_scene(const std::string& scene_path, const std::string& base_path = "");

struct PixelSampler { int32_t kind = SPT_SAMPLER_RECURRENCE, spp = 256, division_x = 1, division_y = 1; };
struct BoxFilter { float radius = 0.5f; };
struct RendererConfig { std::string type; int32_t max_depth = 8; PixelSampler sampler; BoxFilter filter; };
RendererConfig load_renderer_config(const std::string& path);

struct OutputConfig { // src/renderer/mod.rs:9-14
  int32_t width = 512, height = 512;
  std::string output_filename; // empty = none
  std::string used_camera_name; bool has_camera_name = false;
};

// A scene resident on one GPU (spt_context_create + spt_scene_create).
class DeviceScene {
 public:
  DeviceScene(FlatScene& fs, int device);
  ~DeviceScene();
  DeviceScene(const DeviceScene&) = delete;
  DeviceScene& operator=(const DeviceScene&) = delete;
  SptScene* handle() const { return scene_; }
  FlatScene& flat() const { return fs_; }

 private:
  FlatScene& fs_;
  SptContext* ctx_ = nullptr;
  SptScene* scene_ = nullptr;
};

// One process per GPU: rank `rank` of `world_size`, joined over NCCL (multi_gpu.cpp).
class RankGroup {
 public:
  RankGroup(int rank, int world_size, int device, const std::string& id_file);
  ~RankGroup();
  RankGroup(const RankGroup&) = delete;
  RankGroup& operator=(const RankGroup&) = delete;
  std::vector<float> render_and_gather(DeviceScene& scene, const SptRenderDesc& rd, SptStats* stats, double* reduce_ms);
  const int rank, world_size;

 private:
  struct Impl;
  Impl* impl_;
};

class Renderer { // RendererT, src/renderer/mod.rs:16-19
 public:
  virtual ~Renderer() {}
  // Returns the linear mean image (height*width*3 f32, row 0 = top); writes the 8-bit PNG if config names a file.
  virtual std::vector<float> render(DeviceScene& scene, const OutputConfig& config) = 0;
};

class WavefrontPathTracer : public Renderer { // ≙ PathTracer, src/renderer/pt.rs:24-37
 public:
  WavefrontPathTracer(int32_t max_depth, PixelSampler sampler, BoxFilter filter) : max_depth_(max_depth), sampler_(sampler), filter_(filter) {}
  std::vector<float> render(DeviceScene& scene, const OutputConfig& config) override;
  uint64_t seed = 1; int32_t pool_size = 0; uint32_t flags = 0; int32_t rank = 0, world_size = 1;
  RankGroup* group = nullptr; // set = render this rank's tiles and gather the frame on rank 0 over NCCL
  int32_t spp_first = 0, spp_count = 0; bool accumulate = false; // progressive passes (SptRenderDesc::spp_first / spp_count)
  SptStats last_stats{}; double last_reduce_ms = 0.0;

 private:
  int32_t max_depth_; PixelSampler sampler_; BoxFilter filter_;
};

// renderer::create_renderer, src/renderer/mod.rs:26-38 ("pt_wavefront" is the arm this library adds; "pt" is refused:
// there is no CPU integrator here).
Renderer* create_renderer(const std::string& ty, int32_t max_depth, PixelSampler sampler, BoxFilter filter);

void dump_scene(FlatScene& fs, const std::string& path); // canonical binary dump of the flattened scene (tests)

} // namespace spt_host
