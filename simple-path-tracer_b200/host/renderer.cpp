// renderer.cpp — the renderer interface of the reference (src/renderer/mod.rs:9-38) on top of the C ABI: DeviceScene
// owns the context + uploaded scene, WavefrontPathTracer::render ≙ PathTracer::render (src/renderer/pt.rs:228-296) but
// the whole image is produced by one spt_render call.
#include <cstdio>
#include <cstring>
#include <fstream>

#include "host.h"

namespace spt_host {

namespace {
void check(int32_t rc, const char* what) {
  if (rc != SPT_OK) throw HostError(std::string(what) + ": " + spt_last_error());
}
} // namespace

DeviceScene::DeviceScene(FlatScene& fs, int device) : fs_(fs) {
  check(spt_context_create(device, &ctx_), "spt_context_create");
  const int32_t rc = spt_scene_create(ctx_, &fs.desc(), &scene_);
  if (rc != SPT_OK) { const std::string msg = spt_last_error(); spt_context_destroy(ctx_); throw HostError("spt_scene_create: " + msg); }
}

DeviceScene::~DeviceScene() {
  if (scene_) spt_scene_destroy(scene_);
  if (ctx_) spt_context_destroy(ctx_);
}

std::vector<float> WavefrontPathTracer::render(DeviceScene& scene, const OutputConfig& config) {
  SptRenderDesc rd{};
  rd.camera_index = scene.flat().camera_index(config.has_camera_name ? &config.used_camera_name : nullptr);
  rd.width = config.width; rd.height = config.height;
  rd.max_depth = max_depth_;
  rd.sampler_kind = sampler_.kind; rd.spp = sampler_.spp; rd.div_x = sampler_.division_x; rd.div_y = sampler_.division_y;
  rd.filter_radius = filter_.radius;
  rd.seed = seed;
  rd.tile_w = rd.tile_h = 0;
  rd.rank = rank; rd.world_size = world_size;
  rd.pool_size = pool_size; rd.flags = flags | (accumulate ? SPT_RENDER_ACCUMULATE : 0u);
  rd.spp_first = spp_first; rd.spp_count = spp_count;
  const size_t n = (size_t)config.width * config.height * 3;
  std::vector<float> mean;
  std::vector<uint8_t> rgb8(n);
  if (group) {
    mean = group->render_and_gather(scene, rd, &last_stats, &last_reduce_ms);
    if (group->rank != 0) return mean; // empty: the frame lives on rank 0
    for (size_t i = 0; i < n; i++) { // Film::color_to_rgb, film.rs:94-99 (the library does this on the device for spt_render)
      const float v = mean[i] * 255.0f;
      rgb8[i] = (uint8_t)(v > 0.0f ? (v > 255.0f ? 255.0f : v) : 0.0f); // NaN -> 0, like the device conversion
    }
  } else {
    mean.resize(n);
    check(spt_render(scene.handle(), &rd, mean.data(), rgb8.data(), &last_stats), "spt_render");
  }
  if (!config.output_filename.empty()) save_png(config.output_filename, config.width, config.height, rgb8.data());
  return mean;
}

Renderer* create_renderer(const std::string& ty, int32_t max_depth, PixelSampler sampler, BoxFilter filter) {
  if (ty == "pt_wavefront") return new WavefrontPathTracer(max_depth, sampler, filter);
  if (ty == "pt") throw HostError("renderer - type 'pt' is the reference's CPU integrator; this library only provides 'pt_wavefront'");
  throw HostError("renderer - unknown type '" + ty + "'");
}

// Canonical dump of the flattened scene: records of [u32 name_len][name][u64 nbytes][bytes].  Struct arrays are written
// as their raw C layout (all fields are 4 bytes wide); textures without their pixel pointer.
void dump_scene(FlatScene& fs, const std::string& path) {
  std::ofstream fh(path, std::ios::binary);
  if (!fh) throw HostError("can't write '" + path + "'");
  auto rec = [&](const std::string& name, const void* data, size_t nbytes) {
    const uint32_t nl = (uint32_t)name.size(); const uint64_t nb = nbytes;
    fh.write((const char*)&nl, 4); fh.write(name.data(), nl); fh.write((const char*)&nb, 8);
    if (nbytes) fh.write((const char*)data, (std::streamsize)nbytes);
  };
  const SptSceneDesc& d = fs.desc();
  for (size_t i = 0; i < fs.meshes.size(); i++) {
    const Mesh& m = fs.meshes[i];
    const std::string k = "mesh" + std::to_string(i) + ".";
    rec(k + "positions", m.positions.data(), m.positions.size() * 4);
    rec(k + "normals", m.normals.data(), m.normals.size() * 4);
    rec(k + "texcoords", m.texcoords.data(), m.texcoords.size() * 4);
    rec(k + "tangents", m.tangents.data(), m.tangents.size() * 4);
    rec(k + "bitangents", m.bitangents.data(), m.bitangents.size() * 4);
    rec(k + "indices", m.indices.data(), m.indices.size() * 4);
  }
  rec("spheres", d.spheres, sizeof(SptSphereDesc) * d.n_spheres);
  rec("instances", d.instances, sizeof(SptInstanceDesc) * d.n_instances);
  rec("surfaces", d.surfaces, sizeof(SptSurfaceDesc) * d.n_surfaces);
  rec("materials", d.materials, sizeof(SptMaterialDesc) * d.n_materials);
  std::string tex;
  for (int i = 0; i < d.n_textures; i++) tex.append((const char*)&d.textures[i], offsetof(SptTextureDesc, rgba));
  rec("textures", tex.data(), tex.size());
  for (size_t i = 0; i < fs.texture_pixels.size(); i++)
    if (!fs.texture_pixels[i].empty()) rec("texture" + std::to_string(i) + ".rgba", fs.texture_pixels[i].data(), fs.texture_pixels[i].size());
  rec("mediums", d.mediums, sizeof(SptMediumDesc) * d.n_mediums);
  rec("lights", d.lights, sizeof(SptLightDesc) * d.n_lights);
  rec("cameras", d.cameras, sizeof(SptCameraDesc) * d.n_cameras);
  if (d.env) {
    const int32_t wh[2] = {d.env->width, d.env->height};
    rec("env.size", wh, 8); rec("env.scale", d.env->scale, 12); rec("env.rgb", d.env->rgb, (size_t)d.env->width * d.env->height * 12);
  }
  const int32_t misc[3] = {d.env_light_index, d.light_sampler, d.aggregate};
  rec("misc", misc, 12);
  std::string w;
  for (const auto& s : fs.warnings) { w += s; w += '\n'; }
  rec("warnings", w.data(), w.size());
}

} // namespace spt_host
