// main.cpp — command line of the reference binary (src/main.rs:26-68) on the CUDA core:
//
//   spt_host --scene S.json --renderer R.json [--width 512] [--height 512] --output out.png [--camera NAME]
//
// plus host-side options that have no counterpart in the reference:
//   --renderer-type T   renderer arm to create (default pt_wavefront; the file's "pt" names the reference's CPU integrator)
//   --base-path P       resolve relative obj_file / image_file / external JSON against P instead of the scene file
//   --device N  --seed N  --pool N  --spp N (overrides the renderer file)
//   --gpus N            one process per GPU: forks N ranks (GPU r = rank r) that render interleaved tiles and gather the
//                       frame on rank 0 with spt_render_gather (owned tiles only, over NCCL inside the library); rank 0 writes the image
//   --rank R --world-size W [--nccl-id FILE]   what --gpus passes to each rank; without --nccl-id the rank only renders
//                       the tiles it owns (no gather)
//   --frames N          render the frame N times (same seed) and report the last one: steady-state timing without the
//                       first call's allocations and NCCL connection set-up
//   --debug-normal      first-hit normal AOV = the reference built with `--features debug_normal` (pt.rs:113-118)
//   --passes N          progressive rendering: the frame's spp in N passes over consecutive sample ranges (SPT_RENDER_ACCUMULATE);
//                       the image file is rewritten after every pass
//   --checkpoint FILE   after every pass, write the film (int64 fixed-point sums) + how many samples it holds
//   --resume FILE       load such a checkpoint and render only the samples that are still missing
//   --raw FILE          also write the linear f32 mean image (height*width*3, row 0 = top)
//   --dump-desc FILE    write the flattened scene (no GPU needed) and exit
#include <sys/wait.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>

#include "host.h"

using namespace spt_host;

namespace {
std::string ext_of(const std::string& path) { // loader/mod.rs:8-32 dispatches on the file extension
  const size_t slash = path.find_last_of('/'), dot = path.find_last_of('.');
  if (dot == std::string::npos || (slash != std::string::npos && dot < slash)) throw HostError("There is no file extension");
  return path.substr(dot + 1);
}
} // namespace

int main(int argc, char** argv) {
  std::string scene_path, renderer_path, output, camera, base_path, dump, raw, renderer_type = "pt_wavefront";
  bool has_camera = false;
  int width = 512, height = 512, device = -1, pool = 0, spp = 0, rank = 0, world = 1, gpus = 0, frames = 1, passes = 1;
  std::string checkpoint, resume;
  std::string nccl_id;
  bool debug_normal = false;
  uint64_t seed = 1;
  for (int i = 1; i < argc; i++) {
    const std::string a = argv[i];
    auto val = [&]() -> std::string { if (i + 1 >= argc) { fprintf(stderr, "error: %s needs a value\n", a.c_str()); exit(2); } return argv[++i]; };
    if (a == "--scene" || a == "-s") scene_path = val();
    else if (a == "--renderer" || a == "-r") renderer_path = val();
    else if (a == "--width" || a == "-w") width = atoi(val().c_str());
    else if (a == "--height" || a == "-h") height = atoi(val().c_str());
    else if (a == "--output" || a == "-o") output = val();
    else if (a == "--camera" || a == "-c") { camera = val(); has_camera = true; }
    else if (a == "--renderer-type") renderer_type = val();
    else if (a == "--base-path") base_path = val();
    else if (a == "--device") device = atoi(val().c_str());
    else if (a == "--seed") seed = strtoull(val().c_str(), nullptr, 10);
    else if (a == "--pool") pool = atoi(val().c_str());
    else if (a == "--spp") spp = atoi(val().c_str());
    else if (a == "--rank") rank = atoi(val().c_str());
    else if (a == "--world-size") world = atoi(val().c_str());
    else if (a == "--gpus") gpus = atoi(val().c_str());
    else if (a == "--nccl-id") nccl_id = val();
    else if (a == "--frames") frames = std::max(1, atoi(val().c_str()));
    else if (a == "--passes") passes = std::max(1, atoi(val().c_str()));
    else if (a == "--checkpoint") checkpoint = val();
    else if (a == "--resume") resume = val();
    else if (a == "--debug-normal") debug_normal = true; // the reference's cargo feature `debug_normal` (pt.rs:113-118)
    else if (a == "--raw") raw = val();
    else if (a == "--dump-desc") dump = val();
    else { fprintf(stderr, "error: unknown argument '%s'\n", a.c_str()); return 2; }
  }
  if (gpus > 1) { // launcher: one child process per GPU, same command line plus the rank arguments
    const std::string id_file = "/tmp/spt_nccl_id." + std::to_string((long)getpid());
    unlink(id_file.c_str());
    std::vector<pid_t> kids;
    for (int r = 0; r < gpus; r++) {
      const pid_t pid = fork();
      if (pid < 0) { perror("fork"); return 1; }
      if (pid == 0) {
        std::vector<std::string> args;
        for (int i = 0; i < argc; i++) { if (std::string(argv[i]) == "--gpus") { i++; continue; } args.push_back(argv[i]); }
        for (const std::string& x : {std::string("--rank"), std::to_string(r), std::string("--world-size"), std::to_string(gpus), std::string("--nccl-id"), id_file}) args.push_back(x);
        std::vector<char*> av;
        for (auto& x : args) av.push_back(&x[0]);
        av.push_back(nullptr);
        execv("/proc/self/exe", av.data());
        perror("execv"); _exit(127);
      }
      kids.push_back(pid);
    }
    int rc = 0;
    for (pid_t k : kids) { int st = 0; waitpid(k, &st, 0); if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = 1; }
    unlink(id_file.c_str());
    return rc;
  }
  if (device < 0) device = world > 1 && !nccl_id.empty() ? rank : 0;
  try {
    if (scene_path.empty()) throw HostError("--scene is required");
    const std::string sext = ext_of(scene_path);
    if (sext == "gltf") throw HostError("json - 'gltf' import is out of scope of the grafted path");
    if (sext != "json") throw HostError("File extension '" + sext + "' is not recognized");
    FlatScene fs = load_scene(scene_path, base_path);
    for (const auto& w : fs.warnings) fprintf(stderr, "warning: %s\n", w.c_str());
    if (!dump.empty()) { dump_scene(fs, dump); return 0; }
    if (renderer_path.empty() || output.empty()) throw HostError("--renderer and --output are required");
    const std::string rext = ext_of(renderer_path);
    if (rext != "json") throw HostError("File extension '" + rext + "' is not recognized");
    RendererConfig rc = load_renderer_config(renderer_path);
    if (spp > 0) rc.sampler.spp = spp;
    std::unique_ptr<Renderer> renderer(create_renderer(renderer_type.empty() ? rc.type : renderer_type, rc.max_depth, rc.sampler, rc.filter));
    WavefrontPathTracer* wf = dynamic_cast<WavefrontPathTracer*>(renderer.get());
    if (wf) { wf->seed = seed; wf->pool_size = pool; wf->rank = rank; wf->world_size = world; if (debug_normal) wf->flags |= SPT_RENDER_DEBUG_NORMAL; }
    DeviceScene scene(fs, device);
    std::unique_ptr<RankGroup> group;
    if (world > 1 && !nccl_id.empty()) { group.reset(new RankGroup(rank, world, device, nccl_id)); if (wf) wf->group = group.get(); }
    OutputConfig oc; oc.width = width; oc.height = height; oc.output_filename = output; oc.used_camera_name = camera; oc.has_camera_name = has_camera;
    std::vector<float> mean;
    double wall = 0.0;
    if (wf && (passes > 1 || !resume.empty() || !checkpoint.empty())) {
      if (group) throw HostError("--passes / --checkpoint / --resume are single-process options");
      const int total = rc.sampler.kind == SPT_SAMPLER_JITTERED ? rc.sampler.division_x * rc.sampler.division_y : rc.sampler.spp;
      int done = 0;
      if (!resume.empty()) { // header: magic, width, height, spp_done; then width*height*3 int64
        std::ifstream fh(resume, std::ios::binary);
        int32_t hdr[4] = {0, 0, 0, 0};
        if (!fh || !fh.read((char*)hdr, sizeof(hdr)) || hdr[0] != 0x46545053) throw HostError("'" + resume + "' is not a film checkpoint");
        if (hdr[1] != width || hdr[2] != height || hdr[3] <= 0 || hdr[3] > total) throw HostError("checkpoint does not match --width/--height/spp");
        std::vector<int64_t> film((size_t)width * height * 3);
        if (!fh.read((char*)film.data(), (std::streamsize)(film.size() * 8))) throw HostError("truncated checkpoint");
        if (spt_film_load(scene.handle(), film.data(), width, height, hdr[3]) != SPT_OK) throw HostError(std::string("spt_film_load: ") + spt_last_error());
        done = hdr[3];
      }
      const int todo = total - done, per = (todo + passes - 1) / passes;
      const auto t0 = std::chrono::steady_clock::now();
      while (done < total) {
        wf->spp_first = done; wf->spp_count = std::min(per, total - done); wf->accumulate = done > 0;
        mean = renderer->render(scene, oc);
        done += wf->spp_count;
        if (!checkpoint.empty()) {
          std::vector<int64_t> film((size_t)width * height * 3);
          if (spt_film_save(scene.handle(), film.data(), (int64_t)film.size()) != SPT_OK) throw HostError(std::string("spt_film_save: ") + spt_last_error());
          const int32_t hdr[4] = {0x46545053, width, height, done};
          std::ofstream fh(checkpoint + ".tmp", std::ios::binary);
          fh.write((const char*)hdr, sizeof(hdr)); fh.write((const char*)film.data(), (std::streamsize)(film.size() * 8));
          fh.close();
          if (!fh || rename((checkpoint + ".tmp").c_str(), checkpoint.c_str()) != 0) throw HostError("can't write '" + checkpoint + "'");
        }
        fprintf(stderr, "pass done: %d / %d spp\n", done, total);
      }
      wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    } else
    for (int f = 0; f < frames; f++) {
      const auto t0 = std::chrono::steady_clock::now();
      mean = renderer->render(scene, oc);
      wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    }
    if (group && rank != 0) return 0; // the frame and the report live on rank 0
    if (!raw.empty()) { std::ofstream fh(raw, std::ios::binary); fh.write((const char*)mean.data(), (std::streamsize)(mean.size() * 4)); }
    if (wf) {
      const SptStats& st = wf->last_stats;
      printf("{\"paths\": %llu, \"ext_rays\": %llu, \"shadow_rays\": %llu, \"probe_rays\": %llu, \"iterations\": %llu, \"kernel_launches\": %llu, "
             "\"seconds_device\": %.6f, \"seconds_wall\": %.6f, \"mpaths_per_s\": %.3f, \"world_size\": %d, \"reduce_ms\": %.3f, \"frame_mpaths_per_s_wall\": %.3f}\n",
             (unsigned long long)st.paths, (unsigned long long)st.ext_rays, (unsigned long long)st.shadow_rays, (unsigned long long)st.probe_rays,
             (unsigned long long)st.iterations, (unsigned long long)st.kernel_launches, st.seconds_total, wall,
             st.seconds_total > 0 ? st.paths / st.seconds_total / 1e6 : 0.0, group ? world : 1, wf->last_reduce_ms,
             (double)width * height * rc.sampler.spp / wall / 1e6);
    }
  } catch (const std::exception& e) {
    fprintf(stderr, "Error: %s\n", e.what());
    return 1;
  }
  return 0;
}
