// multi_gpu.cpp — one process per GPU for the compiled host.  The ranks rendezvous through a file holding the communicator id
// (spt_comm_unique_id on rank 0), join with spt_comm_create, and every frame is one spt_render_gather: each rank renders the
// interleaved tiles it owns and only those tiles travel — as the integer film, over NVLink by ncclSend / ncclRecv inside the library —
// to rank 0, which reconstructs the frame.  The host binds ONE library: NCCL never appears here.  There is no per-bounce communication
// (SURVEY.md §8e).  The reference's own parallelism is row bands over CPU threads (renderer/pt.rs:243-261).
#include <cuda_runtime.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <thread>

#include "host.h"

namespace spt_host {

namespace {
void cu(cudaError_t e, const char* what) { if (e != cudaSuccess) throw HostError(std::string(what) + ": " + cudaGetErrorString(e)); }
void ok(int32_t st, const char* what) { if (st != SPT_OK) throw HostError(std::string(what) + ": " + spt_last_error()); }
} // namespace

struct RankGroup::Impl { SptContext* ctx = nullptr; SptComm* comm = nullptr; };

RankGroup::RankGroup(int rank_, int world_, int device, const std::string& id_file) : rank(rank_), world_size(world_), impl_(new Impl) {
  uint8_t id[SPT_COMM_ID_BYTES];
  if (rank == 0) {
    ok(spt_comm_unique_id(id), "spt_comm_unique_id");
    const std::string tmp = id_file + ".tmp";
    { std::ofstream fh(tmp, std::ios::binary); fh.write((const char*)id, sizeof(id)); if (!fh) throw HostError("can't write '" + tmp + "'"); }
    if (rename(tmp.c_str(), id_file.c_str()) != 0) throw HostError("can't publish the communicator id file '" + id_file + "'");
  } else {
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      std::ifstream fh(id_file, std::ios::binary);
      if (fh && fh.read((char*)id, sizeof(id))) break;
      if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120.0) throw HostError("timed out waiting for the communicator id file '" + id_file + "'");
      std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
  }
  ok(spt_context_create(device, &impl_->ctx), "spt_context_create");
  ok(spt_comm_create(impl_->ctx, id, rank, world_size, &impl_->comm), "spt_comm_create");
}

RankGroup::~RankGroup() {
  spt_comm_destroy(impl_->comm);
  spt_context_destroy(impl_->ctx);
  delete impl_;
}

// Render this rank's tiles and gather the frame on rank 0.  Returns the full mean image on rank 0, an empty vector elsewhere.
std::vector<float> RankGroup::render_and_gather(DeviceScene& scene, const SptRenderDesc& rd, SptStats* stats, double* reduce_ms) {
  const size_t n = (size_t)rd.width * rd.height * 3;
  float* d_img = nullptr;
  if (rank == 0) cu(cudaMalloc(&d_img, n * sizeof(float)), "cudaMalloc(framebuffer)");
  std::vector<float> out;
  SptStats st; memset(&st, 0, sizeof(st));
  try {
    ok(spt_render_gather(scene.handle(), impl_->comm, &rd, 0, d_img, &st), "spt_render_gather");
    if (stats) *stats = st;
    if (reduce_ms) *reduce_ms = st.seconds_stage[5] * 1e3;
    if (rank == 0) { out.resize(n); cu(cudaMemcpy(out.data(), d_img, n * sizeof(float), cudaMemcpyDeviceToHost), "cudaMemcpy(framebuffer)"); }
  } catch (...) { cudaFree(d_img); throw; }
  cudaFree(d_img);
  return out;
}

} // namespace spt_host
