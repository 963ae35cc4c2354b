// multi_gpu.cpp — one process per GPU for the compiled host: the ranks rendezvous through a file holding the NCCL unique
// id, each renders the interleaved tiles it owns into a device buffer (spt_render_device) and the framebuffer is summed
// onto rank 0 with ONE ncclReduce over NVLink (disjoint tiles, zeros elsewhere: the sum is exact).  There is no per-bounce
// communication.  The reference's own parallelism is row bands over CPU threads (renderer/pt.rs:243-261).
#include <cuda_runtime.h>
#include <nccl.h>
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
void nc(ncclResult_t r, const char* what) { if (r != ncclSuccess) throw HostError(std::string(what) + ": " + ncclGetErrorString(r)); }
} // namespace

struct RankGroup::Impl { ncclComm_t comm = nullptr; cudaStream_t stream = nullptr; };

RankGroup::RankGroup(int rank_, int world_, int device, const std::string& id_file) : rank(rank_), world_size(world_), impl_(new Impl) {
  cu(cudaSetDevice(device), "cudaSetDevice");
  ncclUniqueId id;
  if (rank == 0) {
    nc(ncclGetUniqueId(&id), "ncclGetUniqueId");
    const std::string tmp = id_file + ".tmp";
    { std::ofstream fh(tmp, std::ios::binary); fh.write((const char*)&id, sizeof(id)); if (!fh) throw HostError("can't write '" + tmp + "'"); }
    if (rename(tmp.c_str(), id_file.c_str()) != 0) throw HostError("can't publish the NCCL id file '" + id_file + "'");
  } else {
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
      std::ifstream fh(id_file, std::ios::binary);
      if (fh && fh.read((char*)&id, sizeof(id))) break;
      if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120.0) throw HostError("timed out waiting for the NCCL id file '" + id_file + "'");
      std::this_thread::sleep_for(std::chrono::milliseconds(20));
    }
  }
  nc(ncclCommInitRank(&impl_->comm, world_size, id, rank), "ncclCommInitRank");
  cu(cudaStreamCreateWithFlags(&impl_->stream, cudaStreamNonBlocking), "cudaStreamCreate");
}

RankGroup::~RankGroup() {
  if (impl_->comm) ncclCommDestroy(impl_->comm);
  if (impl_->stream) cudaStreamDestroy(impl_->stream);
  delete impl_;
}

// Render this rank's tiles and gather the frame on rank 0.  Returns the full mean image on rank 0, an empty vector elsewhere.
std::vector<float> RankGroup::render_and_gather(DeviceScene& scene, const SptRenderDesc& rd_in, SptStats* stats, double* reduce_ms) {
  SptRenderDesc rd = rd_in;
  rd.rank = rank; rd.world_size = world_size;
  const size_t n = (size_t)rd.width * rd.height * 3;
  float* d_img = nullptr;
  cu(cudaMalloc(&d_img, n * sizeof(float)), "cudaMalloc(framebuffer)");
  std::vector<float> out;
  try {
    if (spt_render_device(scene.handle(), &rd, d_img, stats) != SPT_OK) throw HostError(std::string("spt_render_device: ") + spt_last_error());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, impl_->stream);
    nc(ncclReduce(d_img, d_img, n, ncclFloat, ncclSum, 0, impl_->comm, impl_->stream), "ncclReduce");
    cudaEventRecord(e1, impl_->stream);
    cu(cudaStreamSynchronize(impl_->stream), "ncclReduce (sync)");
    float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
    if (reduce_ms) *reduce_ms = ms;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rank == 0) { out.resize(n); cu(cudaMemcpy(out.data(), d_img, n * sizeof(float), cudaMemcpyDeviceToHost), "cudaMemcpy(framebuffer)"); }
  } catch (...) { cudaFree(d_img); throw; }
  cudaFree(d_img);
  return out;
}

} // namespace spt_host
