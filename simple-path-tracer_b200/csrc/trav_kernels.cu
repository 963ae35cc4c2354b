// trav_kernels.cu — translation unit of the traversal kernels (extend / connect / ray batches).
#include "launch.h"
#include "trav_kernels.cuh"

namespace spt {
cudaError_t occupancy_traversal(int* e, int* c, int* b) {
  cudaError_t r = cudaOccupancyMaxActiveBlocksPerMultiprocessor(e, k_extend<false>, SPT_BLOCK, 0);
  if (r != cudaSuccess) return r;
  r = cudaOccupancyMaxActiveBlocksPerMultiprocessor(c, k_connect<false>, SPT_BLOCK, 0);
  if (r != cudaSuccess) return r;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(b, k_batch<false, false>, SPT_BLOCK, 0);
}
void launch_extend(const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st) {
  if (count) k_extend<true><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_extend<false><<<grid, SPT_BLOCK, 0, st>>>(sc, P);
}
void launch_connect(const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st) {
  if (count) k_connect<true><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_connect<false><<<grid, SPT_BLOCK, 0, st>>>(sc, P);
}
void launch_batch(const DScene& sc, const float4* rays, uint32_t n, HitIO* hits, uint8_t* occ, uint32_t* next, DevCounters* ctr,
                  bool any, bool count, int grid, cudaStream_t st) {
  if (!any) { if (count) k_batch<false, true><<<grid, SPT_BLOCK, 0, st>>>(sc, rays, n, hits, occ, next, ctr); else k_batch<false, false><<<grid, SPT_BLOCK, 0, st>>>(sc, rays, n, hits, occ, next, nullptr); }
  else { if (count) k_batch<true, true><<<grid, SPT_BLOCK, 0, st>>>(sc, rays, n, hits, occ, next, ctr); else k_batch<true, false><<<grid, SPT_BLOCK, 0, st>>>(sc, rays, n, hits, occ, next, nullptr); }
}
} // namespace spt

// ---- spt_trace_batch_noalloc: the ray batch of seam B2 without a single allocation ---------------------------------------------------
// Added (no counterpart in the reference): callers that cannot call cudaMalloc — code under CUDA
// graph capture, tight serving loops, harnesses that forbid it — get the host-buffer ray batch with its device scratch (ray staging,
// hit staging, the warps' fetch cursor) in ONE file-scope device array of fixed size instead of the scene's grown-on-demand buffers.
// Every GPU that loads the library holds the array until exit.  All calls share it and it keeps the last call's contents: the cursor
// is cleared at the start of every call (it must start at zero), rays and hits are overwritten up to n; calls that overlap on the GPU
// would race on it, so the call synchronises before it returns and must not be issued from two host threads at once.
struct NoAllocScratch {
  uint32_t cursor; uint32_t pad[63];                 // 256 B: the fetch cursor of the launch in its own 256-byte line
  float4 rays[2 * SPT_NOALLOC_MAX_RAYS];             // (o, tmin) (d, tmax) per ray = SptRay
  HitIO hits[SPT_NOALLOC_MAX_RAYS];                  // = SptHit
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];
static_assert(sizeof(SptRay) == 2 * sizeof(float4) && sizeof(SptHit) == sizeof(HitIO), "scratch layout follows the ABI structs");

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per device,
// with that device current, and cached.
static NoAllocScratch* noalloc_scratch_ptr(cudaError_t* err) {
  static NoAllocScratch* ptr[64] = {};
  int dev = -1;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  if (dev < 0 || dev >= 64) { *err = cudaErrorInvalidDevice; return nullptr; }
  if (!ptr[dev]) {
    void* p = nullptr;
    *err = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (*err != cudaSuccess) return nullptr;
    ptr[dev] = (NoAllocScratch*)p;
  }
  return ptr[dev];
}

namespace spt {
// Closest-hit batch of n <= SPT_NOALLOC_MAX_RAYS rays, host buffers in and out, on stream `st` of the CURRENT device; synchronises.
// No cudaMalloc / cudaFree / allocating library call on this path.
cudaError_t trace_batch_noalloc(const DScene& sc, const SptRay* h_rays, uint32_t n, SptHit* h_hits, int grid, cudaStream_t st) {
  cudaError_t e;
  NoAllocScratch* S = noalloc_scratch_ptr(&e);
  if (!S) return e;
  if ((e = cudaMemsetAsync(&S->cursor, 0, sizeof(uint32_t), st)) != cudaSuccess) return e;          // must start at zero
  if ((e = cudaMemcpyAsync(S->rays, h_rays, (size_t)n * sizeof(SptRay), cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
  k_batch<false, false><<<grid, SPT_BLOCK, 0, st>>>(sc, S->rays, n, S->hits, nullptr, &S->cursor, nullptr);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  if ((e = cudaMemcpyAsync(h_hits, S->hits, (size_t)n * sizeof(SptHit), cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
  return cudaStreamSynchronize(st);
}
} // namespace spt
