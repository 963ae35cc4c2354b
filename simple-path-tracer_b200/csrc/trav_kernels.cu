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
