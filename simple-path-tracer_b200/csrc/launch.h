// launch.h — host-callable launchers of the device kernels.  Each group of kernels lives in its own translation unit
// (trav_kernels.cu, shade_*.cu, misc_kernels.cu) so they compile in parallel; spt_api.cu holds only host code.
#pragma once
#include <cuda_runtime.h>
#include "../../include/spt.h"
#include "scene.h"
#include "wavefront_types.h"

#define SPT_TRAV_BLOCK 128  // threads per block of every traversal / shade kernel
#define SPT_LOGIC_BLOCK 256
#ifndef SPT_SHADE_LEAN_MINB
#define SPT_SHADE_LEAN_MINB 4 // resident blocks per SM the lean shade kernels are compiled for (and launched with: one wave).  Measured 4 / 5 / 6: C2 shade stage 22.1 / 23.4 / 23.6 ms (general kernels: 27.2)
#endif
#ifndef SPT_FETCH_CHUNK
#define SPT_FETCH_CHUNK 32  // rays a warp reserves at a time with one atomic: caller ray batches
#endif
// ... and in the wavefront loop: pool slots in extend, shadow-queue entries in connect.  64 measured +2..3 % over 32 on C2..C4 and
// +13 % on C1 at the render-sized pools (96 the same, 128 worse on C5); connect alone is 3 % faster at 32, but not the render
// with its overlapped streams (profiles/README.md)
#ifndef SPT_EXTEND_CHUNK
#define SPT_EXTEND_CHUNK 64
#endif
#ifndef SPT_CONNECT_CHUNK
#define SPT_CONNECT_CHUNK 64
#endif

struct HitIO { float t, u, v; int prim, inst; };

namespace spt {
// resident blocks per SM (occupancy API) of the persistent traversal kernels
cudaError_t occupancy_traversal(int* extend_blocks, int* connect_blocks, int* batch_blocks);
void launch_extend(const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st);
void launch_connect(const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st);
void launch_batch(const DScene& sc, const float4* rays, uint32_t n, HitIO* hits, uint8_t* occ, uint32_t* next, DevCounters* ctr,
                  bool any, bool count, int grid, cudaStream_t st);
// the no-allocation ray batch (trav_kernels.cu: scratch in a file-scope device array of SPT_NOALLOC_MAX_RAYS rays)
cudaError_t trace_batch_noalloc(const DScene& sc, const SptRay* h_rays, uint32_t n, SptHit* h_hits, int grid, cudaStream_t st);
void launch_logic(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st);
void launch_special(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st);
void launch_shade(int mat, const DScene& sc, const RenderParams& P, bool count, bool lean, int grid, cudaStream_t st); // mat = -1: unsorted (never lean)
void launch_generate(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st);
void launch_resolve(const long long* film, int n_pixels, float inv_spp, float* out_mean, uint8_t* out_rgb8, int grid, cudaStream_t st);
void launch_resolve_box(const RenderParams& P, int k0, int n_k, float radius, int radius_int, float* out_mean, uint8_t* out_rgb8, int grid, cudaStream_t st);
void launch_pack_tiles(const long long* film, const uint32_t* tiles, int n_tiles, const RenderParams& P, long long* out, int grid, cudaStream_t st);
void launch_unpack_tiles(long long* film, const uint32_t* all_tiles, int per_rank, int world, int root, const RenderParams& P, const long long* in, int grid, cudaStream_t st);
void launch_aov_normal(const DScene& sc, const RenderParams& P, int grid, cudaStream_t st);
void launch_bxdf(const DScene& sc, const SptBxdfQuery* q, long long n, SptBxdfResult* out, int grid, cudaStream_t st);
// defined per shade translation unit
void launch_shade_a(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st); // 0..3
void launch_shade_b(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st); // 4..6
void launch_shade_c(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st); // 7, -1
// the same kernels compiled with SPT_LEAN (scenes whose textures are all plain scalars; shade_*_lean.cu)
void launch_shade_a_lean(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st);
void launch_shade_b_lean(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st);
void launch_shade_c_lean(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st); // 7 only
} // namespace spt
