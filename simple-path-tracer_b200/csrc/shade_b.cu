// shade_b.cu — shade kernels of the material classes plastic, pbr_metallic, pbr_specular.
// Compiled twice: as is, and with SPT_LEAN = 1 through shade_b_lean.cu (see texture.cuh).
#ifndef SPT_LEAN
#define SPT_LEAN 0
#endif
#include "launch.h"
#include "shade_kernels.cuh"
namespace spt {
#if SPT_LEAN
void launch_shade_b_lean(int mat, const DScene& sc, const RenderParams& P, bool, int grid, cudaStream_t st) {
#else
void launch_shade_b(int mat, const DScene& sc, const RenderParams& P, bool, int grid, cudaStream_t st) {
#endif
  switch (mat) {
    case 4: k_shade<4, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    case 5: k_shade<5, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    default: k_shade<6, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
  }
}
} // namespace spt
