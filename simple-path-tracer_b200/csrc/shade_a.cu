// shade_a.cu — shade kernels of the material classes lambert, pseudo, conductor, dielectric.
// Compiled twice: as is, and with SPT_LEAN = 1 through shade_a_lean.cu (see texture.cuh).
#ifndef SPT_LEAN
#define SPT_LEAN 0
#endif
#include "launch.h"
#include "shade_kernels.cuh"
namespace spt {
#if SPT_LEAN
void launch_shade_a_lean(int mat, const DScene& sc, const RenderParams& P, bool, int grid, cudaStream_t st) {
#else
void launch_shade_a(int mat, const DScene& sc, const RenderParams& P, bool, int grid, cudaStream_t st) {
#endif
  switch (mat) {
    case 0: k_shade<0, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    case 1: k_shade<1, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    case 2: k_shade<2, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    default: k_shade<3, false, SPT_LEAN><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
  }
}
} // namespace spt
