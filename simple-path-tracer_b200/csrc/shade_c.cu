// shade_c.cu — shade kernels with the BSSRDF probe: subsurface, and the unsorted all-materials kernel (SPT_RENDER_NO_SORT).
// Compiled twice: as is, and with SPT_LEAN = 1 through shade_c_lean.cu (subsurface only; see texture.cuh).
#ifndef SPT_LEAN
#define SPT_LEAN 0
#endif
#include "launch.h"
#include "shade_kernels.cuh"
namespace spt {
#if SPT_LEAN
void launch_shade_c_lean(int, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st) {
  if (count) k_shade<7, true, 1><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_shade<7, false, 1><<<grid, SPT_BLOCK, 0, st>>>(sc, P);
}
#else
void launch_shade_c(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st) {
  if (mat == 7) { if (count) k_shade<7, true, 0><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_shade<7, false, 0><<<grid, SPT_BLOCK, 0, st>>>(sc, P); }
  else { if (count) k_shade<-1, true, 0><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_shade<-1, false, 0><<<grid, SPT_BLOCK, 0, st>>>(sc, P); }
}
#endif
} // namespace spt
