// shade_c.cu — shade kernels with the BSSRDF probe: subsurface, and the unsorted all-materials kernel (SPT_RENDER_NO_SORT).
#include "launch.h"
#include "shade_kernels.cuh"
namespace spt {
void launch_shade_c(int mat, const DScene& sc, const RenderParams& P, bool count, int grid, cudaStream_t st) {
  if (mat == 7) { if (count) k_shade<7, true><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_shade<7, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); }
  else { if (count) k_shade<-1, true><<<grid, SPT_BLOCK, 0, st>>>(sc, P); else k_shade<-1, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); }
}
} // namespace spt
