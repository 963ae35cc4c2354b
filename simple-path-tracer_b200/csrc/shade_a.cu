// shade_a.cu — shade kernels of the material classes lambert, pseudo, conductor, dielectric.
#include "launch.h"
#include "shade_kernels.cuh"
namespace spt {
void launch_shade_a(int mat, const DScene& sc, const RenderParams& P, bool, int grid, cudaStream_t st) {
  switch (mat) {
    case 0: k_shade<0, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    case 1: k_shade<1, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    case 2: k_shade<2, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    default: k_shade<3, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
  }
}
} // namespace spt
