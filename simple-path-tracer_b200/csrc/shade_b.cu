// shade_b.cu — shade kernels of the material classes plastic, pbr_metallic, pbr_specular.
#include "launch.h"
#include "shade_kernels.cuh"
namespace spt {
void launch_shade_b(int mat, const DScene& sc, const RenderParams& P, bool, int grid, cudaStream_t st) {
  switch (mat) {
    case 4: k_shade<4, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    case 5: k_shade<5, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
    default: k_shade<6, false><<<grid, SPT_BLOCK, 0, st>>>(sc, P); break;
  }
}
} // namespace spt
