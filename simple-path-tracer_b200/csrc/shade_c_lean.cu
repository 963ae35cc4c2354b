// shade_c_lean.cu — the shade kernels of shade_c.cu compiled with SPT_LEAN = 1 (all textures plain scalars: texture.cuh).
#define SPT_LEAN 1
#include "shade_c.cu"
