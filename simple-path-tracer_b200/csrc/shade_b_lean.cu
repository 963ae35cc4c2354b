// shade_b_lean.cu — the shade kernels of shade_b.cu compiled with SPT_LEAN = 1 (all textures plain scalars: texture.cuh).
#define SPT_LEAN 1
#include "shade_b.cu"
