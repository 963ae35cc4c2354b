// shade_a_lean.cu — the shade kernels of shade_a.cu compiled with SPT_LEAN = 1 (all textures plain scalars: texture.cuh).
#define SPT_LEAN 1
#include "shade_a.cu"
