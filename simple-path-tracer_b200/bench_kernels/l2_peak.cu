// l2_peak.cu — measurement helper, not part of the product ABI: the L2 -> SM read bandwidth of this GPU for 16-byte loads
// over a read-only set that fits in L2 (the roof that bounds BVH node / triangle fetches when the scene is
// L2-resident; MEASURED_PEAKS.json only has the HBM and tensor figures).  Loads bypass L1 (ld.global.cg) so every request
// is served by L2; the buffer is walked once untimed to make it resident.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

__global__ void __launch_bounds__(256) k_l2_read(const uint4* __restrict__ buf, size_t n_vec, int passes, uint4* sink) {
  uint4 acc = make_uint4(0, 0, 0, 0);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int p = 0; p < passes; p++) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    // 4 independent loads in flight per thread
    for (; i + 3 * stride < n_vec; i += 4 * stride) {
      uint4 a, b, c, d;
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(buf + i));
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(buf + i + stride));
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(c.x), "=r"(c.y), "=r"(c.z), "=r"(c.w) : "l"(buf + i + 2 * stride));
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(d.x), "=r"(d.y), "=r"(d.z), "=r"(d.w) : "l"(buf + i + 3 * stride));
      acc.x ^= a.x ^ b.x ^ c.x ^ d.x; acc.y ^= a.y ^ b.y ^ c.y ^ d.y; acc.z ^= a.z ^ b.z ^ c.z ^ d.z; acc.w ^= a.w ^ b.w ^ c.w ^ d.w;
    }
    for (; i < n_vec; i += stride) {
      uint4 a;
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(buf + i));
      acc.x ^= a.x; acc.y ^= a.y; acc.z ^= a.z; acc.w ^= a.w;
    }
  }
  if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x9e3779b9u) *sink = acc; // never true for the zero-filled buffer's xor pattern: keeps the loads alive
}

// Returns 0 on success; *gbs = bytes read per second / 1e9 (best of `reps`).
extern "C" int spt_bench_l2_read_gbs(int device, size_t bytes, int passes, int reps, double* gbs) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 1;
  uint4 *buf = nullptr, *sink = nullptr;
  const size_t n_vec = bytes / sizeof(uint4);
  if (cudaMalloc(&buf, n_vec * sizeof(uint4)) != cudaSuccess || cudaMalloc(&sink, sizeof(uint4)) != cudaSuccess) return 2;
  cudaMemset(buf, 0x5a, n_vec * sizeof(uint4));
  const int grid = prop.multiProcessorCount * 8;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_l2_read<<<grid, 256>>>(buf, n_vec, 2, sink); // warm: make the set L2-resident
  double best = 0.0;
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(e0);
    k_l2_read<<<grid, 256>>>(buf, n_vec, passes, sink);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(buf); cudaFree(sink); return 3; }
    float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
    const double g = (double)n_vec * sizeof(uint4) * passes / (ms * 1e-3) / 1e9;
    if (g > best) best = g;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(buf); cudaFree(sink);
  *gbs = best;
  return 0;
}
