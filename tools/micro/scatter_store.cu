// Micro-benchmark: cost of scattered 4-byte stores on B200 (informs the binning scatter design).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scatter_store scatter_store.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x;
}

// mode 0: every lane a random 4-B slot in [0, n_slots)
// mode 1: groups of G lanes share a random 32-B-aligned run (G consecutive words)
template <int G>
__global__ void scatter(uint32_t* out, uint32_t n_slots, uint32_t iters, uint32_t seed) {
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t it = 0; it < iters; ++it) {
    const uint32_t grp = (tid / G) * 7919u + it * 104729u + seed;
    uint32_t slot = (mix(grp) % (n_slots / G)) * G + (lane % G);
    out[slot] = tid + it;
  }
}

template <int G>
float run(uint32_t* out, uint32_t n_slots, uint32_t total, int grid, int block) {
  uint32_t iters = total / (grid * block);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  scatter<G><<<grid, block>>>(out, n_slots, iters, 1);
  cudaDeviceSynchronize();
  cudaEventRecord(a);
  scatter<G><<<grid, block>>>(out, n_slots, iters, 2);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms * 1e3f;
}

int main() {
  const uint32_t total = 22700000;
  uint32_t* out; cudaMalloc(&out, 256u << 20);
  for (uint32_t mb : {8u, 32u, 91u, 256u}) {
    uint32_t n_slots = mb << 18;
    for (int grid : {284, 592, 1184}) {
      printf("region %3u MB grid %4d x256:  G=1 %7.1f us  G=2 %7.1f  G=4 %7.1f  G=8 %7.1f  G=32 %7.1f\n", mb, grid,
             run<1>(out, n_slots, total, grid, 256), run<2>(out, n_slots, total, grid, 256),
             run<4>(out, n_slots, total, grid, 256), run<8>(out, n_slots, total, grid, 256),
             run<32>(out, n_slots, total, grid, 256));
    }
  }
  return 0;
}
