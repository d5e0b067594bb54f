// Shared helpers for the pings_b200 CUDA sources (sm_100a only).
#pragma once
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/pings_b200.h"

#ifndef __CUDA_ARCH__
#define PG_HOST_ONLY 1
#endif

namespace pg {

extern thread_local char g_last_error[256];

inline int record_cuda_error(cudaError_t e, const char* what, int line) {
  snprintf(g_last_error, sizeof(g_last_error), "%s (line %d): %s", what, line, cudaGetErrorString(e));
  return PG_ERR_CUDA;
}

#define PG_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t _e = (call);                                                 \
    if (_e != cudaSuccess) return pg::record_cuda_error(_e, #call, __LINE__); \
  } while (0)

// After a kernel launch: always catch launch-configuration errors; in debug mode also
// synchronise so that asynchronous faults surface at the offending launch.
#define PG_LAUNCH_CHECK(debug, stream)                                        \
  do {                                                                        \
    cudaError_t _e = cudaGetLastError();                                      \
    if (_e != cudaSuccess) return pg::record_cuda_error(_e, "launch", __LINE__); \
    if (debug) {                                                              \
      _e = cudaStreamSynchronize(stream);                                     \
      if (_e != cudaSuccess) return pg::record_cuda_error(_e, "sync", __LINE__); \
    }                                                                         \
  } while (0)

// Opt-in phase timer (pg_api_core.cu): records a CUDA-event pair around a group of launches.
bool timing_on();
void phase_begin(const char* name, cudaStream_t st);
void phase_end(cudaStream_t st);
struct Phase {
  cudaStream_t st;
  Phase(const char* name, cudaStream_t s) : st(s) { phase_begin(name, s); }
  ~Phase() { phase_end(st); }
};

constexpr int kSmCount = 148;  // B200: 2 dies x 74 SMs

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over a caller-provided byte buffer (128-B aligned slices).
struct Carver {
  char* base;
  size_t off;
  explicit Carver(void* p) : base(reinterpret_cast<char*>(p)), off(0) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 128);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * sizeof(T);
    return p;
  }
  size_t total() const { return align_up(off, 128) + 128; }
};

// Programmatic dependent launch (sm_90+): kernels of a chain are launched with
// programmaticStreamSerialization, so their CTAs are set up while the previous kernel drains; `pdl_wait` blocks until
// the previous grid has completed and its writes are visible, and must precede every dependent access.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Every kernel of a chain calls pdl_wait() as its FIRST statement in every thread (a grid whose CTAs all left
// without waiting would let its own dependent start before the grid two steps back has finished).
template <typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(Args...), unsigned grid, unsigned block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  // Opt-in (PG_PDL=1).  Measured on B200, C2 step, alternating runs on one box (tools/gpu_ab_pdl.sh): the depth-sort
  // phase alone gets shorter (46 -> 40 us) but the whole step gets LONGER (0.529 -> 0.543 ms): the early-launched
  // CTAs sit on the SMs' CTA slots / shared memory while the previous kernel still runs.  Plain launches by default.
  static const bool on = [] { const char* e = getenv("PG_PDL"); return e && e[0] == '1'; }();
  cfg.numAttrs = on ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, args...);
}

__host__ __device__ inline int div_up(int a, int b) { return (a + b - 1) / b; }

}  // namespace pg
