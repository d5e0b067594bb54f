// Shared helpers for the pings_b200 CUDA sources (sm_100a only).
#pragma once
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

__host__ __device__ inline int div_up(int a, int b) { return (a + b - 1) / b; }

}  // namespace pg
