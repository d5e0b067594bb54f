// Gradient exchange of the data-parallel step over NVSwitch multicast: ONE kernel that is barrier +
// in-switch reduction + broadcast, instead of ncclAllReduce.  The reference is single-GPU: this
// has no counterpart there.
//
// Plumbing (allocation of the symmetric buffer, exchange of the handles, the multicast mapping, the signal pads) is
// torch.distributed._symmetric_memory on the host side (pings_b200/parallel.py: MultimemGradAllReduce); the data
// path below is ours:
//   * every rank owns the slice [n4 r / N, n4 (r+1) / N) of the flat fp32 buffer (float4 units);
//   * multimem.ld_reduce on the MULTICAST address of an element makes the switch read that element from all N
//     replicas and return their fp32 sum (one load on the wire per rank instead of N-1 peer reads);
//   * multimem.st writes the sum back to all N replicas at once.
//   Every element is reduced exactly once, by its owner, and the same bits are broadcast: replicas stay identical.
//   * start barrier: CTA b of every rank signals CTA b of every peer through the signal pads (release) and waits
//     for theirs (acquire), so a rank starts pulling only when every peer's kernel has started, i.e. when every
//     peer's backward (earlier on its stream) has finished writing its gradients;
//   * end barrier, same pattern: the kernel of a rank completes only when every CTA of every peer has pushed its
//     slice, so the next kernel on the stream sees the complete sum.
// All ranks must launch the same grid.  A spin that lasts more than ~4 s traps instead of hanging the GPU.
#include "pg_common.cuh"

namespace pg {

__device__ __forceinline__ void mm_put_signal(uint32_t* addr) {
  uint32_t old;
  const long long t0 = clock64();
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 0u && clock64() - t0 > 8000000000ll) __trap();
  } while (old != 0u);
}
__device__ __forceinline__ void mm_wait_signal(uint32_t* addr) {
  uint32_t old;
  const long long t0 = clock64();
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(addr) : "memory");
    if (old != 1u && clock64() - t0 > 8000000000ll) __trap();
  } while (old != 1u);
}
// CTA b of this rank <-> CTA b of every rank (itself included: the self slot keeps the code uniform).
__device__ __forceinline__ void mm_barrier(uint32_t* const* pads, int rank, int world) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    mm_put_signal(pads[peer] + (size_t)blockIdx.x * world + rank);
    mm_wait_signal(pads[rank] + (size_t)blockIdx.x * world + peer);
  }
  __syncthreads();
}

constexpr int kMmThreads = 512;
constexpr int kMmUnroll = 4;

__global__ void __launch_bounds__(kMmThreads) multimem_allreduce_kernel(float4* __restrict__ mc, uint32_t* const* __restrict__ pads,
                                                                        size_t lo, size_t hi, int rank, int world, float scale) {
  mm_barrier(pads, rank, world);
  const size_t stride = (size_t)gridDim.x * kMmThreads;
  for (size_t i0 = lo + (size_t)blockIdx.x * kMmThreads + threadIdx.x; i0 < hi; i0 += stride * kMmUnroll) {
    float4 v[kMmUnroll];
#pragma unroll
    for (int u = 0; u < kMmUnroll; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                     : "l"(mc + i)
                     : "memory");
    }
#pragma unroll
    for (int u = 0; u < kMmUnroll; ++u) {
      const size_t i = i0 + (size_t)u * stride;
      if (i < hi)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + i), "f"(v[u].x * scale),
                     "f"(v[u].y * scale), "f"(v[u].z * scale), "f"(v[u].w * scale)
                     : "memory");
    }
  }
  __threadfence_system();
  mm_barrier(pads, rank, world);
}

}  // namespace pg

using namespace pg;

extern "C" {

int pg_multimem_allreduce_f32(void* multicast_ptr, const void* signal_pad_ptrs_dev, size_t signal_pad_bytes, int64_t numel,
                              int rank, int world, float scale, int max_ctas, void* stream) {
  if (!multicast_ptr || !signal_pad_ptrs_dev || numel < 0 || world < 1 || rank < 0 || rank >= world) return PG_ERR_ARG;
  if (numel % 4 != 0 || (reinterpret_cast<uintptr_t>(multicast_ptr) & 15u)) return PG_ERR_ARG;
  if (world > 32) return PG_ERR_UNSUPPORTED;
  const size_t slots = signal_pad_bytes / (sizeof(uint32_t) * (size_t)world);
  if (slots < 1) return PG_ERR_WORKSPACE;
  // A few CTAs saturate the transfer (28 MB at N = 8: 0.084 ms standalone with 8 CTAs) and leave the SMs to the kernels
  // the exchange overlaps.  C2 step, ms: N = 8: 32 CTAs 0.656, 8 CTAs 0.604 (0.590 on a high-priority stream), 4 CTAs
  // 0.593; N = 4: 16 CTAs 0.589, 12 CTAs 0.570, 8 CTAs 0.568; N = 2 (14 MB slices): 8 CTAs 0.586, 16 CTAs 0.569,
  // 32 CTAs 0.597 (ncclAllReduce 0.565).
  int ctas = max_ctas > 0 ? max_ctas : (world <= 2 ? 16 : 8);
  if ((size_t)ctas > slots) ctas = (int)slots;
  const size_t n4 = (size_t)numel / 4;
  const size_t lo = n4 * (size_t)rank / (size_t)world, hi = n4 * (size_t)(rank + 1) / (size_t)world;
  multimem_allreduce_kernel<<<ctas, kMmThreads, 0, (cudaStream_t)stream>>>(
      static_cast<float4*>(multicast_ptr), static_cast<uint32_t* const*>(signal_pad_ptrs_dev), lo, hi, rank, world, scale);
  PG_LAUNCH_CHECK(0, (cudaStream_t)stream);
  return PG_OK;
}

}  // extern "C"
