// Depth ordering of the visible Gaussians for the band binning: a bucket sort instead of 4 LSD radix passes.
//
// What is ordered: the reference sorts 64-bit keys (tile << 32 | depth bits) with a stable radix sort whose input
// is emitted in Gaussian-index order (R3 rasterizer_impl.cu:70-110, 306-311), so inside a tile the list is ordered
// by (depth bits, Gaussian index).  That composite is UNIQUE per Gaussian, so any correct sort of the pairs
// (key << 32 | idx) reproduces the reference order bit for bit - stability is not needed.
//
//   preprocess   (r3/rs_preprocess_kernel) leaves max(key) and max(~key) of the visible Gaussians in `mm`
//   hist         bin = (key - kmin) >> shift over 2^lognb bins, linear in key-bit space; RED.ADD per key
//   scan         exclusive prefix of the bins (decoupled look-back over <= 128 CTAs) -> scatter cursors; bins are
//                grouped into buckets: bin b belongs to bucket floor(prefix[b] / kBucketTarget), so bucket
//                boundaries fall on bin boundaries and a bucket holds < kBucketTarget + max bin size pairs
//   scatter      pairs[atomicAdd(cursor[bin])] = (key, idx): unordered inside a bin
//   local sort   one warp (= one 32-thread CTA) per bucket: bitonic network held in registers (all comparators
//                ascending, padding with the +inf pair), writes sorted_idx and gathers the rects; buckets above 512
//                pairs (a bin holding hundreds of Gaussians within 2^shift depth-bit values of each other) are
//                sorted by the same warp with an LSD radix sort in global memory over the bits that differ.
// No grid-wide barriers, no MATCH.ANY ranking chains; the culled Gaussians (key 0xFFFFFFFF) never enter.
#pragma once
#include "pg_common.cuh"

namespace pg {

constexpr uint32_t kBucketTarget = 192;
constexpr int kBucketMinLog = 12, kBucketMaxLog = 19;
constexpr int kBucketKeys = 2;   // keys per thread in the histogram / scatter kernels (uint2 loads)

inline int bucket_log_bins(uint64_t n) {
  int l = kBucketMinLog;
  while (l < kBucketMaxLog && (1ull << (l + 2)) < n) ++l;   // about n / 4 bins
  return l;
}
inline size_t bucket_max_buckets(uint64_t n) { return (size_t)(n / kBucketTarget) + 4; }
// u32 words: bins | look-back words (u64 per scan CTA) | misc | bucket starts
inline size_t bucket_sort_scratch_elems(uint64_t n) {
  const size_t nb = (size_t)1 << bucket_log_bins(n);
  return nb + 2 * (nb / 4096) + 16 + bucket_max_buckets(n) + 4;
}

// words at the head of the scratch that must be zero before the histogram: bins | look-back words | misc
inline size_t bucket_sort_zero_words(uint64_t n) {
  const size_t nb = (size_t)1 << bucket_log_bins(n);
  return nb + 2 * (nb / 4096) + 16;
}

struct BucketSortArgs {
  const uint32_t* dkey;        // [n] depth bits, 0xFFFFFFFF = culled
  const uint32_t* mm;          // [2] max(key), max(~key) over the valid keys (0, 0 when there is none)
  uint32_t* bins;              // [2^lognb] counts, then scatter cursors (zeroed by the host)
  unsigned long long* look;    // [2^lognb / 4096] look-back words (zeroed by the host)
  uint32_t* misc;              // [0] number of buckets
  uint32_t* bstart;            // [buckets + 1] first pair of every bucket
  unsigned long long* pairs;   // [n] (key << 32 | idx), grouped by bin after the scatter
  const uint2* rect;           // [n]
  uint32_t* sorted_idx;        // [n] out: Gaussian ids in (depth bits, id) order, the V visible ones
  uint2* srect;                // [n] out: rect[sorted_idx[i]]
  uint32_t* nvalid;            // out: V
  uint32_t n;
  int lognb;
};

__device__ __forceinline__ void bucket_map(const uint32_t* __restrict__ mm, int lognb, uint32_t& kmin, int& shift) {
  const uint32_t kmax = mm[0];
  kmin = ~mm[1];
  const uint32_t range = kmax >= kmin ? kmax - kmin : 0u;
  const int bits = 32 - __clz(range);
  shift = bits > lognb ? bits - lognb : 0;
}

__global__ void __launch_bounds__(256) bucket_hist_kernel(const BucketSortArgs a) {
  pdl_trigger();
  uint32_t kmin; int shift;
  bucket_map(a.mm, a.lognb, kmin, shift);
  const uint32_t t = blockIdx.x * 256 + threadIdx.x;
  const uint32_t i0 = t * kBucketKeys;
  if (i0 >= a.n) return;
  uint32_t k[kBucketKeys];
  if (i0 + kBucketKeys <= a.n) {
    const uint2 v = reinterpret_cast<const uint2*>(a.dkey)[t];
    k[0] = v.x; k[1] = v.y;
  } else {
#pragma unroll
    for (int j = 0; j < kBucketKeys; ++j) k[j] = i0 + j < a.n ? a.dkey[i0 + j] : 0xffffffffu;
  }
#pragma unroll
  for (int j = 0; j < kBucketKeys; ++j)
    if (k[j] != 0xffffffffu) atomicAdd(&a.bins[(k[j] - kmin) >> shift], 1u);
}

// One CTA per 4096 bins.  bins[] becomes the exclusive prefix (the scatter cursor of the bin).
__global__ void __launch_bounds__(256) bucket_scan_kernel(const BucketSortArgs a) {
  __shared__ uint32_t wsum[8];
  __shared__ uint32_t s_prefix;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t base = blockIdx.x * 4096u + threadIdx.x * 16u;
  pdl_trigger();
  pdl_wait();
  uint32_t c[16];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint4 v = reinterpret_cast<const uint4*>(a.bins + base)[q];
    c[4 * q] = v.x; c[4 * q + 1] = v.y; c[4 * q + 2] = v.z; c[4 * q + 3] = v.w;
  }
  uint32_t t = 0;
#pragma unroll
  for (int j = 0; j < 16; ++j) t += c[j];
  uint32_t inc = t;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += v;
  }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  uint32_t wbase = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    if (w < warp) wbase += wsum[w];
    total += wsum[w];
  }
  if (threadIdx.x == 0) {
    __threadfence();
    atomicExch(&a.look[blockIdx.x], (1ull << 32) | total);
  }
  if (warp == 0) {
    uint32_t acc = 0;
    for (uint32_t i = lane; i < blockIdx.x; i += 32) {
      unsigned long long v;
      do { v = *reinterpret_cast<volatile unsigned long long*>(&a.look[i]); } while (!(v >> 32));
      acc += (uint32_t)v;
    }
    acc = __reduce_add_sync(0xffffffffu, acc);
    if (lane == 0) s_prefix = acc;
  }
  __syncthreads();
  uint32_t run = s_prefix + wbase + inc - t;
  uint32_t o[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const uint32_t s = run, e = run + c[j];
    o[j] = s;
    if (c[j]) {
      for (uint32_t k = s / kBucketTarget + 1; k <= e / kBucketTarget; ++k) a.bstart[k] = e;
    }
    run = e;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q)
    reinterpret_cast<uint4*>(a.bins + base)[q] = make_uint4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
  if (blockIdx.x == 0 && threadIdx.x == 0) a.bstart[0] = 0u;
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 255) {
    const uint32_t V = run, nbk = (V + kBucketTarget - 1) / kBucketTarget;
    a.bstart[nbk] = V;   // V % target == 0: the same value the bin rule wrote
    a.misc[0] = nbk;
    *a.nvalid = V;
  }
}

__global__ void __launch_bounds__(256) bucket_scatter_kernel(const BucketSortArgs a) {
  pdl_trigger();
  pdl_wait();
  uint32_t kmin; int shift;
  bucket_map(a.mm, a.lognb, kmin, shift);
  const uint32_t t = blockIdx.x * 256 + threadIdx.x;
  const uint32_t i0 = t * kBucketKeys;
  if (i0 >= a.n) return;
  uint32_t k[kBucketKeys];
  if (i0 + kBucketKeys <= a.n) {
    const uint2 v = reinterpret_cast<const uint2*>(a.dkey)[t];
    k[0] = v.x; k[1] = v.y;
  } else {
#pragma unroll
    for (int j = 0; j < kBucketKeys; ++j) k[j] = i0 + j < a.n ? a.dkey[i0 + j] : 0xffffffffu;
  }
  uint32_t pos[kBucketKeys];
#pragma unroll
  for (int j = 0; j < kBucketKeys; ++j)
    if (k[j] != 0xffffffffu) pos[j] = atomicAdd(&a.bins[(k[j] - kmin) >> shift], 1u);
#pragma unroll
  for (int j = 0; j < kBucketKeys; ++j)
    if (k[j] != 0xffffffffu) a.pairs[pos[j]] = ((unsigned long long)k[j] << 32) | (unsigned long long)(i0 + j);
}

__device__ __forceinline__ void ce64(unsigned long long& lo, unsigned long long& hi) {
  const unsigned long long x = lo, y = hi;
  const bool sw = y < x;
  lo = sw ? y : x;
  hi = sw ? x : y;
}
// keep min(a, o) when `lower`, max(a, o) otherwise
__device__ __forceinline__ unsigned long long pick64(unsigned long long a, unsigned long long o, bool lower) {
  return ((o < a) == lower) ? o : a;
}

// Bitonic network over 32*E elements held in registers, element i = lane*E + r.  Every comparator puts the minimum
// at the lower index, so +inf padding at the top never moves.  Stage sizes k <= E stay inside a lane (fully
// unrolled, E is 8 or 16); the cross-lane part is a LOOP over the lane-level stage size with the XOR masks as
// run-time values, so the code stays a few hundred instructions and is shared by every warp of the grid
// (a fully unrolled network is ~2,000 instructions executed once per warp: instruction-fetch bound).
template <int E>
__device__ __forceinline__ void reg_strides(unsigned long long (&a)[E]) {   // strides E/2 .. 1 inside the lane
#pragma unroll
  for (int j = E / 2; j >= 1; j >>= 1) {
#pragma unroll
    for (int r = 0; r < E; ++r)
      if ((r & j) == 0) ce64(a[r], a[r | j]);
  }
}

template <int E>
__device__ __forceinline__ void warp_bitonic(unsigned long long (&a)[E], int lane) {
#pragma unroll
  for (int k = 2; k <= E; k <<= 1) {
#pragma unroll
    for (int r = 0; r < E; ++r) {   // flip: partner = r ^ (k - 1)
      const int p = r ^ (k - 1);
      if (p > r) ce64(a[r], a[p]);
    }
#pragma unroll
    for (int j = k / 4; j >= 1; j >>= 1) {
#pragma unroll
      for (int r = 0; r < E; ++r)
        if ((r & j) == 0) ce64(a[r], a[r | j]);
    }
  }
#pragma unroll 1
  for (int lk = 2; lk <= 32; lk <<= 1) {   // stage size k = lk * E
    {
      // flip: partner element = (lane ^ (lk - 1), E - 1 - r)
      const int lm = lk - 1;
      const bool lower = (lane & (lk >> 1)) == 0;
#pragma unroll
      for (int r = 0; r < E / 2; ++r) {
        const int p = E - 1 - r;
        const unsigned long long o1 = __shfl_xor_sync(0xffffffffu, a[p], lm);
        const unsigned long long o2 = __shfl_xor_sync(0xffffffffu, a[r], lm);
        a[r] = pick64(a[r], o1, lower);
        a[p] = pick64(a[p], o2, lower);
      }
    }
#pragma unroll 1
    for (int lj = lk >> 2; lj >= 1; lj >>= 1) {   // strides lj * E
      const bool lower = (lane & lj) == 0;
#pragma unroll
      for (int r = 0; r < E; ++r) {
        const unsigned long long o = __shfl_xor_sync(0xffffffffu, a[r], lj);
        a[r] = pick64(a[r], o, lower);
      }
    }
    reg_strides<E>(a);
  }
}

template <int E>
__device__ __forceinline__ void bucket_warp_sort(const BucketSortArgs& a, uint32_t s, uint32_t n, int lane) {
  unsigned long long v[E];
#pragma unroll
  for (int r = 0; r < E; ++r) {
    const uint32_t i = lane * E + r;
    v[r] = i < n ? a.pairs[s + i] : ~0ull;
  }
  warp_bitonic<E>(v, lane);
#pragma unroll
  for (int r = 0; r < E; ++r) {
    const uint32_t i = lane * E + r;
    if (i < n) {
      const uint32_t id = (uint32_t)v[r];
      a.sorted_idx[s + i] = id;
      a.srect[s + i] = a.rect[id];
    }
  }
}

// Fallback for a bucket above 512 pairs: a bin that holds many Gaussians within 2^shift depth-bit values of each other
// (a facade seen head-on: tens of thousands of Gaussians inside a millimetre of depth).  One warp sorts the bucket with
// a stable LSD radix sort over the bits that actually differ - composite = (key - bucket min) << idx_bits | idx - in
// global memory, ping-ponging between the pair array and the bucket's own slice of `srect` (same 8 bytes per item,
// written for real afterwards).  O(n) per 8-bit digit with ballot ranking; a bitonic network in global memory (the
// first version of this path) needed ~100 ms for 50,000 pairs, this ~1 ms.
__device__ __forceinline__ unsigned long long* warp_lsd_sort_pairs(unsigned long long* src, unsigned long long* dst, uint32_t n,
                                                                   int lane, uint32_t kmin, int key_bits, int idx_bits,
                                                                   uint32_t* cnt /* shared [256] */) {
  const uint32_t below = (1u << lane) - 1u;
  const unsigned long long idx_mask = (1ull << idx_bits) - 1ull;
  for (int sh = 0; sh < key_bits + idx_bits; sh += 8) {
    for (int k = lane; k < 256; k += 32) cnt[k] = 0u;
    __syncwarp();
    for (uint32_t i = lane; i < n; i += 32) {
      const unsigned long long p = src[i];
      const unsigned long long c = ((unsigned long long)((uint32_t)(p >> 32) - kmin) << idx_bits) | (p & idx_mask);
      atomicAdd(&cnt[(uint32_t)(c >> sh) & 255u], 1u);
    }
    __syncwarp();
    {   // exclusive scan of the 256 counters: lane owns 8 consecutive ones
      uint32_t v[8], t = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) { v[k] = cnt[lane * 8 + k]; t += v[k]; }
      uint32_t inc = t;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
      }
      uint32_t run = inc - t;
      __syncwarp();
#pragma unroll
      for (int k = 0; k < 8; ++k) { cnt[lane * 8 + k] = run; run += v[k]; }
    }
    __syncwarp();
    for (uint32_t base = 0; base < n; base += 32) {
      const uint32_t i = base + lane;
      const bool valid = i < n;
      unsigned long long p = 0ull;
      uint32_t d = 0u;
      if (valid) {
        p = src[i];
        const unsigned long long c = ((unsigned long long)((uint32_t)(p >> 32) - kmin) << idx_bits) | (p & idx_mask);
        d = (uint32_t)(c >> sh) & 255u;
      }
      uint32_t m = __ballot_sync(0xffffffffu, valid);      // lanes holding the same digit
#pragma unroll
      for (int bit = 0; bit < 8; ++bit) {
        const uint32_t bb = __ballot_sync(0xffffffffu, (d >> bit) & 1u);
        m &= ((d >> bit) & 1u) ? bb : ~bb;
      }
      uint32_t pos = 0u;
      if (valid) pos = cnt[d] + __popc(m & below);
      __syncwarp();
      if (valid && (m & below) == 0u) cnt[d] += __popc(m);
      __syncwarp();
      if (valid) dst[pos] = p;
    }
    __syncwarp();
    unsigned long long* t = src; src = dst; dst = t;
  }
  return src;   // where the sorted bucket lives
}

// One bucket per 32-thread CTA: the bucket size is then uniform for the compiler as well (no WARPSYNC around the
// shuffles).  Buckets of up to 512 pairs are sorted in registers, larger ones by the LSD fallback above.
__global__ void __launch_bounds__(32) bucket_local_sort_kernel(const BucketSortArgs a) {
  __shared__ uint32_t cnt[256];
  const int lane = threadIdx.x;
  const uint32_t b = blockIdx.x;
  pdl_wait();
  if (b >= a.misc[0]) return;
  const uint32_t s = a.bstart[b], n = a.bstart[b + 1] - s;
  if (n == 0) return;
  if (n <= 256) {
    bucket_warp_sort<8>(a, s, n, lane);
  } else if (n <= 512) {
    bucket_warp_sort<16>(a, s, n, lane);
  } else {
    unsigned long long* d = a.pairs + s;
    uint32_t kmin = 0xffffffffu, kmax = 0u;
    for (uint32_t i = lane; i < n; i += 32) {
      const uint32_t k = (uint32_t)(d[i] >> 32);
      kmin = min(kmin, k);
      kmax = max(kmax, k);
    }
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    const int key_bits = 32 - __clz(kmax - kmin);
    const int idx_bits = max(1, 32 - __clz(a.n - 1u));
    unsigned long long* tmp = reinterpret_cast<unsigned long long*>(a.srect + s);
    unsigned long long* res = warp_lsd_sort_pairs(d, tmp, n, lane, kmin, key_bits, idx_bits, cnt);
    if (res != d) {
      for (uint32_t i = lane; i < n; i += 32) d[i] = res[i];
      __syncwarp();
    }
    for (uint32_t i = lane; i < n; i += 32) {
      const uint32_t id = (uint32_t)d[i];
      a.sorted_idx[s + i] = id;
      a.srect[s + i] = a.rect[id];
    }
  }
}

// Host side: four launches on `st`.  `scratch` holds bucket_sort_scratch_elems(n) words, of which the first
// bucket_sort_zero_words(n) were zeroed by the caller (before the kernel that writes `mm`).
inline cudaError_t depth_bucket_sort(const uint32_t* dkey, const uint32_t* mm, uint32_t* scratch, unsigned long long* pairs,
                                     const uint2* rect, uint32_t* sorted_idx, uint2* srect, uint32_t* nvalid, uint32_t n,
                                     cudaStream_t st) {
  BucketSortArgs a;
  a.dkey = dkey; a.mm = mm; a.n = n; a.lognb = bucket_log_bins(n);
  const size_t nb = (size_t)1 << a.lognb, nscan = nb / 4096;
  a.bins = scratch;
  a.look = reinterpret_cast<unsigned long long*>(scratch + nb);
  a.misc = scratch + nb + 2 * nscan;
  a.bstart = a.misc + 16;
  a.pairs = pairs; a.rect = rect; a.sorted_idx = sorted_idx; a.srect = srect; a.nvalid = nvalid;
  const uint32_t g4 = (uint32_t)(((size_t)n + 256 * kBucketKeys - 1) / (256 * kBucketKeys));
  bucket_hist_kernel<<<g4, 256, 0, st>>>(a);
  cudaError_t e = launch_pdl(bucket_scan_kernel, (unsigned)nscan, 256u, 0, st, a);
  if (e != cudaSuccess) return e;
  e = launch_pdl(bucket_scatter_kernel, g4, 256u, 0, st, a);
  if (e != cudaSuccess) return e;
  e = launch_pdl(bucket_local_sort_kernel, (unsigned)bucket_max_buckets(n), 32u, 0, st, a);
  if (e != cudaSuccess) return e;
  return cudaGetLastError();
}

}  // namespace pg
