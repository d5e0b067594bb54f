// CUB-free device primitives: exclusive scan (u32) and a stable LSD radix-sort pass
// (hist -> scan -> scatter) for (key, u32 value) pairs.  Replaces the reference's
// cub::DeviceScan::InclusiveSum / cub::DeviceRadixSort::SortPairs call sites
// (R3 cuda_rasterizer/rasterizer_impl.cu:280,306-311).
//
// Design notes (B200): all kernels are pure HBM streams; a block owns a contiguous chunk of
// 4096 items (256 threads x 16), loads are 32-lane coalesced, ranking inside a warp uses
// MATCH.ANY so equal digits cost one shared-memory update per distinct digit (tile keys
// arrive in long equal runs, so plain shared atomics would serialise).
#pragma once
#include "pg_common.cuh"
#include <cooperative_groups.h>

namespace pg {

constexpr int kScanThreads = 256;
constexpr int kScanIpt = 16;
constexpr int kScanChunk = kScanThreads * kScanIpt;  // 4096

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// Block-wide exclusive scan of one value per thread (256 threads). Returns exclusive prefix,
// writes block total to *total (all threads receive it).
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* total) {
  __shared__ uint32_t wsum[8];
  __shared__ uint32_t tot;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = warp_incl_scan(v, lane);
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < 8 ? wsum[lane] : 0u;
    uint32_t wi = warp_incl_scan(w, lane);
    if (lane < 8) wsum[lane] = wi - w;
    if (lane == 7) tot = wi;
  }
  __syncthreads();
  uint32_t r = inc - v + wsum[warp];
  *total = tot;
  __syncthreads();
  return r;
}

// ---- scan: level 1 reduce ------------------------------------------------------------------
template <bool GATHER>
__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(
    const uint32_t* __restrict__ in, const uint32_t* __restrict__ idx, uint32_t n,
    uint32_t* __restrict__ block_sums, const uint32_t* __restrict__ n_dev = nullptr) {
  if (n_dev) n = min(n, *n_dev);   // device-side length (<= the launch bound n)
  const uint32_t base = blockIdx.x * kScanChunk;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < kScanIpt; ++k) {
    uint32_t i = base + k * kScanThreads + threadIdx.x;
    if (i < n) s += GATHER ? in[idx[i]] : in[i];
  }
  __shared__ uint32_t red[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < 8; ++w) t += red[w];
    block_sums[blockIdx.x] = t;
  }
}

// ---- scan: level 2, single block, exclusive in place ---------------------------------------
__global__ void __launch_bounds__(1024) scan_sums_kernel(uint32_t* __restrict__ sums, uint32_t m,
                                                          uint32_t* __restrict__ total_out) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (uint32_t base = 0; base < m; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = i < m ? sums[i] : 0u;
    uint32_t inc = warp_incl_scan(v, lane);
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = wsum[lane];
      uint32_t wi = warp_incl_scan(w, lane);
      wsum[lane] = wi - w;
    }
    __syncthreads();
    uint32_t carry = carry_s;
    uint32_t ex = inc - v + wsum[warp] + carry;
    if (i < m) sums[i] = ex;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = ex + v;
    __syncthreads();
  }
  if (total_out && threadIdx.x == 0) *total_out = carry_s;
}

// ---- scan: level 1 downsweep (exclusive) ----------------------------------------------------
template <bool GATHER>
__global__ void __launch_bounds__(kScanThreads) scan_down_kernel(
    const uint32_t* __restrict__ in, const uint32_t* __restrict__ idx, uint32_t n,
    const uint32_t* __restrict__ block_sums, uint32_t* __restrict__ out,
    const uint32_t* __restrict__ n_dev = nullptr) {
  if (n_dev) {
    n = min(n, *n_dev);
    if (blockIdx.x * kScanChunk >= n) return;   // uniform per block
  }
  // thread t owns items [base + t*16, base + t*16 + 16)
  const uint32_t base = blockIdx.x * kScanChunk + threadIdx.x * kScanIpt;
  uint32_t v[kScanIpt];
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < kScanIpt; ++k) {
    uint32_t i = base + k;
    uint32_t x = 0;
    if (i < n) x = GATHER ? in[idx[i]] : in[i];
    v[k] = s;
    s += x;
  }
  uint32_t tot;
  uint32_t ex = block_excl_scan_256(s, &tot) + block_sums[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanIpt; ++k) {
    uint32_t i = base + k;
    if (i < n) out[i] = v[k] + ex;
  }
}

inline size_t scan_scratch_elems(size_t n) { return (n + kScanChunk - 1) / kScanChunk + 1; }

// Exclusive scan of n u32 (optionally gathered through idx). `scratch` holds
// scan_scratch_elems(n) u32. If total_out != nullptr the grand total is written there.
// in == out is allowed when idx == nullptr.
inline cudaError_t exclusive_scan_u32(const uint32_t* in, const uint32_t* idx, uint32_t n,
                                      uint32_t* out, uint32_t* scratch, uint32_t* total_out,
                                      cudaStream_t st) {
  if (n == 0) {
    if (total_out) return cudaMemsetAsync(total_out, 0, sizeof(uint32_t), st);
    return cudaSuccess;
  }
  const uint32_t nblk = (n + kScanChunk - 1) / kScanChunk;
  if (idx) scan_reduce_kernel<true><<<nblk, kScanThreads, 0, st>>>(in, idx, n, scratch);
  else     scan_reduce_kernel<false><<<nblk, kScanThreads, 0, st>>>(in, idx, n, scratch);
  scan_sums_kernel<<<1, 1024, 0, st>>>(scratch, nblk, total_out);
  if (idx) scan_down_kernel<true><<<nblk, kScanThreads, 0, st>>>(in, idx, n, scratch, out);
  else     scan_down_kernel<false><<<nblk, kScanThreads, 0, st>>>(in, idx, n, scratch, out);
  return cudaGetLastError();
}

// In-place exclusive scan whose true length *n_dev (<= n_bound) is only known on the device: the grid
// is sized for n_bound, blocks past the true length contribute 0 and write nothing.
inline cudaError_t exclusive_scan_u32_dyn(uint32_t* data, uint32_t n_bound, const uint32_t* n_dev,
                                          uint32_t* scratch, cudaStream_t st) {
  if (n_bound == 0) return cudaSuccess;
  const uint32_t nblk = (n_bound + kScanChunk - 1) / kScanChunk;
  scan_reduce_kernel<false><<<nblk, kScanThreads, 0, st>>>(data, nullptr, n_bound, scratch, n_dev);
  scan_sums_kernel<<<1, 1024, 0, st>>>(scratch, nblk, nullptr);
  scan_down_kernel<false><<<nblk, kScanThreads, 0, st>>>(data, nullptr, n_bound, scratch, data, n_dev);
  return cudaGetLastError();
}

// ---- radix pass ---------------------------------------------------------------------------
constexpr int kSortThreads = 256;
constexpr int kSortWarps = 8;
constexpr int kSortIpt = 16;
constexpr int kSortChunk = kSortThreads * kSortIpt;  // 4096
constexpr int kSortWarpItems = 32 * kSortIpt;         // 512

// hist[d * nblk + blk] = number of items of block `blk` with digit d.
// n_dev (optional): device-side item count (<= n, the launch bound).  DROP: keys equal to all-ones are skipped
// (compacting first pass of the depth sort: culled Gaussians).
template <typename KeyT, bool DROP = false>
__global__ void __launch_bounds__(kSortThreads) radix_hist_kernel(
    const KeyT* __restrict__ keys, uint32_t n, int shift, uint32_t nbins,
    uint32_t* __restrict__ hist, uint32_t nblk, const uint32_t* __restrict__ n_dev = nullptr) {
  __shared__ uint32_t h[256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (n_dev) n = min(n, *n_dev);
  h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * kSortChunk + warp * kSortWarpItems;
  const uint32_t mask = nbins - 1;
  uint32_t d[kSortIpt];
#pragma unroll
  for (int it = 0; it < kSortIpt; ++it) {
    uint32_t i = base + it * 32 + lane;
    d[it] = 0xffffffffu;
    if (i < n) {
      const KeyT kk = keys[i];
      if (!(DROP && kk == (KeyT)~(KeyT)0)) d[it] = (uint32_t)(kk >> shift) & mask;
    }
  }
#pragma unroll
  for (int it = 0; it < kSortIpt; ++it)
    if (d[it] != 0xffffffffu) atomicAdd(&h[d[it]], 1u);
  __syncthreads();
  if (threadIdx.x < nbins) hist[threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}

// Stable scatter. `offs` is the exclusive scan of the digit-major histogram matrix.
// IOTA: values are generated as the item index (first pass of an argsort).
template <typename KeyT, bool IOTA, bool WRITE_KEYS, bool DROP = false>
__global__ void __launch_bounds__(kSortThreads) radix_scatter_kernel(
    const KeyT* __restrict__ kin, const uint32_t* __restrict__ vin, KeyT* __restrict__ kout,
    uint32_t* __restrict__ vout, uint32_t n, int shift, uint32_t nbins,
    const uint32_t* __restrict__ offs, uint32_t nblk, const uint32_t* __restrict__ n_dev = nullptr) {
  __shared__ uint32_t wh[kSortWarps][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (n_dev) {
    n = min(n, *n_dev);
    if (blockIdx.x * kSortChunk >= n) return;   // uniform per block
  }
#pragma unroll
  for (int w = 0; w < kSortWarps; ++w) wh[w][threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * kSortChunk + warp * kSortWarpItems;
  const uint32_t mask = nbins - 1;
  KeyT k[kSortIpt];
  uint32_t v[kSortIpt];
#pragma unroll
  for (int it = 0; it < kSortIpt; ++it) {
    uint32_t i = base + it * 32 + lane;
    if (i < n) {
      k[it] = kin[i];
      v[it] = IOTA ? i : vin[i];
    } else {
      k[it] = 0;
      v[it] = 0;
    }
  }
  // per-warp digit counts
#pragma unroll
  for (int it = 0; it < kSortIpt; ++it) {
    uint32_t i = base + it * 32 + lane;
    uint32_t d = (i < n && !(DROP && k[it] == (KeyT)~(KeyT)0)) ? ((uint32_t)(k[it] >> shift) & mask) : 0xffffffffu;
    if (d != 0xffffffffu) atomicAdd(&wh[warp][d], 1u);
  }
  __syncthreads();
  // digit = threadIdx.x: CTA-local start of the digit (exclusive scan of the CTA's digit counts) and the distance
  // from a CTA-local slot to the global position (offs = global start of this CTA's items of the digit)
  __shared__ KeyT sk[kSortChunk];
  __shared__ uint32_t sv[kSortChunk];
  __shared__ uint32_t gdelta[256];
  uint32_t cta_items;
  {
    uint32_t mine = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) mine += wh[w][threadIdx.x];
    uint32_t run = block_excl_scan_256(mine, &cta_items);
    if (threadIdx.x < nbins) gdelta[threadIdx.x] = offs[threadIdx.x * nblk + blockIdx.x] - run;
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) {
      uint32_t c = wh[w][threadIdx.x];
      wh[w][threadIdx.x] = run;
      run += c;
    }
  }
  __syncthreads();
  const uint32_t lt = (1u << lane) - 1u;
  // exchange through shared memory so that the global stores go out in per-digit runs (see depth_sort_coop_kernel)
  uint32_t peers[kSortIpt];
#pragma unroll
  for (int it = 0; it < kSortIpt; ++it) {   // 16 independent MATCH.ANY in flight (see depth_sort_coop_kernel)
    const uint32_t i = base + it * 32 + lane;
    const bool valid = i < n && !(DROP && k[it] == (KeyT)~(KeyT)0);
    peers[it] = __match_any_sync(0xffffffffu, valid ? ((uint32_t)(k[it] >> shift) & mask) : 0xffffffffu);
  }
#pragma unroll
  for (int it = 0; it < kSortIpt; ++it) {
    uint32_t i = base + it * 32 + lane;
    bool valid = i < n && !(DROP && k[it] == (KeyT)~(KeyT)0);
    uint32_t d = valid ? ((uint32_t)(k[it] >> shift) & mask) : 0xffffffffu;
    if (valid) {
      const uint32_t slot = wh[warp][d] + __popc(peers[it] & lt);
      sk[slot] = k[it];
      sv[slot] = v[it];
    }
    __syncwarp();
    if (valid && lane == (__ffs(peers[it]) - 1)) wh[warp][d] += __popc(peers[it]);
    __syncwarp();
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < kSortIpt; ++q) {
    const uint32_t slot = q * kSortThreads + threadIdx.x;
    if (slot < cta_items) {
      const KeyT key = sk[slot];
      const uint32_t pos = slot + gdelta[(uint32_t)(key >> shift) & mask];
      if (WRITE_KEYS) kout[pos] = key;
      vout[pos] = sv[slot];
    }
  }
}

inline uint32_t sort_num_blocks(uint64_t n) { return (uint32_t)((n + kSortChunk - 1) / kSortChunk); }
// u32 elements of scratch needed by one radix pass over n items (hist matrix + scan scratch).
inline size_t sort_scratch_elems(uint64_t n) {
  size_t hm = (size_t)sort_num_blocks(n) * 256;
  return hm + scan_scratch_elems(hm) + 64;
}

// One stable pass on bits [shift, shift+bits) (bits <= 8).
// n_dev: optional device-side item count (the grids are sized for n).  DROP: skip all-ones keys; the number of
// kept items is then written to *kept_out (device) by the histogram scan.
template <typename KeyT, bool IOTA, bool WRITE_KEYS, bool DROP = false>
inline cudaError_t radix_pass(const KeyT* kin, const uint32_t* vin, KeyT* kout, uint32_t* vout,
                              uint32_t n, int shift, int bits, uint32_t* scratch, cudaStream_t st,
                              const uint32_t* n_dev = nullptr, uint32_t* kept_out = nullptr) {
  if (n == 0) {
    if (kept_out) return cudaMemsetAsync(kept_out, 0, sizeof(uint32_t), st);
    return cudaSuccess;
  }
  const uint32_t nblk = sort_num_blocks(n);
  const uint32_t nbins = 1u << bits;
  uint32_t* hist = scratch;
  uint32_t* scan_scratch = scratch + (size_t)nblk * 256;
  radix_hist_kernel<KeyT, DROP><<<nblk, kSortThreads, 0, st>>>(kin, n, shift, nbins, hist, nblk, n_dev);
  cudaError_t e = exclusive_scan_u32(hist, nullptr, nbins * nblk, hist, scan_scratch, kept_out, st);
  if (e != cudaSuccess) return e;
  radix_scatter_kernel<KeyT, IOTA, WRITE_KEYS, DROP><<<nblk, kSortThreads, 0, st>>>(
      kin, vin, kout, vout, n, shift, nbins, hist, nblk, n_dev);
  return cudaGetLastError();
}

// ---- fused depth argsort (one cooperative launch) ---------------------------------------------
// The P-sized depth sort is latency bound (32 MB of traffic, but 12 launches + 4 three-kernel scans
// when run pass by pass).  For P <= kCoopSortMaxBlocks * 4096 the four LSD passes, the depth-order
// gather of the tile rects and the exclusive scan of the binning weights run as ONE cooperative
// kernel: a CTA keeps its 4096 (key, value) pairs in registers between the counting and the scatter
// half of a pass.  Cross-CTA digit prefixes without a scan phase: every CTA stores its 256 digit counts as
// a row of the [block][digit] matrix and adds them (RED.ADD) into the row of its group of 8 CTAs and into
// the digit totals; after ONE grid barrier a CTA sums the rows of the groups before its own (<= 36) and of
// the CTAs before it inside its group (<= 7) - 2 grid barriers per pass instead of 3 (13 -> 8 in total).
//
// COMPACT mode (band binning): culled Gaussians (key 0xFFFFFFFF) are dropped in pass 0, so passes 1-3, the
// rect gather and everything downstream work on the V valid Gaussians only (V stays on the device, *nvalid):
// 5 M-Gaussian maps have V ~ 1 M per view.
constexpr uint32_t kCoopSortMaxBlocks = 296;  // 2 CTAs per SM
constexpr uint32_t kSortGroup = 8;            // CTAs per group row
constexpr uint32_t kSortMaxGroups = (kCoopSortMaxBlocks + kSortGroup - 1) / kSortGroup;   // 37
// u32 words of accumulation scratch per pass: group rows + digit totals; zeroed by the host before the launch
constexpr size_t kSortAccumWords = (size_t)(kSortMaxGroups + 1) * 256;

struct DepthSortArgs {
  const uint32_t* dkey;   // [n] keys (read only)
  uint32_t* keyA;         // ping
  uint32_t* valA;
  uint32_t* keyB;         // pong (keys of the last pass are never written)
  uint32_t* valB;         // final argsort
  uint32_t* hist;         // [2][nblk][256] double-buffered row matrix
  uint32_t* accum;        // [4][kSortAccumWords] zero-initialised group rows + totals, one set per pass
  uint32_t* bsum;         // [nblk] tail scan block sums
  const uint32_t* wsrc;   // [n] per-item weight, gathered through the argsort (NULL: skip the scan)
  const uint2* rect;      // [n] per-item rect, gathered through the argsort
  uint2* srect;           // [n] out
  uint32_t* offsets;      // [n] out: exclusive scan of wsrc in sorted order
  uint32_t* total_out;    // nullable: grand total of wsrc
  uint32_t* nvalid;       // out: number of sorted items (n, or the number of valid keys in compact mode)
  uint32_t n, nblk;
  int compact;            // drop keys == 0xFFFFFFFF in pass 0
};

inline size_t depth_sort_scratch_elems(uint64_t n) {
  const size_t nblk = sort_num_blocks(n);
  const size_t coop = 2 * nblk * 256 + 4 * kSortAccumWords + nblk + 64;
  const size_t plain = sort_scratch_elems(n) + 64;
  return coop > plain ? coop : plain;
}

__global__ void __launch_bounds__(kSortThreads) depth_sort_coop_kernel(const DepthSortArgs a) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ uint32_t wh[kSortWarps][256];
  __shared__ uint32_t sk[kSortChunk], sv[kSortChunk];   // exchange buffers (key, value) in CTA-local sorted order
  __shared__ uint32_t gdelta[256];                      // digit -> global position minus CTA-local slot
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t blk = blockIdx.x, nblk = a.nblk;
  const uint32_t base = blk * kSortChunk + warp * kSortWarpItems;
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t grp = blk / kSortGroup;
  uint32_t cta_items = 0;
  uint32_t n = a.n;   // items of the current pass (pass 0: all; later: the valid ones in compact mode)
#pragma unroll 1
  for (int pass = 0; pass < 4; ++pass) {
    const int shift = pass * 8;
    const uint32_t* kin = pass == 0 ? a.dkey : ((pass & 1) ? a.keyA : a.keyB);
    const uint32_t* vin = (pass & 1) ? a.valA : a.valB;
    uint32_t* kout = (pass & 1) ? a.keyB : a.keyA;
    uint32_t* vout = (pass & 1) ? a.valB : a.valA;
    uint32_t* hist = a.hist + (size_t)(pass & 1) * nblk * 256;
    uint32_t* gsum = a.accum + (size_t)pass * kSortAccumWords;         // [groups][256]
    uint32_t* dtot = gsum + (size_t)kSortMaxGroups * 256;               // [256]
#pragma unroll
    for (int w = 0; w < kSortWarps; ++w) wh[w][threadIdx.x] = 0;
    __syncthreads();
    uint32_t k[kSortIpt], v[kSortIpt];
    const bool drop = a.compact && pass == 0;
#pragma unroll
    for (int it = 0; it < kSortIpt; ++it) {
      const uint32_t i = base + it * 32 + lane;
      // .cg: pass >= 1 reads what other CTAs wrote earlier in this launch
      k[it] = i < n ? __ldcg(kin + i) : 0xffffffffu;
      v[it] = i < n ? (pass == 0 ? i : __ldcg(vin + i)) : 0u;
    }
    // counting: plain shared-memory atomics (no ranks needed here).  MATCH.ANY costs ~450 cycles per call on sm_100
    // (one internal round per distinct digit in the warp; ncu source view of the previous version: the instructions
    // consuming its result held 40 % of the stall samples), so it is used once per item per pass, for the ranks
    // only, and all 16 calls are issued back to back there.
    uint32_t valid = 0;   // bit `it`: item takes part in this pass
#pragma unroll
    for (int it = 0; it < kSortIpt; ++it) {
      const uint32_t i = base + it * 32 + lane;
      const bool ok = i < n && !(drop && k[it] == 0xffffffffu);
      valid |= ok ? (1u << it) : 0u;
      if (ok) atomicAdd(&wh[warp][(k[it] >> shift) & 255u], 1u);
    }
    __syncthreads();
    {
      uint32_t c = 0;
#pragma unroll
      for (int w = 0; w < kSortWarps; ++w) c += wh[w][threadIdx.x];
      hist[(size_t)blk * 256 + threadIdx.x] = c;
      if (c) {
        atomicAdd(gsum + (size_t)grp * 256 + threadIdx.x, c);
        atomicAdd(dtot + threadIdx.x, c);
      }
    }
    grid.sync();
    {
      // digit = threadIdx.x: items of this digit in earlier blocks (group rows before mine + rows of the CTAs
      // before me in my group), and in all blocks
      uint32_t below = 0;
      for (uint32_t g = 0; g < grp; ++g) below += __ldcg(gsum + (size_t)g * 256 + threadIdx.x);
      for (uint32_t b = grp * kSortGroup; b < blk; ++b) below += __ldcg(hist + (size_t)b * 256 + threadIdx.x);
      const uint32_t all = __ldcg(dtot + threadIdx.x);
      uint32_t tot;
      const uint32_t gstart = block_excl_scan_256(all, &tot) + below;   // global start of this CTA's items of the digit
      if (pass == 0) {
        n = tot;   // number of valid items (== a.n unless compact)
        if (blk == 0 && threadIdx.x == 0 && a.nvalid) *a.nvalid = tot;
      }
      uint32_t mine = 0;
#pragma unroll
      for (int w = 0; w < kSortWarps; ++w) mine += wh[w][threadIdx.x];
      uint32_t run = block_excl_scan_256(mine, &cta_items);             // CTA-local start of the digit
      gdelta[threadIdx.x] = gstart - run;
#pragma unroll
      for (int w = 0; w < kSortWarps; ++w) {
        const uint32_t c = wh[w][threadIdx.x];
        wh[w][threadIdx.x] = run;
        run += c;
      }
    }
    __syncthreads();
    // exchange through shared memory: items are placed at their CTA-local sorted slot first, so that the global
    // stores below go out in runs of one digit (16 consecutive items on average) instead of 4-byte scatters -
    // scattered 4-B stores cost a 32-B sector each and were the floor of the pass
    uint32_t peers[kSortIpt];
#pragma unroll
    for (int it = 0; it < kSortIpt; ++it)   // 16 independent MATCH.ANY in flight
      peers[it] = __match_any_sync(0xffffffffu, ((valid >> it) & 1u) ? ((k[it] >> shift) & 255u) : 0xffffffffu);
#pragma unroll
    for (int it = 0; it < kSortIpt; ++it) {
      const bool ok = (valid >> it) & 1u;
      const uint32_t d = (k[it] >> shift) & 255u;
      if (ok) {
        const uint32_t slot = wh[warp][d] + __popc(peers[it] & lt);
        sk[slot] = k[it];
        sv[slot] = v[it];
      }
      __syncwarp();
      if (ok && lane == (__ffs(peers[it]) - 1)) wh[warp][d] += __popc(peers[it]);
      __syncwarp();
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < kSortIpt; ++q) {
      const uint32_t slot = q * kSortThreads + threadIdx.x;
      if (slot < cta_items) {
        const uint32_t key = sk[slot];
        const uint32_t pos = slot + gdelta[(key >> shift) & 255u];
        if (pass < 3) kout[pos] = key;
        vout[pos] = sv[slot];
      }
    }
    grid.sync();
  }
  // tail: srect[i] = rect[sorted[i]]; offsets = exclusive scan of wsrc[sorted[i]]
  if (!a.wsrc) {   // band binning needs the depth-ordered rects only
#pragma unroll
    for (int q = 0; q < kSortIpt; ++q) {
      const uint32_t i = blk * kSortChunk + q * kSortThreads + threadIdx.x;
      if (i < n) a.srect[i] = a.rect[__ldcg(a.valB + i)];
    }
    return;
  }
  const uint32_t tbase = blk * kSortChunk + threadIdx.x * kSortIpt;
  uint32_t pre[kSortIpt];
  uint32_t s = 0;
#pragma unroll
  for (int q = 0; q < kSortIpt; ++q) {
    const uint32_t i = tbase + q;
    uint32_t x = 0;
    if (i < n) {
      const uint32_t id = __ldcg(a.valB + i);
      a.srect[i] = a.rect[id];
      x = a.wsrc[id];
    }
    pre[q] = s;
    s += x;
  }
  uint32_t mine;
  const uint32_t ex = block_excl_scan_256(s, &mine);
  if (threadIdx.x == 0) a.bsum[blk] = mine;
  grid.sync();
  uint32_t part = 0;
  for (uint32_t b = threadIdx.x; b < blk; b += kSortThreads) part += __ldcg(a.bsum + b);
  uint32_t before;
  block_excl_scan_256(part, &before);
#pragma unroll
  for (int q = 0; q < kSortIpt; ++q) {
    const uint32_t i = tbase + q;
    if (i < n) a.offsets[i] = pre[q] + ex + before;
  }
  if (a.total_out && blk == nblk - 1 && threadIdx.x == 0) *a.total_out = before + mine;
}

// True when the cooperative kernel can hold all `nblk` CTAs resident on the current device.
inline bool depth_sort_coop_fits(uint32_t nblk) {
  static int capacity = -1;  // per process; every rank drives one device
  if (capacity < 0) {
    int dev = 0, sms = 0, per_sm = 0, coop = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, depth_sort_coop_kernel, kSortThreads, 0);
    capacity = coop ? sms * per_sm : 0;
  }
  return nblk > 0 && nblk <= kCoopSortMaxBlocks && (int)nblk <= capacity;
}

inline cudaError_t depth_sort_coop(DepthSortArgs a, cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(a.accum, 0, 4 * kSortAccumWords * sizeof(uint32_t), st);
  if (e != cudaSuccess) return e;
  void* params[] = {(void*)&a};
  return cudaLaunchCooperativeKernel((const void*)depth_sort_coop_kernel, dim3(a.nblk), dim3(kSortThreads), params, 0, st);
}

}  // namespace pg
