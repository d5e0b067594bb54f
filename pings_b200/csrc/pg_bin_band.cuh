// Band binning: builds the per-tile depth-ordered Gaussian lists (the reference's duplicateWithKeys +
// 64-bit SortPairs + identifyTileRanges, R3 cuda_rasterizer/rasterizer_impl.cu:70-138,295-330) in two
// counting-sort levels whose stores are coalesced.
//
// Why two levels: a one-level scatter writes 4 bytes per (Gaussian, tile) instance to R random
// sectors; tools/micro/scatter_store.cu measures 170 us for 22.7 M such stores on B200 against 26 us
// when 8 lanes share a sector.  So
//   A. every depth-ordered Gaussian emits one 4-byte record (its id) into each BAND it touches - a
//      band is 32 horizontally adjacent tiles of one tile row, lists are [row][band] - with a chunked
//      counting sort whose writes are warp-compacted runs (kernels band_rows_kernel<false/true>);
//   B. a warp expands 128 consecutive records of one band: lane t owns tile t of the band, walks the
//      records in order (so its list stays depth-ordered without atomics or ranking), appends the ids
//      it is covered by to a shared-memory stage and the warp flushes the 32 runs with coalesced
//      stores (band_expand_kernel<false/true>; count pass first, one scan of the small
//      [band][tile][segment] matrix in between).
// All sizes that depend on the data (records per band, segments) live on the device; the grids of
// stage B are persistent and stride over the device-side segment count.
#pragma once
#include "pg_sort.cuh"

namespace pg {

constexpr int kBandTiles = 32;        // tiles per band == lanes of the expanding warp
constexpr int kBandMaxLists = 1024;   // [row][band] lists; more -> older paths
constexpr int kBandSegRecs = 128;     // records per warp segment in stage B
constexpr int kBandStageCap = 2048;   // staged instances per warp (8 KB)
constexpr int kBandWarps = 8;
constexpr int kBandStageWords = kBandStageCap + 32 + kBandStageCap / 4;   // ids + deltas + tile bytes

inline int band_count(int gx) { return (gx + kBandTiles - 1) / kBandTiles; }
inline int band_lists(int gx, int gy) { return band_count(gx) * gy; }
inline bool band_path(int gx, int gy) { return band_lists(gx, gy) <= kBandMaxLists; }
// stage-A chunks (warps): a multiple of 8, matrix [lists][chunks] kept <= 1 Mi entries
inline int band_row_chunks(int nL) {
  int c = ((1 << 20) / (nL > 0 ? nL : 1)) & ~7;
  return c > 4096 ? 4096 : (c < 256 ? 256 : c);
}
// upper bounds used to size the workspace from num_rendered alone (records <= instances)
inline size_t band_seg_bound(int64_t R, int nL) { return (size_t)(R > 0 ? R : 0) / kBandSegRecs + (size_t)nL; }
inline size_t band_tilemat_bound(int64_t R, int nL) { return band_seg_bound(R, nL) * kBandTiles + 1; }

struct BandWs {
  uint32_t* recs;      // [<= R] Gaussian ids, list-major, depth-ordered inside a list
  uint32_t* matA;      // [nL * chunks + 1] stage-A counts -> record offsets
  uint32_t* lstart;    // [nL + 1] first record of list l (lstart[nL] = number of records)
  uint32_t* segbase;   // [nL + 1] first segment of list l (segbase[nL] = number of segments)
  uint32_t* scan_len;  // [1] 32 * segments + 1
  uint32_t* tilemat;   // [32 * segments + 1] stage-B counts -> output offsets; index 32*segbase[l] + x*nseg_l + s
  uint32_t* mscan;     // scan scratch
};

// ---- stage A ----------------------------------------------------------------------------------
// One warp per chunk of `per_warp` consecutive depth ranks; `cur` = per-warp counters/cursors of
// all lists in shared memory.  For 32 Gaussians at a time the warp loops over the (row, band)
// cells of their joint bounding box and ballots coverage: a cell's 32-bit mask gives both the count
// and, in the WRITE pass, the stable rank of every covering lane.
template <bool WRITE>
__global__ void __launch_bounds__(kBandWarps * 32) band_rows_kernel(
    int P, int per_warp, int NB, int nL, int nchunks, const uint32_t* __restrict__ sorted_idx,
    const uint2* __restrict__ srect, uint32_t* __restrict__ matA, uint32_t* __restrict__ recs) {
  extern __shared__ uint32_t band_sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunk = blockIdx.x * kBandWarps + warp;
  uint32_t* cur = band_sm + (size_t)warp * nL;
  for (int l = lane; l < nL; l += 32) cur[l] = WRITE ? matA[(size_t)l * nchunks + chunk] : 0u;
  __syncwarp();
  const uint32_t lt = (1u << lane) - 1u;
  const int begin = chunk * per_warp, end = min(P, begin + per_warp);
  for (int i0 = begin; i0 < end; i0 += 32) {
    const int i = i0 + lane;
    uint2 rc = make_uint2(0u, 0u);
    if (i < end) rc = srect[i];
    const uint32_t x0 = rc.x & 0xffffu, x1 = rc.x >> 16, y0 = rc.y & 0xffffu, y1 = rc.y >> 16;
    const bool ne = x1 > x0 && y1 > y0;
    if (!__any_sync(0xffffffffu, ne)) continue;
    uint32_t id = 0;
    if (WRITE && ne) id = sorted_idx[i];
    const uint32_t b0 = x0 / kBandTiles, b1 = ne ? (x1 - 1) / kBandTiles : 0u;
    const uint32_t rmin = __reduce_min_sync(0xffffffffu, ne ? y0 : 0xffffu);
    const uint32_t rmax = __reduce_max_sync(0xffffffffu, ne ? y1 : 0u);
    const uint32_t bmin = __reduce_min_sync(0xffffffffu, ne ? b0 : 0xffffu);
    const uint32_t bmax = __reduce_max_sync(0xffffffffu, ne ? b1 : 0u);
    for (uint32_t r = rmin; r < rmax; ++r) {
      const bool in_row = ne && y0 <= r && r < y1;
      for (uint32_t b = bmin; b <= bmax; ++b) {
        const bool cov = in_row && b0 <= b && b <= b1;
        const uint32_t m = __ballot_sync(0xffffffffu, cov);
        if (m == 0) continue;
        const uint32_t l = r * NB + b;
        if (WRITE) {
          const uint32_t base = cur[l];
          if (cov) recs[base + __popc(m & lt)] = id;
          __syncwarp();
          if (lane == 0) cur[l] = base + __popc(m);
        } else {
          if (lane == 0) cur[l] += __popc(m);
        }
      }
    }
    __syncwarp();
  }
  if (!WRITE) {
    __syncwarp();
    for (int l = lane; l < nL; l += 32) matA[(size_t)l * nchunks + chunk] = cur[l];
  }
}

// ---- header: list starts, segment table (one CTA of 1024 threads, nL <= 1024) ---------------------
__global__ void __launch_bounds__(1024) band_header_kernel(int nL, int nchunks, const uint32_t* __restrict__ matA,
                                                            uint32_t* __restrict__ lstart, uint32_t* __restrict__ segbase,
                                                            uint32_t* __restrict__ scan_len, uint32_t* __restrict__ tilemat) {
  __shared__ uint32_t wsum[32];
  const int l = threadIdx.x, lane = l & 31, warp = l >> 5;
  uint32_t nseg = 0;
  if (l < nL) {
    const uint32_t a = matA[(size_t)l * nchunks], b = matA[(size_t)(l + 1) * nchunks];  // scanned, [nL*nchunks] = total
    lstart[l] = a;
    if (l == nL - 1) lstart[nL] = b;
    nseg = (b - a + kBandSegRecs - 1) / kBandSegRecs;
  }
  const uint32_t inc = warp_incl_scan(nseg, lane);
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = wsum[lane];
    const uint32_t wi = warp_incl_scan(w, lane);
    wsum[lane] = wi - w;
  }
  __syncthreads();
  const uint32_t ex = inc - nseg + wsum[warp];
  if (l < nL) segbase[l] = ex;
  if (l == nL - 1) {
    const uint32_t S = ex + nseg;
    segbase[nL] = S;
    *scan_len = S * kBandTiles + 1;
    tilemat[(size_t)S * kBandTiles] = 0;   // scanned into the grand total (== num_rendered)
  }
}

// Lane t appends the ids of the records (one per lane, coverage mask cm) that cover its tile, in lane
// order, to its run of the stage starting at k; returns the advanced k.
__device__ __forceinline__ uint32_t band_stage_group(uint32_t cm, uint32_t id, uint32_t lane_bit, int lane,
                                                     uint32_t* stage, uint8_t* stage_tile, uint32_t k) {
#pragma unroll
  for (int jj = 0; jj < 32; ++jj) {
    const bool hit = (__shfl_sync(0xffffffffu, cm, jj) & lane_bit) != 0u;
    const uint32_t idj = __shfl_sync(0xffffffffu, id, jj);
    if (hit) {
      stage[k] = idj;
      stage_tile[k] = (uint8_t)lane;
      ++k;
    }
  }
  return k;
}

// Flat flush: the stage is tile-major, so consecutive lanes write consecutive words of a run.
__device__ __forceinline__ void band_flush(const uint32_t* stage, const uint8_t* stage_tile, const uint32_t* stage_delta,
                                           uint32_t tot, int lane, uint32_t* __restrict__ out) {
  __syncwarp();
  for (uint32_t q = lane; q < tot; q += 32) out[q + stage_delta[stage_tile[q]]] = stage[q];
  __syncwarp();
}

// ---- stage B ------------------------------------------------------------------------------------
// Persistent warps stride over the segments.  Lane t owns tile `band*32 + t`.
template <bool WRITE>
__global__ void __launch_bounds__(kBandWarps * 32) band_expand_kernel(
    int nL, int NB, const uint32_t* __restrict__ lstart, const uint32_t* __restrict__ segbase,
    const uint32_t* __restrict__ recs, const uint2* __restrict__ rect, uint32_t* __restrict__ tilemat,
    uint32_t* __restrict__ out) {
  extern __shared__ uint32_t band_sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t* sb = band_sm;                                      // [nL + 1]
  const int sb_words = (nL + 1 + 31) & ~31;
  // WRITE only: per warp [cap] ids, [32] run deltas, [cap] owning tile (bytes)
  uint32_t* stage = band_sm + sb_words + warp * kBandStageWords;
  uint32_t* stage_delta = stage + kBandStageCap;
  uint8_t* stage_tile = reinterpret_cast<uint8_t*>(stage_delta + 32);
  for (int k = threadIdx.x; k <= nL; k += blockDim.x) sb[k] = segbase[k];
  __syncthreads();
  const uint32_t S = sb[nL];
  for (uint32_t g = blockIdx.x * kBandWarps + warp; g < S; g += gridDim.x * kBandWarps) {
    int lo = 0, hi = nL;   // sb[lo] <= g < sb[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (sb[mid] <= g) lo = mid; else hi = mid;
    }
    const uint32_t l = (uint32_t)lo, s = g - sb[lo], nseg = sb[lo + 1] - sb[lo];
    const uint32_t band0 = (l % (uint32_t)NB) * kBandTiles;
    const uint32_t first = lstart[l] + s * kBandSegRecs;
    const uint32_t n = min((uint32_t)kBandSegRecs, lstart[l + 1] - first);
    const uint32_t* rp = recs + first;
    const size_t e = (size_t)sb[lo] * kBandTiles + (size_t)lane * nseg + s;
    uint32_t c = 0, goff = 0, start = 0, k = 0;
    bool staged = false;
    if (WRITE) {
      goff = tilemat[e];
      c = tilemat[e + 1] - goff;
      const uint32_t inc = warp_incl_scan(c, lane);
      start = inc - c;
      staged = __shfl_sync(0xffffffffu, inc, 31) <= (uint32_t)kBandStageCap;
      k = staged ? start : 0u;
    }
    const uint32_t lane_bit = 1u << lane;
    for (uint32_t j0 = 0; j0 < n; j0 += 32) {
      uint32_t id = 0, cm = 0;   // cm = coverage mask of this lane's record over the band's 32 tiles
      if (j0 + lane < n) {
        id = rp[j0 + lane];
        const uint32_t rx = rect[id].x;
        const uint32_t x0 = max(rx & 0xffffu, band0) - band0;
        const uint32_t x1 = min(rx >> 16, band0 + kBandTiles) - band0;   // 1..32
        cm = (0xffffffffu >> (32u - (x1 - x0))) << x0;
      }
      if (!WRITE) {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) c += (__shfl_sync(0xffffffffu, cm, jj) & lane_bit) ? 1u : 0u;
      } else if (staged) {
        k = band_stage_group(cm, id, lane_bit, lane, stage, stage_tile, k);
      } else {
        // dense segment (> cap instances): stage and flush these 32 records on their own (<= 1024 instances)
        uint32_t cg = 0;
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) cg += (__shfl_sync(0xffffffffu, cm, jj) & lane_bit) ? 1u : 0u;
        const uint32_t inc = warp_incl_scan(cg, lane);
        band_stage_group(cm, id, lane_bit, lane, stage, stage_tile, inc - cg);
        stage_delta[lane] = goff + k - (inc - cg);   // k = instances of this tile already written
        band_flush(stage, stage_tile, stage_delta, __shfl_sync(0xffffffffu, inc, 31), lane, out);
        k += cg;
      }
    }
    if (!WRITE) {
      tilemat[e] = c;
    } else if (staged) {
      stage_delta[lane] = goff - start;   // staged position -> output position, for this lane's tile
      band_flush(stage, stage_tile, stage_delta, __shfl_sync(0xffffffffu, start + c, 31), lane, out);
    }
  }
}

// ranges[tile] = [start, end) in the output list; (0, 0) for empty tiles like the reference's memset
__global__ void __launch_bounds__(256) band_ranges_kernel(int T, int gx, int NB, const uint32_t* __restrict__ segbase,
                                                           const uint32_t* __restrict__ tilemat, uint2* __restrict__ ranges) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int r = t / gx, x = t - r * gx;
  const int l = r * NB + x / kBandTiles, xl = x % kBandTiles;
  const uint32_t sb = segbase[l], nseg = segbase[l + 1] - sb;
  uint2 out = make_uint2(0u, 0u);
  if (nseg) {
    const size_t e = (size_t)sb * kBandTiles + (size_t)xl * nseg;
    const uint32_t a = tilemat[e], b = tilemat[e + nseg];
    if (b > a) out = make_uint2(a, b);
  }
  ranges[t] = out;
}

inline cudaError_t band_bin(const BandWs& w, int P, int gx, int gy, int64_t R, const uint32_t* sorted_idx,
                            const uint2* srect, const uint2* rect, uint32_t* out, uint2* ranges, cudaStream_t st) {
  const int NB = band_count(gx), nL = NB * gy, nchunks = band_row_chunks(nL), T = gx * gy;
  const int per_warp = ((P + nchunks - 1) / nchunks + 31) & ~31;
  const size_t smA = (size_t)kBandWarps * nL * sizeof(uint32_t);
  const int sb_words = (nL + 1 + 31) & ~31;
  const size_t smC = (size_t)sb_words * sizeof(uint32_t);
  const size_t smW = smC + (size_t)kBandWarps * kBandStageWords * sizeof(uint32_t);
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(band_expand_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  int sms = 148;
  {
    static int cached = 0;
    if (!cached) {
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    }
    sms = cached > 0 ? cached : 148;
  }
  const uint32_t nA = (uint32_t)((size_t)nL * nchunks);
  cudaError_t e = cudaMemsetAsync(w.matA + nA, 0, sizeof(uint32_t), st);
  if (e != cudaSuccess) return e;
  band_rows_kernel<false><<<nchunks / kBandWarps, kBandWarps * 32, smA, st>>>(P, per_warp, NB, nL, nchunks, sorted_idx,
                                                                              srect, w.matA, nullptr);
  e = exclusive_scan_u32(w.matA, nullptr, nA + 1, w.matA, w.mscan, nullptr, st);
  if (e != cudaSuccess) return e;
  band_header_kernel<<<1, 1024, 0, st>>>(nL, nchunks, w.matA, w.lstart, w.segbase, w.scan_len, w.tilemat);
  band_rows_kernel<true><<<nchunks / kBandWarps, kBandWarps * 32, smA, st>>>(P, per_warp, NB, nL, nchunks, sorted_idx,
                                                                             srect, w.matA, w.recs);
  band_expand_kernel<false><<<sms * 8, kBandWarps * 32, smC, st>>>(nL, NB, w.lstart, w.segbase, w.recs, rect,
                                                                    w.tilemat, nullptr);
  e = exclusive_scan_u32_dyn(w.tilemat, (uint32_t)band_tilemat_bound(R, nL), w.scan_len, w.mscan, st);
  if (e != cudaSuccess) return e;
  band_ranges_kernel<<<(T + 255) / 256, 256, 0, st>>>(T, gx, NB, w.segbase, w.tilemat, ranges);
  band_expand_kernel<true><<<sms * 2, kBandWarps * 32, smW, st>>>(nL, NB, w.lstart, w.segbase, w.recs, rect,
                                                                   w.tilemat, out);
  return cudaGetLastError();
}

}  // namespace pg
