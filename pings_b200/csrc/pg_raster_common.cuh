// Layout of the rasterizer workspaces + kernels shared by the 3DGS and surfel variants:
// instance emission, tile ranges, depth ordering.
//
// Pipeline (differs from the reference's preprocess -> scan -> duplicate -> 64-bit sort,
// R3 cuda_rasterizer/rasterizer_impl.cu:198-344, but yields bit-identical point_list/ranges):
//   1. preprocess writes one packed record per Gaussian + a u32 depth key + its tile rect
//   2. Gaussians are argsorted by depth key (4 LSD passes over P items, not R instances):
//      the low 32 bits of the reference's 64-bit key are identical for all instances of a
//      Gaussian, so those LSD passes commute with instance duplication
//   3. (tiles <= kBinMaxTiles) chunked counting sort keyed by tile with IMPLICIT keys: the
//      depth-ordered Gaussians are cut into <= NC chunks, one warp per chunk walks its
//      Gaussians in order and bumps a per-tile counter in shared memory for every tile of the
//      rect (count pass), the [tile][chunk] count matrix is scanned, and the same walk writes
//      the ids at off[tile]++ (scatter pass).  No key array, no pass over R items other than
//      the final 4-byte id writes; tile ranges fall out of the scanned matrix.
//      (larger grids) instances are emitted in depth order with a 16-bit tile key, then stably
//      sorted on the tile bits only (1-2 passes over R items of 6 bytes)
#pragma once
#include "pg_common.cuh"
#include <cstdlib>
#include "pg_sort.cuh"
#include "pg_bucket_sort.cuh"
#include "pg_bin_band.cuh"

namespace pg {

constexpr int kTile = 16;
constexpr uint32_t kBigRect = 24;  // rects with >= this many tiles are walked cooperatively when binning

// Cost model of the binning walk (bin_walk_kernel): a packed instance of a small rect costs about 8x
// a cooperatively walked one, plus a fixed per-Gaussian cost.  Chunks are balanced on this weight.
__host__ __device__ __forceinline__ uint32_t bin_weight(uint32_t tiles) {
  if (tiles == 0) return 0u;
  return tiles < kBigRect ? 8u * tiles + 16u : tiles + 32u;
}

// Tail of the two preprocess kernels: block-reduced instance count and the key range of the visible Gaussians
// (max(key), max(~key) so that a zero fill initialises both) for the bucket sort; 3 atomics per CTA.
// total[0] = instances (u64), total[1] viewed as u32[2] = {max key, max ~key}.
__device__ __forceinline__ void preprocess_block_tail(uint32_t my_tiles, uint32_t key, unsigned long long* total) {
  __shared__ uint32_t red[3][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool valid = my_tiles != 0;
  uint32_t s = __reduce_add_sync(0xffffffffu, my_tiles);
  uint32_t hi = __reduce_max_sync(0xffffffffu, valid ? key : 0u);
  uint32_t lo = __reduce_max_sync(0xffffffffu, valid ? ~key : 0u);
  if (lane == 0) { red[0][warp] = s; red[1][warp] = hi; red[2][warp] = lo; }
  __syncthreads();
  if (warp == 0) {
    const int nw = (int)(blockDim.x >> 5);
    s = __reduce_add_sync(0xffffffffu, lane < nw ? red[0][lane] : 0u);
    hi = __reduce_max_sync(0xffffffffu, lane < nw ? red[1][lane] : 0u);
    lo = __reduce_max_sync(0xffffffffu, lane < nw ? red[2][lane] : 0u);
    if (lane == 0 && s) {
      atomicAdd(total, (unsigned long long)s);
      uint32_t* mm = reinterpret_cast<uint32_t*>(total + 1);
      atomicMax(mm, hi);
      atomicMax(mm + 1, lo);
    }
  }
}

struct GeomWs {
  float4* rec;            // [3P] packed per-Gaussian record (see pg_raster3.cu)
  float4* ext;            // [P] surfel: (n_view.xyz, zx) or null (pg_rasters_fwd.cuh)
  float* zyv;             // [P] surfel: zy or null
  uint32_t* dkey;         // [P] depth bits, 0xFFFFFFFF when culled
  uint32_t* tiles;        // [P] tiles touched
  uint32_t* wts;          // [P] binning cost weight of the Gaussian (bin_weight(tiles))
  uint32_t* wtotal;       // [1] sum of wts (device scalar written by the scan)
  uint32_t* nvalid;       // [1] number of depth-sorted Gaussians: the visible ones on the band path (culled dropped), else P
  uint2* rect;            // [P] x0 | x1<<16 , y0 | y1<<16 (0,0 when culled)
  uint2* srect;           // [P] rect in depth order (rect[sorted_idx[i]])
  uint32_t* keyA;         // [P] sort ping
  uint32_t* keyB;         // [P] sort pong
  uint32_t* valA;         // [P]
  uint32_t* sorted_idx;   // [P] Gaussian ids in depth order (culled ones last)
  uint32_t* offsets;      // [P] exclusive scan of tiles in depth order
  uint8_t* clamped;       // [3P] SH clamp flags
  uint32_t* sort_scratch; // radix/scan scratch for P items
  unsigned long long* total;  // [2] number of instances; [1] = {max key, max ~key} of the visible Gaussians (u32 x 2)
  double* tau_acc;        // [8] pose-gradient accumulators (6 used) + block counter
  float4* gacc;           // [3P] (3DGS) / [4P] (surfel) backward gradient accumulation records
  size_t gacc_count;      // number of float4 in gacc
};

inline size_t carve_geom(int variant, int P, void* base, GeomWs* ws) {
  Carver c(base);
  GeomWs w;
  size_t p = (size_t)(P > 0 ? P : 1);
  w.rec = c.take<float4>(3 * p);
  w.ext = variant == PG_VARIANT_SURFEL ? c.take<float4>(p) : nullptr;
  w.zyv = variant == PG_VARIANT_SURFEL ? c.take<float>(p) : nullptr;
  w.dkey = c.take<uint32_t>(p);
  w.tiles = c.take<uint32_t>(p);
  w.wts = c.take<uint32_t>(p);
  w.wtotal = c.take<uint32_t>(4);
  w.nvalid = w.wtotal ? w.wtotal + 1 : nullptr;
  w.rect = c.take<uint2>(p);
  w.srect = c.take<uint2>(p);
  w.keyA = c.take<uint32_t>(p);
  w.keyB = c.take<uint32_t>(p);
  w.valA = c.take<uint32_t>(p);
  w.sorted_idx = c.take<uint32_t>(p);
  w.offsets = c.take<uint32_t>(p);
  w.clamped = c.take<uint8_t>(3 * p);
  // `total` (one 128-B line) sits immediately before the sort scratch: on the band path ONE memset clears the
  // instance count, the key range and the bucket sort's bins / look-back words before the preprocess kernel
  w.total = c.take<unsigned long long>(16);
  {
    const size_t s1 = depth_sort_scratch_elems(p), s2 = bucket_sort_scratch_elems(p);
    w.sort_scratch = c.take<uint32_t>(s1 > s2 ? s1 : s2);
  }
  w.tau_acc = c.take<double>(8);
  w.gacc_count = (variant == PG_VARIANT_SURFEL ? 4 : 3) * p;
  w.gacc = c.take<float4>(w.gacc_count);
  if (ws) *ws = w;
  return c.total();
}

struct ImageWs {
  float* final_T;       // [W*H]
  uint32_t* n_contrib;  // [W*H]
  uint2* ranges;        // [tiles]
  float* final_D;       // [W*H] surfel only
};

inline size_t carve_image(int variant, int W, int H, void* base, ImageWs* ws) {
  Carver c(base);
  ImageWs w;
  size_t n = (size_t)W * H;
  size_t tiles = (size_t)div_up(W, kTile) * div_up(H, kTile);
  w.final_T = c.take<float>(n);
  w.n_contrib = c.take<uint32_t>(n);
  w.ranges = c.take<uint2>(tiles);
  w.final_D = variant == PG_VARIANT_SURFEL ? c.take<float>(n) : nullptr;
  if (ws) *ws = w;
  return c.total();
}

// ---- chunked counting sort geometry ---------------------------------------------------------
constexpr int kBinMaxTiles = 12288;          // 48 KB of u32 counters per warp
constexpr int kBinMatrixBudget = 4 << 20;    // max entries of the [tile][chunk] matrix (16 MB)
inline bool bin_fast_path(int num_tiles) { return num_tiles <= kBinMaxTiles; }
// Upper bound of the number of chunks for a tile grid (independent of P so that the workspace
// size only depends on the image size).
inline int bin_max_chunks(int num_tiles) {
  int nc = kBinMatrixBudget / (num_tiles > 0 ? num_tiles : 1);
  return nc > 4096 ? 4096 : (nc < 148 ? 148 : nc);
}
// Chunks are balanced by WEIGHT (near Gaussians cover hundreds of tiles, far ones a few, and small
// rects cost more per instance): chunk c owns the depth-ordered Gaussians whose exclusive weight
// prefix lies in [c*S, (c+1)*S), S = max(kBinMinChunk, ceil(wtotal / NC)) computed on the device.
constexpr uint32_t kBinMinChunk = 2048;
inline int bin_warps_per_cta(int num_tiles) {
  int w = (160 * 1024) / (num_tiles * 4);
  return w > 8 ? 8 : (w < 1 ? 1 : w);
}

struct BinWs {
  uint16_t* tkeyA;   // [R] tile keys ping          (radix path only)
  uint16_t* tkeyB;   // [R] pong                    (radix path only)
  uint32_t* valA;    // [R] sorted ids (counting path) / ping
  uint32_t* valB;    // [R]                         (radix path only)
  uint32_t* scratch; // radix scratch for R items   (radix path only)
  uint32_t* matrix;  // [tiles * chunks + 1] count -> offset matrix, tile-major (counting path)
  uint32_t* mscan;   // scan scratch for the matrix
  BandWs band;       // band binning (pg_bin_band.cuh), the default path
};

inline size_t carve_binning(int64_t R, int W, int H, void* base, BinWs* ws) {
  Carver c(base);
  BinWs w;
  memset(&w, 0, sizeof(w));
  size_t r = (size_t)(R > 0 ? R : 1);
  const int gx = div_up(W, kTile), gy = div_up(H, kTile);
  const int T = gx * gy;
  if (band_path(gx, gy)) {
    const int nL = band_lists(gx, gy);
    w.valA = c.take<uint32_t>(r);
    w.band.recs = c.take<uint32_t>(r);
    w.band.matA = c.take<uint32_t>((size_t)nL * band_row_chunks(nL));
    w.band.lcount = c.take<uint32_t>(nL);
    w.band.lstart = c.take<uint32_t>(nL + 1);
    w.band.segbase = c.take<uint32_t>(nL + 1);
    w.band.tilemat = c.take<uint32_t>(band_tilemat_bound(R, nL));
    w.band.tiletot = c.take<uint32_t>((size_t)nL * kBandTiles);
    w.band.tilestart = c.take<uint32_t>((size_t)nL * kBandTiles);
    w.band.ticket = c.take<uint32_t>(4);
  } else if (bin_fast_path(T)) {
    w.valA = c.take<uint32_t>(r);
    const size_t m = (size_t)T * bin_max_chunks(T) + 1;
    w.matrix = c.take<uint32_t>(m);
    w.mscan = c.take<uint32_t>(scan_scratch_elems(m) + 64);
  } else {
    w.tkeyA = c.take<uint16_t>(r);
    w.tkeyB = c.take<uint16_t>(r);
    w.valA = c.take<uint32_t>(r);
    w.valB = c.take<uint32_t>(r);
    w.scratch = c.take<uint32_t>(sort_scratch_elems(r));
  }
  if (ws) *ws = w;
  return c.total();
}

// Number of tile-id bits and how they split into <= 8-bit passes.
inline int tile_key_bits(int num_tiles) {
  int b = 1;
  while ((1 << b) < num_tiles) ++b;
  return b;
}
// After the tile sort the sorted list lives in valA when the pass count is even, valB when odd.
inline int tile_sort_passes(int num_tiles) { return tile_key_bits(num_tiles) > 8 ? 2 : 1; }

// ---- instance emission (duplicateWithKeys, R3 rasterizer_impl.cu:70-111) -------------------
// One warp walks 32 depth-ordered Gaussians; each non-empty one is emitted cooperatively so
// that writes of a rect are contiguous.
__global__ void __launch_bounds__(256) emit_instances_kernel(
    int P, const uint32_t* __restrict__ sorted_idx, const uint32_t* __restrict__ offsets,
    const uint32_t* __restrict__ tiles, const uint2* __restrict__ rect, int grid_x,
    uint16_t* __restrict__ tkeys, uint32_t* __restrict__ vals) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t id = 0, cnt = 0, off = 0;
  uint2 rc = make_uint2(0, 0);
  if (i < P) {
    id = sorted_idx[i];
    cnt = tiles[id];
    if (cnt) {
      off = offsets[i];
      rc = rect[id];
    }
  }
  uint32_t m = __ballot_sync(0xffffffffu, cnt != 0);
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const uint32_t s_id = __shfl_sync(0xffffffffu, id, src);
    const uint32_t s_cnt = __shfl_sync(0xffffffffu, cnt, src);
    const uint32_t s_off = __shfl_sync(0xffffffffu, off, src);
    const uint32_t s_rx = __shfl_sync(0xffffffffu, rc.x, src);
    const uint32_t s_ry = __shfl_sync(0xffffffffu, rc.y, src);
    const uint32_t x0 = s_rx & 0xffffu, x1 = s_rx >> 16;
    const uint32_t y0 = s_ry & 0xffffu;
    const uint32_t w = x1 - x0;
    for (uint32_t t = lane; t < s_cnt; t += 32) {
      const uint32_t yy = t / w, xx = t - yy * w;
      tkeys[s_off + t] = (uint16_t)((y0 + yy) * grid_x + x0 + xx);
      vals[s_off + t] = s_id;
    }
  }
}

// ---- tile ranges (identifyTileRanges, R3 rasterizer_impl.cu:116-138) ------------------------
__global__ void __launch_bounds__(256) tile_ranges_kernel(uint32_t L, const uint16_t* __restrict__ tkeys,
                                                          uint2* __restrict__ ranges) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const uint32_t cur = tkeys[i];
  if (i == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = tkeys[i - 1];
    if (cur != prev) {
      ranges[prev].y = i;
      ranges[cur].x = i;
    }
  }
  if (i == L - 1) ranges[cur].y = L;
}

// ---- chunked counting sort by tile (implicit keys) ------------------------------------------
// One warp per chunk of `per_chunk` depth-ordered Gaussians.  WRITE = false: cnt[tile]++ for every
// tile of every rect, then matrix[tile * nchunks + chunk] = cnt[tile].  WRITE = true: off[tile] is
// loaded from the scanned matrix and every (Gaussian, tile) stores its id at off[tile]++.
// Gaussians are visited strictly in order and the tiles of one rect are distinct, so plain
// shared-memory read-modify-writes keep the per-tile lists depth-ordered (stable) without atomics.
// First i in [0, n) with a[i] >= target (n if none); warp-cooperative 32-ary search.
__device__ __forceinline__ int warp_lower_bound(const uint32_t* __restrict__ a, int n, uint32_t target, int lane) {
  int lo = 0, hi = n;
  while (hi > lo) {
    const int step = (hi - lo + 31) / 32;
    const int idx = lo + lane * step;
    const bool pred = idx < hi ? (a[idx] >= target) : true;
    const uint32_t b = __ballot_sync(0xffffffffu, pred);
    if (b == 0) {
      lo = lo + 31 * step + 1;
    } else {
      const int f = __ffs(b) - 1;
      if (f == 0) {
        hi = lo;
      } else {
        const int nhi = min(hi, lo + f * step);
        lo = lo + (f - 1) * step + 1;
        hi = nhi;
      }
    }
  }
  return lo;
}

template <bool WRITE>
__global__ void __launch_bounds__(256) bin_walk_kernel(
    int P, int nchunks, int num_tiles, int grid_x, const uint32_t* __restrict__ wtotal_ptr,
    const uint32_t* __restrict__ sorted_idx, const uint32_t* __restrict__ offsets,
    const uint2* __restrict__ srect, uint32_t* __restrict__ matrix, uint32_t* __restrict__ out) {
  extern __shared__ uint32_t bin_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunk = blockIdx.x * (blockDim.x >> 5) + warp;
  if (chunk >= nchunks) return;
  uint32_t* ctr = bin_smem + (size_t)warp * num_tiles;
  for (int t = lane; t < num_tiles; t += 32) ctr[t] = WRITE ? matrix[(size_t)t * nchunks + chunk] : 0u;
  __syncwarp();
  const uint32_t wtotal = *wtotal_ptr;
  const uint32_t per_chunk = max(kBinMinChunk, (wtotal + (uint32_t)nchunks - 1u) / (uint32_t)nchunks);
  // (Gaussians at or after the first prefix == wtotal are culled: nothing to walk)
  const uint64_t lo_w = (uint64_t)chunk * per_chunk, hi_w = lo_w + per_chunk;
  const int begin = lo_w >= wtotal ? P : warp_lower_bound(offsets, P, (uint32_t)lo_w, lane);
  const int end = lo_w >= wtotal ? P : warp_lower_bound(offsets, P, (uint32_t)min(hi_w, (uint64_t)wtotal), lane);
  const uint32_t lt = (1u << lane) - 1u;
  // software pipeline: the (id -> tiles, rect) gathers of group g+1 are in flight while group g is walked
  uint32_t n_id = 0;
  uint2 n_rc = make_uint2(0, 0);
  if (begin + lane < end) {
    n_id = sorted_idx[begin + lane];
    n_rc = srect[begin + lane];
  }
  for (int i0 = begin; i0 < end; i0 += 32) {
    const uint32_t id = n_id;
    const uint2 rc = n_rc;
    n_id = 0;
    n_rc = make_uint2(0, 0);
    if (i0 + 32 + lane < end) {
      n_id = sorted_idx[i0 + 32 + lane];
      n_rc = srect[i0 + 32 + lane];
    }
    // the instances of the group's 32 Gaussians form one flattened sequence (Gaussian order, then
    // tile order); lanes take 32 consecutive instances per batch
    const uint32_t x0 = rc.x & 0xffffu, w = (rc.x >> 16) - x0, y0 = rc.y & 0xffffu;
    const uint32_t cnt = w * ((rc.y >> 16) - y0);
    const uint32_t incl = warp_incl_scan(cnt, lane);
    const uint32_t pre = incl - cnt;
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    const float invw = w ? 1.0f / (float)w : 0.f;
    const uint32_t tile00 = y0 * grid_x + x0;
    // Gaussians covering >= kBigRect tiles are walked cooperatively (their tiles are distinct: no
    // ranking needed); the runs of smaller ones in between are packed 32 instances per batch.
    const uint32_t bigmask = __ballot_sync(0xffffffffu, cnt >= kBigRect);
    int j0 = 0;
    while (j0 < 32) {
      const uint32_t rest = j0 ? (bigmask & ~((1u << j0) - 1u)) : bigmask;
      const int jb = rest ? (__ffs(rest) - 1) : 32;
      // ---- packed run [j0, jb) ----
      const uint32_t run_lo = __shfl_sync(0xffffffffu, pre, j0);
      const uint32_t run_hi = jb < 32 ? __shfl_sync(0xffffffffu, pre, jb & 31) : total;
      for (uint32_t s0 = run_lo; s0 < run_hi; s0 += 32) {
        const uint32_t q = s0 + lane;
        const bool valid = q < run_hi;
        // j = last lane with pre_j <= q (5-step search over the prefix values held across lanes);
        // zero-count Gaussians share pre with their successor and are skipped by taking the last one
        int j = 0;
#pragma unroll
        for (int b = 16; b > 0; b >>= 1) {
          const uint32_t pj = __shfl_sync(0xffffffffu, pre, j + b);
          if (pj <= q) j += b;
        }
        const uint32_t g_pre = __shfl_sync(0xffffffffu, pre, j);
        const uint32_t g_id = __shfl_sync(0xffffffffu, id, j);
        const uint32_t g_w = __shfl_sync(0xffffffffu, w, j);
        const float g_invw = __shfl_sync(0xffffffffu, invw, j);
        const uint32_t g_t0 = __shfl_sync(0xffffffffu, tile00, j);
        const uint32_t t = q - g_pre;
        const uint32_t yy = (uint32_t)(((float)t + 0.5f) * g_invw);  // exact: see DESIGN.md (t < 2^22)
        const uint32_t xx = t - yy * g_w;
        const uint32_t tile = valid ? g_t0 + yy * grid_x + xx : 0xffffffffu;
        // instances of one batch that hit the same tile (different Gaussians) are ranked in lane order
        const uint32_t peers = __match_any_sync(0xffffffffu, tile);
        const int leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (valid && lane == leader) {
          base = ctr[tile];
          ctr[tile] = base + __popc(peers);
        }
        base = __shfl_sync(0xffffffffu, base, leader);
        if (WRITE && valid) out[base + __popc(peers & lt)] = g_id;
        __syncwarp();
      }
      if (jb >= 32) break;
      // ---- cooperative walk of Gaussian jb ----
      {
        const uint32_t s_id = __shfl_sync(0xffffffffu, id, jb);
        const uint32_t s_cnt = __shfl_sync(0xffffffffu, cnt, jb);
        const uint32_t s_w = __shfl_sync(0xffffffffu, w, jb);
        const float s_invw = __shfl_sync(0xffffffffu, invw, jb);
        const uint32_t s_t0 = __shfl_sync(0xffffffffu, tile00, jb);
        for (uint32_t t = lane; t < s_cnt; t += 32) {
          const uint32_t yy = (uint32_t)(((float)t + 0.5f) * s_invw);
          const uint32_t xx = t - yy * s_w;
          const uint32_t tile = s_t0 + yy * grid_x + xx;
          const uint32_t p = ctr[tile];
          ctr[tile] = p + 1;
          if (WRITE) out[p] = s_id;
        }
        __syncwarp();
      }
      j0 = jb + 1;
    }
  }
  if (!WRITE) {
    for (int t = lane; t < num_tiles; t += 32) matrix[(size_t)t * nchunks + chunk] = ctr[t];
  }
}

// ranges[t] = [offset of (t, chunk 0), offset of (t+1, chunk 0)); empty tiles stay (0,0) like the
// reference's memset + identifyTileRanges (R3 rasterizer_impl.cu:313-321).
__global__ void __launch_bounds__(256) bin_ranges_kernel(int num_tiles, int nchunks, uint32_t total,
                                                         const uint32_t* __restrict__ matrix,
                                                         uint2* __restrict__ ranges) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= num_tiles) return;
  const uint32_t a = matrix[(size_t)t * nchunks];
  const uint32_t b = t + 1 < num_tiles ? matrix[(size_t)(t + 1) * nchunks] : total;
  ranges[t] = a == b ? make_uint2(0u, 0u) : make_uint2(a, b);
}

__global__ void __launch_bounds__(256) gather_rect_kernel(int P, const uint32_t* __restrict__ sorted_idx,
                                                          const uint2* __restrict__ rect, uint2* __restrict__ srect,
                                                          const uint32_t* __restrict__ n_dev = nullptr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = n_dev ? min(P, (int)*n_dev) : P;
  if (i < n) srect[i] = rect[sorted_idx[i]];
}

// Depth argsort + offsets (stage 1 tail).  Leaves Gaussian ids in ws.sorted_idx.
// `offsets` = exclusive scan in depth order of the binning weights (counting path; total -> ws.wtotal)
// or of the tile counts (radix path).
// Bytes to clear at ws.total before the preprocess kernel (`bucket`: the depth order will use the bucket sort).
inline size_t geom_zero_bytes(int P, bool bucket) {
  return bucket ? 128 + bucket_sort_zero_words((uint64_t)(P > 0 ? P : 1)) * sizeof(uint32_t) : 2 * sizeof(unsigned long long);
}

// PG_SORT_IMPL=radix keeps the 4-pass LSD radix sort on the band path (A/B measurements, parity tests of both).
inline bool sort_force_radix() {
  const char* e = getenv("PG_SORT_IMPL");
  return e && e[0] == 'r';
}

inline cudaError_t depth_order_and_offsets(const GeomWs& ws, int P, bool fast, cudaStream_t st, bool need_offsets = true) {
  cudaError_t e;
  const uint32_t n = (uint32_t)P;
  const uint32_t nblk = sort_num_blocks(n);
  const bool compact = !need_offsets;   // band path: only the visible Gaussians are sorted (count in ws.nvalid)
  if (compact && !sort_force_radix()) {
    // keyA | keyB | valA are carved back to back (128-B padding): >= 8 bytes per Gaussian for the (key, id) pairs
    // (bins and look-back words were zeroed together with ws.total before the preprocess kernel: geom_zero_bytes)
    return depth_bucket_sort(ws.dkey, reinterpret_cast<const uint32_t*>(ws.total + 1), ws.sort_scratch,
                             reinterpret_cast<unsigned long long*>(ws.keyA), ws.rect, ws.sorted_idx, ws.srect, ws.nvalid, n, st);
  }
  if (depth_sort_coop_fits(nblk)) {
    DepthSortArgs a;
    a.dkey = ws.dkey; a.keyA = ws.keyA; a.valA = ws.valA; a.keyB = ws.keyB; a.valB = ws.sorted_idx;
    a.hist = ws.sort_scratch; a.accum = ws.sort_scratch + (size_t)2 * nblk * 256; a.bsum = a.accum + 4 * kSortAccumWords;
    a.wsrc = need_offsets ? (fast ? ws.wts : ws.tiles) : nullptr; a.rect = ws.rect; a.srect = ws.srect; a.offsets = ws.offsets;
    a.total_out = fast ? ws.wtotal : nullptr;
    a.nvalid = ws.nvalid;
    a.n = n; a.nblk = nblk; a.compact = compact ? 1 : 0;
    return depth_sort_coop(a, st);
  }
  // 4 stable passes, 8 bits each: (dkey,iota)->(A,valA)->(B,sorted)->(A,valA)->(-,sorted); in compact mode the
  // first pass drops the culled keys and the later ones run on the device-side count
  const uint32_t* nd = compact ? ws.nvalid : nullptr;
  if (compact) e = radix_pass<uint32_t, true, true, true>(ws.dkey, nullptr, ws.keyA, ws.valA, n, 0, 8, ws.sort_scratch, st, nullptr, ws.nvalid);
  else e = radix_pass<uint32_t, true, true>(ws.dkey, nullptr, ws.keyA, ws.valA, n, 0, 8, ws.sort_scratch, st);
  if (e != cudaSuccess) return e;
  e = radix_pass<uint32_t, false, true>(ws.keyA, ws.valA, ws.keyB, ws.sorted_idx, n, 8, 8, ws.sort_scratch, st, nd);
  if (e != cudaSuccess) return e;
  e = radix_pass<uint32_t, false, true>(ws.keyB, ws.sorted_idx, ws.keyA, ws.valA, n, 16, 8, ws.sort_scratch, st, nd);
  if (e != cudaSuccess) return e;
  e = radix_pass<uint32_t, false, false>(ws.keyA, ws.valA, ws.keyB, ws.sorted_idx, n, 24, 8, ws.sort_scratch, st, nd);
  if (e != cudaSuccess) return e;
  gather_rect_kernel<<<div_up(P, 256), 256, 0, st>>>(P, ws.sorted_idx, ws.rect, ws.srect, nd);
  if (!need_offsets) return cudaGetLastError();
  if (fast) return exclusive_scan_u32(ws.wts, ws.sorted_idx, n, ws.offsets, ws.sort_scratch, ws.wtotal, st);
  return exclusive_scan_u32(ws.tiles, ws.sorted_idx, n, ws.offsets, ws.sort_scratch, nullptr, st);
}

// Emit + tile sort + ranges (stage 2 head). On return *point_list points at the sorted ids.
inline cudaError_t bin_instances(const GeomWs& g, const BinWs& b, const ImageWs& im, int P, int W, int H,
                                 int64_t R, const uint32_t** point_list, cudaStream_t st,
                                 unsigned long long cap = ~0ull) {
  const int gx = div_up(W, kTile), gy = div_up(H, kTile);
  const int T = gx * gy;
  cudaError_t e = cudaSuccess;
  *point_list = b.valA;
  // the band binning writes the range of every tile itself
  if (R > 0 && P > 0 && band_path(gx, gy)) return band_bin(b.band, P, g.nvalid, gx, gy, R, g.sorted_idx, g.srect, g.rect, b.valA, im.ranges, st, g.total, cap);
  e = cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * (size_t)T, st);
  if (e != cudaSuccess) return e;
  if (R <= 0 || P <= 0) return cudaSuccess;
  if (bin_fast_path(T)) {
    const int nchunks = bin_max_chunks(T);
    const int wpc = bin_warps_per_cta(T);
    const size_t smem = (size_t)wpc * T * sizeof(uint32_t);
    // opt in to > 48 KB dynamic shared memory (per device; cheap enough to repeat every call)
    e = cudaFuncSetAttribute(bin_walk_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(bin_walk_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    const int ctas = div_up(nchunks, wpc);
    bin_walk_kernel<false><<<ctas, wpc * 32, smem, st>>>(P, nchunks, T, gx, g.wtotal, g.sorted_idx, g.offsets,
                                                          g.srect, b.matrix, nullptr);
    const uint32_t m = (uint32_t)((size_t)T * nchunks);
    e = exclusive_scan_u32(b.matrix, nullptr, m, b.matrix, b.mscan, nullptr, st);
    if (e != cudaSuccess) return e;
    bin_ranges_kernel<<<div_up(T, 256), 256, 0, st>>>(T, nchunks, (uint32_t)R, b.matrix, im.ranges);
    bin_walk_kernel<true><<<ctas, wpc * 32, smem, st>>>(P, nchunks, T, gx, g.wtotal, g.sorted_idx, g.offsets,
                                                         g.srect, b.matrix, b.valA);
    return cudaGetLastError();
  }
  emit_instances_kernel<<<div_up(P, 256), 256, 0, st>>>(P, g.sorted_idx, g.offsets, g.tiles, g.rect, gx,
                                                        b.tkeyA, b.valA);
  const int bits = tile_key_bits(T);
  const uint16_t* sorted_keys;
  if (bits <= 8) {
    e = radix_pass<uint16_t, false, true>(b.tkeyA, b.valA, b.tkeyB, b.valB, (uint32_t)R, 0, bits, b.scratch, st);
    if (e != cudaSuccess) return e;
    sorted_keys = b.tkeyB;
    *point_list = b.valB;
  } else {
    const int lo = (bits + 1) / 2, hi = bits - lo;
    e = radix_pass<uint16_t, false, true>(b.tkeyA, b.valA, b.tkeyB, b.valB, (uint32_t)R, 0, lo, b.scratch, st);
    if (e != cudaSuccess) return e;
    e = radix_pass<uint16_t, false, true>(b.tkeyB, b.valB, b.tkeyA, b.valA, (uint32_t)R, lo, hi, b.scratch, st);
    if (e != cudaSuccess) return e;
    sorted_keys = b.tkeyA;
    *point_list = b.valA;
  }
  tile_ranges_kernel<<<(uint32_t)((R + 255) / 256), 256, 0, st>>>((uint32_t)R, sorted_keys, im.ranges);
  return cudaGetLastError();
}

inline const uint32_t* sorted_point_list(const BinWs& b, int W, int H) {
  const int T = div_up(W, kTile) * div_up(H, kTile);
  if (band_path(div_up(W, kTile), div_up(H, kTile)) || bin_fast_path(T)) return b.valA;
  return tile_sort_passes(T) == 2 ? b.valA : b.valB;
}

}  // namespace pg
