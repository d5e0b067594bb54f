// Layout of the rasterizer workspaces + kernels shared by the 3DGS and surfel variants:
// instance emission, tile ranges, depth ordering.
//
// Pipeline (differs from the reference's preprocess -> scan -> duplicate -> 64-bit sort,
// R3 cuda_rasterizer/rasterizer_impl.cu:198-344, but yields bit-identical point_list/ranges):
//   1. preprocess writes one packed record per Gaussian + a u32 depth key + its tile rect
//   2. Gaussians are argsorted by depth key (4 LSD passes over P items, not R instances):
//      the low 32 bits of the reference's 64-bit key are identical for all instances of a
//      Gaussian, so those LSD passes commute with instance duplication
//   3. instances are emitted in depth order with a 16-bit tile key, then stably sorted on the
//      tile bits only (1-2 passes over R items of 6 bytes instead of 6 passes of 12 bytes)
#pragma once
#include "pg_common.cuh"
#include "pg_sort.cuh"

namespace pg {

constexpr int kTile = 16;

struct GeomWs {
  float4* rec;            // [3P] packed per-Gaussian record (see pg_raster3.cu)
  float4* ext;            // [P] surfel: (n_view.xyz, zx) or null (pg_rasters_fwd.cuh)
  float* zyv;             // [P] surfel: zy or null
  uint32_t* dkey;         // [P] depth bits, 0xFFFFFFFF when culled
  uint32_t* tiles;        // [P] tiles touched
  uint2* rect;            // [P] x0 | x1<<16 , y0 | y1<<16
  uint32_t* keyA;         // [P] sort ping
  uint32_t* keyB;         // [P] sort pong
  uint32_t* valA;         // [P]
  uint32_t* sorted_idx;   // [P] Gaussian ids in depth order (culled ones last)
  uint32_t* offsets;      // [P] exclusive scan of tiles in depth order
  uint8_t* clamped;       // [3P] SH clamp flags
  uint32_t* sort_scratch; // radix/scan scratch for P items
  unsigned long long* total;  // [1] number of instances
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
  w.rect = c.take<uint2>(p);
  w.keyA = c.take<uint32_t>(p);
  w.keyB = c.take<uint32_t>(p);
  w.valA = c.take<uint32_t>(p);
  w.sorted_idx = c.take<uint32_t>(p);
  w.offsets = c.take<uint32_t>(p);
  w.clamped = c.take<uint8_t>(3 * p);
  w.sort_scratch = c.take<uint32_t>(sort_scratch_elems(p));
  w.total = c.take<unsigned long long>(2);
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

struct BinWs {
  uint16_t* tkeyA;   // [R] tile keys ping
  uint16_t* tkeyB;   // [R] pong
  uint32_t* valA;    // [R]
  uint32_t* valB;    // [R]
  uint32_t* scratch; // radix scratch for R items
};

inline size_t carve_binning(int64_t R, void* base, BinWs* ws) {
  Carver c(base);
  BinWs w;
  size_t r = (size_t)(R > 0 ? R : 1);
  w.tkeyA = c.take<uint16_t>(r);
  w.tkeyB = c.take<uint16_t>(r);
  w.valA = c.take<uint32_t>(r);
  w.valB = c.take<uint32_t>(r);
  w.scratch = c.take<uint32_t>(sort_scratch_elems(r));
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

// Depth argsort + offsets (stage 1 tail).  Leaves Gaussian ids in ws.sorted_idx.
inline cudaError_t depth_order_and_offsets(const GeomWs& ws, int P, cudaStream_t st) {
  cudaError_t e;
  const uint32_t n = (uint32_t)P;
  // 4 stable passes, 8 bits each: (dkey,iota)->(A,valA)->(B,sorted)->(A,valA)->(-,sorted)
  e = radix_pass<uint32_t, true, true>(ws.dkey, nullptr, ws.keyA, ws.valA, n, 0, 8, ws.sort_scratch, st);
  if (e != cudaSuccess) return e;
  e = radix_pass<uint32_t, false, true>(ws.keyA, ws.valA, ws.keyB, ws.sorted_idx, n, 8, 8, ws.sort_scratch, st);
  if (e != cudaSuccess) return e;
  e = radix_pass<uint32_t, false, true>(ws.keyB, ws.sorted_idx, ws.keyA, ws.valA, n, 16, 8, ws.sort_scratch, st);
  if (e != cudaSuccess) return e;
  e = radix_pass<uint32_t, false, false>(ws.keyA, ws.valA, ws.keyB, ws.sorted_idx, n, 24, 8, ws.sort_scratch, st);
  if (e != cudaSuccess) return e;
  return exclusive_scan_u32(ws.tiles, ws.sorted_idx, n, ws.offsets, ws.sort_scratch, nullptr, st);
}

// Emit + tile sort + ranges (stage 2 head). On return *point_list points at the sorted ids.
inline cudaError_t bin_instances(const GeomWs& g, const BinWs& b, const ImageWs& im, int P, int W, int H,
                                 int64_t R, const uint32_t** point_list, cudaStream_t st) {
  const int gx = div_up(W, kTile), gy = div_up(H, kTile);
  const int T = gx * gy;
  cudaError_t e = cudaMemsetAsync(im.ranges, 0, sizeof(uint2) * (size_t)T, st);
  if (e != cudaSuccess) return e;
  *point_list = b.valA;
  if (R <= 0 || P <= 0) return cudaSuccess;
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
  return tile_sort_passes(T) == 2 ? b.valA : b.valB;
}

}  // namespace pg
