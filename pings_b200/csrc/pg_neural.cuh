// Neural-point kernels: voxel-hash kNN, warp-per-query top-k + inverse-distance interpolation
// (forward / backward / stability side effects) and the fused SDF query.
//
// Reference behaviour (a chain of ~55 eager torch ops): NP = model/neural_gaussians.py
//   radius_neighborhood_search  NP:1061-1115     query_feature  NP:506-725
//   Mapper.sdf                  utils/mapper.py:2273-2289
//
// Parity notes.  Index decisions hang on fp32 results, so the arithmetic follows torch's CPU
// evaluation exactly: IEEE division p / res then floor (NP:1069), int64 hash with C remainder and
// Python negative-index wrap (NP:1078-1084), dist2 = (dx*dx + dy*dy) + dz*dz WITHOUT fused
// multiply-add (torch squares, then sums the three terms left to right), strict `>` collision
// test against max_valid_dist2 rounded to fp32 (NP:1102-1105), index -1 reading the LAST row.
//
// B200 mapping: one warp owns one query; its 32 lanes issue the K (81) slot-table gathers of the
// neighbourhood as independent 8-byte loads (the only HBM-latency-bound step: one 32-B sector per
// probe out of an 800 MB table), keep the candidates in registers and select the nn_k nearest with
// shuffle arg-min rounds instead of a full 81-wide sort; features are gathered with one coalesced
// row read per kept neighbour.
#pragma once
#include "pg_common.cuh"

namespace pg {

struct NpMapDev {
  long long buffer_size;
  unsigned long long bs_magic;   // floor(2^64 / buffer_size): exact remainder without the 64-bit division routine
  const long long* table;
  const int* table32;            // optional int32 mirror of `table` (same slots; half the footprint), or NULL
  const float* pts;
  long long n_points;
  const long long* dx;
  int K;
  long long p0, p1, p2;
  float res, max_valid;
  const int* ts_create;
  const float* travel;
  int cur_ts;
  float diff_local;
  int time_filtering;
};

__device__ __forceinline__ long long np_wrap(long long i, long long n) { return i < 0 ? i + n : i; }

// One probe of the neighbourhood: returns the global index (or -1) and the squared distance.
// Hash contribution of neighbourhood offset k.  int64 arithmetic wraps, so
// (g + dx) . p == g . p + dx . p exactly and the per-offset part can be hoisted out of the query loop.
__device__ __forceinline__ long long np_hash_delta(const NpMapDev& m, int k) {
  return (long long)((unsigned long long)m.dx[3 * k + 0] * (unsigned long long)m.p0 +
                     (unsigned long long)m.dx[3 * k + 1] * (unsigned long long)m.p1 +
                     (unsigned long long)m.dx[3 * k + 2] * (unsigned long long)m.p2);
}
__device__ __forceinline__ long long np_hash_base(const NpMapDev& m, long long gx, long long gy, long long gz) {
  return (long long)((unsigned long long)gx * (unsigned long long)m.p0 + (unsigned long long)gy * (unsigned long long)m.p1 +
                     (unsigned long long)gz * (unsigned long long)m.p2);
}

// h mod buffer_size in [0, buffer_size): torch.fmod on int64 (NP:1078-1084) is the C remainder, and the negative
// results then wrap as Python indices (NP:1084) - together the floored modulo.  Barrett: q = hi64(|h| *
// floor(2^64/d)) is at most 1 below the true quotient, so r = |h| - q*d < 2d.
__device__ __forceinline__ long long np_slot_of(const NpMapDev& m, long long h) {
  const unsigned long long d = (unsigned long long)m.buffer_size;
  const unsigned long long ah = h < 0 ? 0ull - (unsigned long long)h : (unsigned long long)h;
  unsigned long long r = ah - __umul64hi(ah, m.bs_magic) * d;
  if (r >= d) r -= d;
  if (r >= d) r -= d;
  return (h < 0 && r != 0) ? (long long)(d - r) : (long long)r;
}

// Probe of slot `hv` of the table.
__device__ __forceinline__ void np_probe_slot(const NpMapDev& m, long long hv, float qx, float qy, float qz,
                                              long long& idx, float& dist2) {
  idx = m.table32 ? (long long)m.table32[hv] : m.table[hv];
  if (m.time_filtering) {
    const int ts = m.ts_create[np_wrap(idx, m.n_points)];
    const float diff = fabsf(m.travel[m.cur_ts] - m.travel[ts]);
    if (!(diff < m.diff_local)) idx = -1;
  }
  const long long pi = np_wrap(idx, m.n_points);
  const float sx = __fsub_rn(m.pts[3 * pi + 0], qx);
  const float sy = __fsub_rn(m.pts[3 * pi + 1], qy);
  const float sz = __fsub_rn(m.pts[3 * pi + 2], qz);
  dist2 = __fadd_rn(__fadd_rn(__fmul_rn(sx, sx), __fmul_rn(sy, sy)), __fmul_rn(sz, sz));
  if (idx == -1) dist2 = m.max_valid;
  if (dist2 > m.max_valid) idx = -1;
}

// `h` = un-reduced hash of the probed cell.
__device__ __forceinline__ void np_probe_h(const NpMapDev& m, long long h, float qx, float qy, float qz,
                                           long long& idx, float& dist2) {
  np_probe_slot(m, np_slot_of(m, h), qx, qy, qz, idx, dist2);
}

__device__ __forceinline__ long long np_cell(float q, float res) { return (long long)floorf(__fdiv_rn(q, res)); }

__device__ __forceinline__ void np_probe(const NpMapDev& m, long long gx, long long gy, long long gz, int k,
                                         float qx, float qy, float qz, long long& idx, float& dist2) {
  const long long h = (long long)((unsigned long long)np_hash_base(m, gx, gy, gz) + (unsigned long long)np_hash_delta(m, k));
  np_probe_h(m, h, qx, qy, qz, idx, dist2);
}

// ---- radius_neighborhood_search: one thread per (query, probe) --------------------------------
__global__ void __launch_bounds__(256) np_radius_search_kernel(NpMapDev m, long long N, const float* __restrict__ q,
                                                               float* __restrict__ out_d, long long* __restrict__ out_i) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= N * m.K) return;
  const long long n = t / m.K;
  const int k = (int)(t - n * m.K);
  const float qx = q[3 * n], qy = q[3 * n + 1], qz = q[3 * n + 2];
  long long idx;
  float d;
  np_probe(m, np_cell(qx, m.res), np_cell(qy, m.res), np_cell(qz, m.res), k, qx, qy, qz, idx, d);
  out_d[t] = d;
  out_i[t] = idx;
}

// ---- warp-per-query kNN + top-k ---------------------------------------------------------------
struct NpQueryDev {
  int nn_k, query_locally, use_only_measured, use_only_valid, weighted_first, after_pgo;
  int F_geo, F_color;
  const uint8_t* free_mask;
  const uint8_t* valid_mask;
  const long long* g2l;
  const float* points;
  const float* quats;
  const float* cert;
  const float* geo;
  const float* color;
};

constexpr float kNpInvalidDist = 9e3f;  // NP:562
constexpr float kNpEps = 1e-15f;        // NP:642
constexpr int kNpMaxK = 32;             // nn_k <= 32 (one kept neighbour per lane)

// Runs the probes of one query on the calling warp and leaves the nn_k nearest valid neighbours in
// lanes [0, nn_k): (sel_d, sel_i) with sel_i the FINAL (local) index or -1. Returns nn_count.
// Per-lane hash deltas of the probes c*32 + lane (query independent: hoist out of persistent loops).
// dr = the same deltas reduced to [0, buffer_size): (h0 + dh) mod d == (h0 mod d + dh mod d) mod d as long as the
// int64 sum does not wrap, so a query pays ONE 64-bit Barrett reduction (its base cell) and every probe a 32-bit
// add + conditional subtract instead of ~30 instructions of 64-bit arithmetic (buffer_size < 2^31).
template <int CPL>
__device__ __forceinline__ void np_lane_deltas(const NpMapDev& m, int lane, long long (&dh)[CPL], uint32_t (&dr)[CPL]) {
#pragma unroll
  for (int c = 0; c < CPL; ++c) {
    dh[c] = (c * 32 + lane < m.K) ? np_hash_delta(m, c * 32 + lane) : 0;
    dr[c] = (uint32_t)np_slot_of(m, dh[c]);
  }
}

// Since round 2 (same scheme as sdf_query_half_kernel, pg_sdf.cuh): the slot-table loads of a lane are issued back to
// back, the OCCUPIED slots (about 8 of 81) are compacted by ballot + popc into per-warp shared memory in probe order,
// distances / masks / local indices are computed only for those, and the nn_k nearest are picked by rank counting over
// the surviving candidates - rank = number of candidates with a smaller (distance bits, probe position) key, i.e. the
// stable order of torch.sort over the probe axis (NP:565); keys are unique.  Before: every lane did the distance
// arithmetic of its CPL probes and the selection took nn_k arg-min rounds of 2 REDUX + bookkeeping (a third of the
// instructions of a query).  Callers run with 8 warps per CTA (kNpWarps / kSdfWarps).
constexpr int kNpKnnWarps = 8;

template <int CPL>
__device__ __forceinline__ int np_warp_knn(const NpMapDev& m, const NpQueryDev& a, const long long (&dh)[CPL],
                                           const uint32_t (&dr)[CPL], float qx, float qy, float qz, int lane, float& sel_d,
                                           int& sel_i) {
  constexpr int kSlots = CPL * 32;
  __shared__ uint32_t s_hit[kNpKnnWarps][2][kSlots];          // occupied slots: point index, probe position
  __shared__ unsigned long long s_key[kNpKnnWarps][kSlots];  // candidates: dist bits << 32 | probe position
  __shared__ uint32_t s_val[kNpKnnWarps][kSlots];            // candidates: final (local) index
  __shared__ float s_od[kNpKnnWarps][kNpMaxK];
  __shared__ int s_oi[kNpKnnWarps][kNpMaxK];
  const int wp = (threadIdx.x >> 5) & (kNpKnnWarps - 1);
  uint32_t* hit_i = s_hit[wp][0];
  uint32_t* hit_p = s_hit[wp][1];
  unsigned long long* ck = s_key[wp];
  uint32_t* cv = s_val[wp];
  const long long h0 = np_hash_base(m, np_cell(qx, m.res), np_cell(qy, m.res), np_cell(qz, m.res));
  // fast path: no int64 wrap possible in h0 + dh (|h0| < 2^62, |dh| < 2^40) and 32-bit slots
  const bool fast = m.buffer_size < (1ll << 31) && h0 > -(1ll << 62) && h0 < (1ll << 62);
  const uint32_t b0 = fast ? (uint32_t)np_slot_of(m, h0) : 0u;
  const uint32_t d32 = (uint32_t)m.buffer_size;
  // ---- phase 1: slot-table probes ----
  long long pidx[CPL];
#pragma unroll
  for (int c = 0; c < CPL; ++c) {
    pidx[c] = -1;
    if (c * 32 + lane < m.K) {
      long long slot;
      if (fast) {
        uint32_t sl = b0 + dr[c];
        if (sl >= d32) sl -= d32;
        slot = (long long)sl;
      } else {
        slot = np_slot_of(m, (long long)((unsigned long long)h0 + (unsigned long long)dh[c]));
      }
      pidx[c] = m.table32 ? (long long)m.table32[slot] : m.table[slot];
    }
  }
  if (m.time_filtering) {
#pragma unroll
    for (int c = 0; c < CPL; ++c) {
      if (pidx[c] >= 0) {
        const float diff = fabsf(m.travel[m.cur_ts] - m.travel[m.ts_create[pidx[c]]]);
        if (!(diff < m.diff_local)) pidx[c] = -1;
      }
    }
  }
  // ---- phase 2: compact the occupied slots, in probe order ----
  const uint32_t below = (1u << lane) - 1u;
  uint32_t nhit = 0;
#pragma unroll
  for (int c = 0; c < CPL; ++c) {
    const bool hit = pidx[c] >= 0;
    const uint32_t bm = __ballot_sync(0xffffffffu, hit);
    if (hit) {
      const uint32_t r = nhit + __popc(bm & below);
      hit_i[r] = (uint32_t)pidx[c];
      hit_p[r] = (uint32_t)(c * 32 + lane);
    }
    nhit += __popc(bm);
  }
  __syncwarp();
  // ---- phase 3: distances (np_probe_slot's arithmetic), masks, local index; compact the survivors ----
  uint32_t ncand = 0;
  for (uint32_t t0 = 0; t0 < nhit; t0 += 32) {
    const uint32_t t = t0 + lane;
    long long idx = -1;
    float d = 0.f;
    uint32_t pos = 0;
    if (t < nhit) {
      idx = (long long)hit_i[t];
      pos = hit_p[t];
      const float sx = __fsub_rn(m.pts[3 * idx + 0], qx);
      const float sy = __fsub_rn(m.pts[3 * idx + 1], qy);
      const float sz = __fsub_rn(m.pts[3 * idx + 2], qz);
      d = __fadd_rn(__fadd_rn(__fmul_rn(sx, sx), __fmul_rn(sy, sy)), __fmul_rn(sz, sz));
      if (d > m.max_valid) idx = -1;
      if (idx >= 0 && a.use_only_measured && a.free_mask[idx]) idx = -1;
      if (idx >= 0 && a.use_only_valid && !a.valid_mask[idx]) idx = -1;
      if (idx >= 0 && a.query_locally) idx = a.g2l[idx];
    }
    const bool ok = idx >= 0;
    const uint32_t bm = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const uint32_t r = ncand + __popc(bm & below);
      ck[r] = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned long long)pos;
      cv[r] = (uint32_t)idx;
    }
    ncand += __popc(bm);
  }
  __syncwarp();
  const int cnt = (int)ncand;
  // ---- phase 4: nn_k nearest by rank counting ----
  for (uint32_t t = lane; t < ncand; t += 32) {
    const unsigned long long key = ck[t];
    uint32_t rank = 0;
    for (uint32_t k = 0; k < ncand; ++k) rank += ck[k] < key ? 1u : 0u;
    if (rank < (uint32_t)a.nn_k) {
      s_od[wp][rank] = __uint_as_float((uint32_t)(key >> 32));
      s_oi[wp][rank] = (int)cv[t];
    }
  }
  __syncwarp();
  sel_d = kNpInvalidDist;
  sel_i = -1;
  if (lane < min(a.nn_k, cnt)) {
    sel_d = s_od[wp][lane];
    sel_i = s_oi[wp][lane];
  }
  __syncwarp();   // the warp's next query overwrites the scratch
  return cnt;
}

// IDW weights of the kept neighbours (NP:642-662): lane j < nn_k holds w_j.
__device__ __forceinline__ float np_idw_weight(float sel_d, int sel_i, int nn_k, int cnt, int lane) {
  float u = 0.f;
  if (lane < nn_k) {
    u = sel_i >= 0 ? __fdiv_rn(1.0f, sel_d + kNpEps) : 0.f;
    if (cnt == 0) u = kNpEps;
  }
  float s = u;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  float w = lane < nn_k ? __fdiv_rn(u, s) : 0.f;
  if (sel_i < 0) w = 0.f;
  return w;
}

// Neighbour vector q - p_j, optionally rotated into the point frame (utils/tools.py:743-751).
__device__ __forceinline__ float3 np_neighbor_vec(const NpQueryDev& a, int idx, float qx, float qy, float qz) {
  float3 v = make_float3(qx - a.points[3 * (size_t)idx], qy - a.points[3 * (size_t)idx + 1],
                         qz - a.points[3 * (size_t)idx + 2]);
  if (a.after_pgo) {
    const float4 qq = reinterpret_cast<const float4*>(a.quats)[idx];
    const float w = qq.x, x = -qq.y, y = -qq.z, z = -qq.w;
    const float tx = 2.f * (y * v.z - z * v.y), ty = 2.f * (z * v.x - x * v.z), tz = 2.f * (x * v.y - y * v.x);
    v = make_float3(v.x + w * tx + (y * tz - z * ty), v.y + w * ty + (z * tx - x * tz), v.z + w * tz + (x * ty - y * tx));
  }
  return v;
}
// Transposed rotation (gradient of the above w.r.t. its input vector).
__device__ __forceinline__ float3 np_neighbor_vec_grad(const NpQueryDev& a, int idx, float3 g) {
  if (!a.after_pgo) return g;
  const float4 qq = reinterpret_cast<const float4*>(a.quats)[idx];
  const float w = qq.x, x = qq.y, y = qq.z, z = qq.w;  // conjugate of the forward rotation
  const float tx = 2.f * (y * g.z - z * g.y), ty = 2.f * (z * g.x - x * g.z), tz = 2.f * (x * g.y - y * g.x);
  return make_float3(g.x + w * tx + (y * tz - z * ty), g.y + w * ty + (z * tx - x * tz), g.z + w * tz + (x * ty - y * tx));
}

constexpr int kNpWarps = 4;  // warps (queries in flight) per CTA

// Writes one feature family of query n: weighted sum (weighted_first) or the per-neighbour rows.
__device__ __forceinline__ void np_write_features(const NpQueryDev& a, const float* __restrict__ table, int F,
                                                  float* __restrict__ out, long long n, float qx, float qy, float qz,
                                                  float sel_w, int sel_i, int lane) {
  const int Fo = F + 3;
  if (a.weighted_first) {
    for (int f0 = 0; f0 < Fo; f0 += 32) {
      const int f = f0 + lane;
      float acc = 0.f;
      for (int j = 0; j < a.nn_k; ++j) {
        const int idx = __shfl_sync(0xffffffffu, sel_i, j);
        const float w = __shfl_sync(0xffffffffu, sel_w, j);
        if (idx < 0 || f >= Fo) continue;
        float v;
        if (f < F) {
          v = table[(size_t)idx * F + f];
        } else {
          const float3 nv = np_neighbor_vec(a, idx, qx, qy, qz);
          v = f == F ? nv.x : (f == F + 1 ? nv.y : nv.z);
        }
        acc += v * w;
      }
      if (f < Fo) out[n * Fo + f] = acc;
    }
  } else {
    for (int j = 0; j < a.nn_k; ++j) {
      const int idx = __shfl_sync(0xffffffffu, sel_i, j);
      for (int f = lane; f < Fo; f += 32) {
        float v = 0.f;
        if (idx >= 0) {
          if (f < F) {
            v = table[(size_t)idx * F + f];
          } else {
            const float3 nv = np_neighbor_vec(a, idx, qx, qy, qz);
            v = f == F ? nv.x : (f == F + 1 ? nv.y : nv.z);
          }
        }
        out[(n * a.nn_k + j) * Fo + f] = v;
      }
    }
  }
}

template <int CPL>
__global__ void __launch_bounds__(kNpWarps * 32) np_query_kernel(NpMapDev m, NpQueryDev a, long long N,
                                                                const float* __restrict__ q, float* __restrict__ geo_vec,
                                                                float* __restrict__ color_vec, float* __restrict__ weight,
                                                                long long* __restrict__ nn_counts,
                                                                float* __restrict__ certainty, int* __restrict__ knn_idx,
                                                                float* __restrict__ knn_d) {
  const int lane = threadIdx.x & 31;
  const long long n = (long long)blockIdx.x * kNpWarps + (threadIdx.x >> 5);
  if (n >= N) return;
  const float qx = q[3 * n], qy = q[3 * n + 1], qz = q[3 * n + 2];
  float sel_d;
  int sel_i;
  long long dh[CPL];
  uint32_t dr[CPL];
  np_lane_deltas<CPL>(m, lane, dh, dr);
  const int cnt = np_warp_knn<CPL>(m, a, dh, dr, qx, qy, qz, lane, sel_d, sel_i);
  const float w = np_idw_weight(sel_d, sel_i, a.nn_k, cnt, lane);
  if (lane < a.nn_k) {
    weight[n * a.nn_k + lane] = w;
    if (knn_idx) knn_idx[n * a.nn_k + lane] = sel_i;
    if (knn_d) knn_d[n * a.nn_k + lane] = sel_d;
  }
  if (certainty) {
    float c = (lane < a.nn_k && sel_i >= 0) ? a.cert[sel_i] * w : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) certainty[n] = c;
  }
  if (lane == 0) nn_counts[n] = cnt;
  if (geo_vec) np_write_features(a, a.geo, a.F_geo, geo_vec, n, qx, qy, qz, w, sel_i, lane);
  if (color_vec) np_write_features(a, a.color, a.F_color, color_vec, n, qx, qy, qz, w, sel_i, lane);
}

// ---- stability side effects (NP:664-689) -------------------------------------------------------
__global__ void __launch_bounds__(256) np_accumulate_kernel(long long total, int nn_k, const int* __restrict__ knn_idx,
                                                            const float* __restrict__ weight, const int* __restrict__ query_ts,
                                                            float* __restrict__ cert, int* __restrict__ ts_update) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int idx = knn_idx[t];
  if (idx < 0) return;
  atomicAdd(&cert[idx], weight[t]);
  if (query_ts && ts_update) atomicMax(&ts_update[idx], query_ts[t / nn_k]);
}

// ---- backward of query_feature ----------------------------------------------------------------
// Warp per query. With s_j = dL/dw_j:  weighted_first:  s_j = <d_out[:F], feat_j> + <d_out[F:], nv_j> (+ d_weight_j)
//                                      otherwise:       s_j = d_weight_j
// w_j = u_j / S, u_j = 1/(d_j + eps)  =>  dL/du_j = (s_j - sum_i w_i s_i) / S,
// d u_j / d q = 2 u_j^2 (p_j - q);  the neighbour vectors add  sum_j R_j^T (w_j d_out[F:])  (or d_out[n,j,F:]).
__global__ void __launch_bounds__(kNpWarps * 32) np_query_bwd_kernel(
    NpQueryDev a, long long N, const float* __restrict__ q, const int* __restrict__ knn_idx,
    const float* __restrict__ knn_d, const float* __restrict__ weight, const long long* __restrict__ nn_counts,
    const float* __restrict__ d_geo, const float* __restrict__ d_color, const float* __restrict__ d_weight,
    float* __restrict__ d_geo_feat, float* __restrict__ d_color_feat, float* __restrict__ d_query) {
  const int lane = threadIdx.x & 31;
  const long long n = (long long)blockIdx.x * kNpWarps + (threadIdx.x >> 5);
  if (n >= N) return;
  const float qx = q[3 * n], qy = q[3 * n + 1], qz = q[3 * n + 2];
  int sel_i = -1;
  float sel_d = kNpInvalidDist, sel_w = 0.f;
  if (lane < a.nn_k) {
    sel_i = knn_idx[n * a.nn_k + lane];
    sel_d = knn_d[n * a.nn_k + lane];
    sel_w = weight[n * a.nn_k + lane];
  }
  float s_mine = (d_weight && lane < a.nn_k) ? d_weight[n * a.nn_k + lane] : 0.f;
  float3 gq = make_float3(0.f, 0.f, 0.f);  // accumulated on lane 0 after reductions

  for (int fam = 0; fam < 2; ++fam) {
    const float* dvec = fam == 0 ? d_geo : d_color;
    if (!dvec) continue;
    const int F = fam == 0 ? a.F_geo : a.F_color;
    const int Fo = F + 3;
    const float* table = fam == 0 ? a.geo : a.color;
    float* dtable = fam == 0 ? d_geo_feat : d_color_feat;
    for (int j = 0; j < a.nn_k; ++j) {
      const int idx = __shfl_sync(0xffffffffu, sel_i, j);
      const float w = __shfl_sync(0xffffffffu, sel_w, j);
      if (idx < 0) continue;
      const float* drow = a.weighted_first ? dvec + n * Fo : dvec + (n * a.nn_k + j) * Fo;
      float part = 0.f;
      for (int f = lane; f < F; f += 32) {
        const float g = drow[f];
        if (a.weighted_first) {
          part += g * table[(size_t)idx * F + f];
          if (dtable) atomicAdd(&dtable[(size_t)idx * F + f], g * w);
        } else if (dtable) {
          atomicAdd(&dtable[(size_t)idx * F + f], g);
        }
      }
      const float3 gv = make_float3(drow[F], drow[F + 1], drow[F + 2]);
      if (a.weighted_first) {
        const float3 nv = np_neighbor_vec(a, idx, qx, qy, qz);
        if (lane == 0) part += gv.x * nv.x + gv.y * nv.y + gv.z * nv.z;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
        if (lane == j) s_mine += part;
      }
      const float ws = a.weighted_first ? w : 1.f;
      const float3 r = np_neighbor_vec_grad(a, idx, make_float3(gv.x * ws, gv.y * ws, gv.z * ws));
      gq.x += r.x; gq.y += r.y; gq.z += r.z;
    }
  }
  if (!d_query) return;
  // IDW weight path
  const bool valid = lane < a.nn_k && sel_i >= 0;
  const float u = valid ? __fdiv_rn(1.0f, sel_d + kNpEps) : 0.f;
  float S = u, ws = valid ? sel_w * s_mine : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    S += __shfl_xor_sync(0xffffffffu, S, o);
    ws += __shfl_xor_sync(0xffffffffu, ws, o);
  }
  float3 gw = make_float3(0.f, 0.f, 0.f);
  if (valid && nn_counts[n] > 0) {
    const float du = (s_mine - ws) / S;
    const float c = du * 2.f * u * u;
    gw = make_float3(c * (a.points[3 * (size_t)sel_i] - qx), c * (a.points[3 * (size_t)sel_i + 1] - qy),
                     c * (a.points[3 * (size_t)sel_i + 2] - qz));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    gw.x += __shfl_xor_sync(0xffffffffu, gw.x, o);
    gw.y += __shfl_xor_sync(0xffffffffu, gw.y, o);
    gw.z += __shfl_xor_sync(0xffffffffu, gw.z, o);
  }
  if (lane == 0) {
    d_query[3 * n + 0] = gq.x + gw.x;
    d_query[3 * n + 1] = gq.y + gw.y;
    d_query[3 * n + 2] = gq.z + gw.z;
  }
}

}  // namespace pg
