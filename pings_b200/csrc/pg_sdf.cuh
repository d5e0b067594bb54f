// Fused inference SDF query = Mapper.sdf under no_grad (utils/mapper.py:2273-2289):
//   hash kNN -> top-k -> IDW -> decoder MLP -> * sdf_scale (-> weighted mean over k).
// One warp per query, persistent CTAs: the decoder weights are staged (transposed) in shared
// memory once per CTA, the interpolated feature vector never leaves the SM, so the kernel's only
// HBM traffic is the slot-table probes, the neighbour rows and 4 B of output per query.
#pragma once
#include "pg_mlp.cuh"
#include "pg_neural.cuh"

namespace pg {

constexpr int kSdfWarps = 8;

// floats of shared memory: transposed weights + biases of every layer, + per-warp 2 x Dmax scratch
__host__ __device__ inline size_t sdf_smem_floats(const MlpDev& m) {
  size_t w = 0;
  for (int l = 0; l <= m.levels; ++l) w += (size_t)mlp_layer_in(m, l) * mlp_layer_out(m, l) + mlp_layer_out(m, l);
  return w + (size_t)kSdfWarps * 2 * mlp_max_dim(m) + 32;
}

// Runs the MLP for the vector held in xa[0..in_dim) (per-warp shared scratch); returns out[0] on all lanes.
__device__ __forceinline__ float sdf_warp_mlp(const MlpDev& m, const float* __restrict__ wsm, float* xa, float* xb,
                                              int lane) {
  const float* wp = wsm;
  float* cur = xa;
  float* nxt = xb;
  float result = 0.f;
  for (int l = 0; l <= m.levels; ++l) {
    const int Din = mlp_layer_in(m, l), Dout = mlp_layer_out(m, l);
    const float* Wt = wp;            // [Din][Dout]
    const float* b = wp + Din * Dout;
    wp = b + Dout;
    if (l == m.levels && Dout == 1) {
      float part = 0.f;
      for (int f = lane; f < Din; f += 32) part += cur[f] * Wt[f];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      result = part + b[0];
    } else {
      for (int u = lane; u < Dout; u += 32) {
        float acc = 0.f;
        for (int f = 0; f < Din; ++f) acc += cur[f] * Wt[f * Dout + u];
        acc += b[u];
        nxt[u] = l < m.levels ? mlp_act(acc, m.leaky) : acc;
      }
      __syncwarp();
      if (l == m.levels) result = nxt[0];
      float* t = cur; cur = nxt; nxt = t;
    }
  }
  __syncwarp();
  return result;
}

template <int CPL>
__global__ void __launch_bounds__(kSdfWarps * 32) sdf_query_kernel(NpMapDev m, NpQueryDev a, MlpDev mlp, float sdf_scale,
                                                                  long long N, const float* __restrict__ q,
                                                                  float* __restrict__ sdf, float* __restrict__ sdf_std,
                                                                  long long* __restrict__ nn_counts) {
  extern __shared__ float smem[];
  // stage weights transposed: for each layer [Din][Dout] then bias[Dout] (zeros when absent)
  {
    float* wp = smem;
    for (int l = 0; l <= mlp.levels; ++l) {
      const int Din = mlp_layer_in(mlp, l), Dout = mlp_layer_out(mlp, l);
      for (int i = threadIdx.x; i < Din * Dout; i += blockDim.x) {
        const int f = i / Dout, u = i - f * Dout;
        wp[i] = mlp.W[l][(size_t)u * Din + f];
      }
      for (int u = threadIdx.x; u < Dout; u += blockDim.x) wp[Din * Dout + u] = mlp.B[l] ? mlp.B[l][u] : 0.f;
      wp += Din * Dout + Dout;
    }
  }
  size_t woff = 0;
  for (int l = 0; l <= mlp.levels; ++l) woff += (size_t)mlp_layer_in(mlp, l) * mlp_layer_out(mlp, l) + mlp_layer_out(mlp, l);
  const int Dmax = mlp_max_dim(mlp);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* xa = smem + woff + (size_t)warp * 2 * Dmax;
  float* xb = xa + Dmax;
  __syncthreads();
  const int F = a.F_geo, Fo = F + 3;
  long long dh[CPL];
  uint32_t dr[CPL];
  np_lane_deltas<CPL>(m, lane, dh, dr);
  for (long long n = (long long)blockIdx.x * kSdfWarps + warp; n < N; n += (long long)gridDim.x * kSdfWarps) {
    const float qx = q[3 * n], qy = q[3 * n + 1], qz = q[3 * n + 2];
    float sel_d;
    int sel_i;
    const int cnt = np_warp_knn<CPL>(m, a, dh, dr, qx, qy, qz, lane, sel_d, sel_i);
    const float w = np_idw_weight(sel_d, sel_i, a.nn_k, cnt, lane);
    if (lane == 0 && nn_counts) nn_counts[n] = cnt;
    if (a.weighted_first) {
      for (int f0 = 0; f0 < Fo; f0 += 32) {
        const int f = f0 + lane;
        float acc = 0.f;
        for (int j = 0; j < a.nn_k; ++j) {
          const int idx = __shfl_sync(0xffffffffu, sel_i, j);
          const float wj = __shfl_sync(0xffffffffu, w, j);
          if (idx < 0 || f >= Fo) continue;
          float v;
          if (f < F) {
            v = a.geo[(size_t)idx * F + f];
          } else {
            const float3 nv = np_neighbor_vec(a, idx, qx, qy, qz);
            v = f == F ? nv.x : (f == F + 1 ? nv.y : nv.z);
          }
          acc += v * wj;
        }
        if (f < Fo) xa[f] = acc;
      }
      __syncwarp();
      const float s = sdf_warp_mlp(mlp, smem, xa, xb, lane) * sdf_scale;
      if (lane == 0) sdf[n] = s;
    } else {
      float sj_mine = 0.f;  // lane j keeps the prediction of neighbour j
      for (int j = 0; j < a.nn_k; ++j) {
        const int idx = __shfl_sync(0xffffffffu, sel_i, j);
        for (int f = lane; f < Fo; f += 32) {
          float v = 0.f;
          if (idx >= 0) {
            if (f < F) {
              v = a.geo[(size_t)idx * F + f];
            } else {
              const float3 nv = np_neighbor_vec(a, idx, qx, qy, qz);
              v = f == F ? nv.x : (f == F + 1 ? nv.y : nv.z);
            }
          }
          xa[f] = v;
        }
        __syncwarp();
        const float s = sdf_warp_mlp(mlp, smem, xa, xb, lane) * sdf_scale;
        if (lane == j) sj_mine = s;
      }
      float mean = lane < a.nn_k ? sj_mine * w : 0.f;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mean += __shfl_xor_sync(0xffffffffu, mean, o);
      if (lane == 0) sdf[n] = mean;
      if (sdf_std) {
        const float d = sj_mine - mean;
        float var = lane < a.nn_k ? w * d * d : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
        if (lane == 0) sdf_std[n] = sqrtf(var);
      }
    }
  }
}


// ---- batched variant: warp-cooperative kNN, then one query per LANE for IDW + gather + MLP -----------------------
// The warp-per-query kernel above spends more instructions after the kNN than in it (ncu r01: 1,348 warp
// instructions per query): its MLP runs 32 lanes wide on ONE query, so a 11 -> 64 -> 1 decoder costs ~100 warp
// instructions per query.  Here a warp takes 32 queries: the kNN of each is still warp-cooperative (81 probes over
// 32 lanes), its nn_k results are parked in shared memory, and then lane l finishes query l on its own - weights,
// feature gather and the MLP with the (transposed) weights broadcast from shared memory, 8 output units per pass
// (1 activation load + 2 LDS.128 + 8 FMA per 8 MACs).  A 704-MAC layer then costs ~30 warp instructions per query.
// (two-layer decoders in -> HID -> 1 with HID a multiple of 8 up to 64: geo_mlp_level 1, the PIN-SLAM and KITTI-GS
// configurations; the hidden layer then lives in registers and only the input vector is staged per lane)
__host__ __device__ inline size_t sdf_batch_warp_floats(const MlpDev& m, int nn_k) {
  return (size_t)32 * nn_k * 3 + 32 + (size_t)32 * m.in_dim;
}
__host__ __device__ inline size_t sdf_batch_smem_floats(const MlpDev& m, int nn_k) {
  size_t w = (size_t)m.in_dim * m.hidden + m.hidden + m.hidden + 1;
  w = (w + 3) & ~(size_t)3;
  return w + (size_t)kSdfWarps * sdf_batch_warp_floats(m, nn_k);
}
__host__ __device__ inline bool sdf_batch_supported(const MlpDev& m) {
  return m.levels == 1 && m.out_dim == 1 && (m.hidden == 64 || m.hidden == 32);
}

// in -> HID -> 1 for the vector stored lane-strided in x[f * 32 + lane]; weights in shared memory:
// Wt1 [Din][HID] (transposed), b1 [HID], w2 [HID], b2 [1].
template <int HID>
__device__ __forceinline__ float sdf_lane_mlp2(const float* __restrict__ wsm, int Din, bool leaky, const float* x, int lane) {
  const float* Wt = wsm;
  const float* b1 = wsm + Din * HID;
  const float* w2 = b1 + HID;
  float acc[HID];
#pragma unroll
  for (int k = 0; k < HID; ++k) acc[k] = b1[k];
  for (int f = 0; f < Din; ++f) {
    const float xv = x[f * 32 + lane];
#pragma unroll
    for (int k4 = 0; k4 < HID / 4; ++k4) {
      const float4 w = *reinterpret_cast<const float4*>(Wt + f * HID + 4 * k4);   // warp-uniform address: broadcast
      acc[4 * k4 + 0] = fmaf(xv, w.x, acc[4 * k4 + 0]);
      acc[4 * k4 + 1] = fmaf(xv, w.y, acc[4 * k4 + 1]);
      acc[4 * k4 + 2] = fmaf(xv, w.z, acc[4 * k4 + 2]);
      acc[4 * k4 + 3] = fmaf(xv, w.w, acc[4 * k4 + 3]);
    }
  }
  float out = w2[HID];   // b2
#pragma unroll
  for (int k = 0; k < HID; ++k) out = fmaf(mlp_act(acc[k], leaky), w2[k], out);
  return out;
}

template <int CPL, int HID>
__global__ void __launch_bounds__(kSdfWarps * 32) sdf_query_batch_kernel(NpMapDev m, NpQueryDev a, MlpDev mlp, float sdf_scale,
                                                                        long long N, const float* __restrict__ q,
                                                                        float* __restrict__ sdf, float* __restrict__ sdf_std,
                                                                        long long* __restrict__ nn_counts) {
  extern __shared__ __align__(16) float smem[];
  const int Din = mlp.in_dim;
  {   // stage: Wt1 [Din][HID] (transposed), b1 [HID], w2 [HID], b2
    for (int i = threadIdx.x; i < Din * HID; i += blockDim.x) {
      const int f = i / HID, u = i - f * HID;
      smem[i] = mlp.W[0][(size_t)u * Din + f];
    }
    for (int u = threadIdx.x; u < HID; u += blockDim.x) {
      smem[Din * HID + u] = mlp.B[0] ? mlp.B[0][u] : 0.f;
      smem[Din * HID + HID + u] = mlp.W[1][u];
    }
    if (threadIdx.x == 0) smem[Din * HID + 2 * HID] = mlp.B[1] ? mlp.B[1][0] : 0.f;
  }
  const size_t woff = ((size_t)Din * HID + 2 * HID + 1 + 3) & ~(size_t)3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nn_k = a.nn_k;
  float* ws = smem + woff + (size_t)warp * sdf_batch_warp_floats(mlp, nn_k);
  float* kd = ws;                                          // [32][nn_k] distances
  int* ki = reinterpret_cast<int*>(ws + 32 * nn_k);        // [32][nn_k] final indices
  float* ks = ws + 64 * nn_k;                              // [32][nn_k] per-neighbour predictions (!weighted_first)
  int* kc = reinterpret_cast<int*>(ws + 96 * nn_k);        // [32] neighbour counts
  float* xa = ws + 96 * nn_k + 32;                         // [Din][32] decoder input, lane-strided
  __syncthreads();
  const int F = a.F_geo, Fo = F + 3;
  long long dh[CPL];
  uint32_t dr[CPL];
  np_lane_deltas<CPL>(m, lane, dh, dr);
  const long long nbatch = (N + 31) / 32;
  for (long long bt = (long long)blockIdx.x * kSdfWarps + warp; bt < nbatch; bt += (long long)gridDim.x * kSdfWarps) {
    const long long n0 = bt * 32;
    const int nq = (int)min((long long)32, N - n0);
    // phase 1: warp-cooperative kNN of the batch's queries
    const float mqx = lane < nq ? q[3 * (n0 + lane)] : 0.f, mqy = lane < nq ? q[3 * (n0 + lane) + 1] : 0.f,
                mqz = lane < nq ? q[3 * (n0 + lane) + 2] : 0.f;
    for (int qi = 0; qi < nq; ++qi) {
      const float qx = __shfl_sync(0xffffffffu, mqx, qi), qy = __shfl_sync(0xffffffffu, mqy, qi),
                  qz = __shfl_sync(0xffffffffu, mqz, qi);
      float sel_d;
      int sel_i;
      const int cnt = np_warp_knn<CPL>(m, a, dh, dr, qx, qy, qz, lane, sel_d, sel_i);
      if (lane < nn_k) {
        kd[qi * nn_k + lane] = sel_d;
        ki[qi * nn_k + lane] = sel_i;
      }
      if (lane == 0) kc[qi] = cnt;
    }
    __syncwarp();
    // phase 2: lane = query
    if (lane < nq) {
      const long long n = n0 + lane;
      const int cnt = kc[lane];
      if (nn_counts) nn_counts[n] = cnt;
      // IDW weights (NP:642-662)
      float S = 0.f;
      for (int j = 0; j < nn_k; ++j) {
        float u = ki[lane * nn_k + j] >= 0 ? __fdiv_rn(1.0f, kd[lane * nn_k + j] + kNpEps) : 0.f;
        if (cnt == 0) u = kNpEps;
        S += u;
      }
      auto weight_of = [&](int j) {
        const int idx = ki[lane * nn_k + j];
        float u = idx >= 0 ? __fdiv_rn(1.0f, kd[lane * nn_k + j] + kNpEps) : 0.f;
        if (cnt == 0) u = kNpEps;
        return idx >= 0 ? __fdiv_rn(u, S) : 0.f;
      };
      if (a.weighted_first) {
        for (int f = 0; f < Fo; ++f) xa[f * 32 + lane] = 0.f;
        for (int j = 0; j < nn_k; ++j) {
          const int idx = ki[lane * nn_k + j];
          if (idx < 0) continue;
          const float wj = weight_of(j);
          const float* row = a.geo + (size_t)idx * F;
          for (int f = 0; f < F; ++f) xa[f * 32 + lane] = fmaf(row[f], wj, xa[f * 32 + lane]);
          const float3 nv = np_neighbor_vec(a, idx, mqx, mqy, mqz);
          xa[F * 32 + lane] = fmaf(nv.x, wj, xa[F * 32 + lane]);
          xa[(F + 1) * 32 + lane] = fmaf(nv.y, wj, xa[(F + 1) * 32 + lane]);
          xa[(F + 2) * 32 + lane] = fmaf(nv.z, wj, xa[(F + 2) * 32 + lane]);
        }
        sdf[n] = sdf_lane_mlp2<HID>(smem, Din, mlp.leaky != 0, xa, lane) * sdf_scale;
      } else {
        float mean = 0.f;
        for (int j = 0; j < nn_k; ++j) {
          const int idx = ki[lane * nn_k + j];
          for (int f = 0; f < Fo; ++f) xa[f * 32 + lane] = 0.f;
          if (idx >= 0) {
            const float* row = a.geo + (size_t)idx * F;
            for (int f = 0; f < F; ++f) xa[f * 32 + lane] = row[f];
            const float3 nv = np_neighbor_vec(a, idx, mqx, mqy, mqz);
            xa[F * 32 + lane] = nv.x;
            xa[(F + 1) * 32 + lane] = nv.y;
            xa[(F + 2) * 32 + lane] = nv.z;
          }
          const float sj = sdf_lane_mlp2<HID>(smem, Din, mlp.leaky != 0, xa, lane) * sdf_scale;
          const float wj = weight_of(j);
          mean = fmaf(sj, wj, mean);
          ks[lane * nn_k + j] = sj;   // keep the per-neighbour prediction for the std pass
        }
        sdf[n] = mean;
        if (sdf_std) {
          float var = 0.f;
          for (int j = 0; j < nn_k; ++j) {
            const float d = ks[lane * nn_k + j] - mean;
            var = fmaf(weight_of(j) * d, d, var);
          }
          sdf_std[n] = sqrtf(var);
        }
      }
    }
    __syncwarp();
  }
}

// ---- half-warp variant ---------------------------------------------------------------------------------------------
// Two queries per warp (16 lanes each): the probe phase has the same work per query (K = 81 probes = 6 per lane), but
// everything that is one instruction per WARP in the kernel above - the arg-min rounds (2 REDUX + bookkeeping per
// round), the IDW normalisation, the gather loop, the decoder - now serves two queries.  The decoder is specialised
// for the shape of Decoder.sdf with weighted_first (DEC:62-83: in = F + 3 <= 16, ONE hidden layer of 64, out = 1):
// lane s of a half owns hidden units s, s+16, s+32, s+48 with their 4 x 16 first-layer weights in REGISTERS for the
// life of the persistent warp, and the inputs arrive by shuffle (x_f lives in lane f), so a query's decoder is
// 16 SHFL + 64 FMA + 4 SHFL, no shared memory.  Same arithmetic order as the kernel above for the hidden layer
// (sequential over f, then the bias); the output sum is a 4-term partial per lane + 4 butterfly steps.
constexpr int kSdfHalfCpl = 6;      // probes per lane: K <= 96
constexpr int kSdfHalfIn = 16;      // in_dim = F_geo + 3 <= 16
constexpr int kSdfHalfHid = 64;
constexpr int kSdfHalfSlots = 16 * kSdfHalfCpl;   // probe slots of a half-warp

__host__ __device__ inline bool sdf_half_supported(const MlpDev& m, const NpQueryDev& a, int K) {
  return a.weighted_first && m.levels == 1 && m.hidden == kSdfHalfHid && m.out_dim == 1 && m.in_dim <= kSdfHalfIn &&
         m.in_dim == a.F_geo + 3 && K <= 16 * kSdfHalfCpl && a.nn_k <= 16;   // (+ n_points < 2^31, checked by the caller)
}

__global__ void __launch_bounds__(kSdfWarps * 32) sdf_query_half_kernel(NpMapDev m, NpQueryDev a, MlpDev mlp, float sdf_scale,
                                                                        long long N, const float* __restrict__ q,
                                                                        float* __restrict__ sdf,
                                                                        long long* __restrict__ nn_counts) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int s = lane & 15, hb = lane & 16;          // sub-lane, first lane of this half
  const uint32_t hmask = 0xffffu << hb;
  const int F = a.F_geo, Fo = F + 3, Din = mlp.in_dim;
  // first-layer weights in shared memory as float4 over this lane's 4 hidden units: w1s[f][s] = W1[s + 16 t][f],
  // t = 0..3 (zeros beyond Din: x_f is zero there too); consecutive lanes read consecutive 16-byte words.
  // (In registers - 64 per thread - the kernel ran at 2 CTAs per SM and lost to the probe latency: 427 vs 570 M q/s.)
  __shared__ float4 w1s[kSdfHalfIn * 16];
  // per half-warp scratch: occupied slots (index, probe position), surviving candidates (key, index), the nn_k kept
  __shared__ uint32_t sh_hit[2 * kSdfHalfSlots * kSdfWarps * 2];
  __shared__ unsigned long long sh_key[kSdfHalfSlots * kSdfWarps * 2];
  __shared__ uint32_t sh_val[kSdfHalfSlots * kSdfWarps * 2];
  __shared__ float sh_outd[16 * kSdfWarps * 2];
  __shared__ int sh_outi[16 * kSdfWarps * 2];
  for (int i = threadIdx.x; i < kSdfHalfIn * 16; i += blockDim.x) {
    const int f = i >> 4, ss = i & 15;
    float v[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) v[t] = f < Din ? mlp.W[0][(size_t)(ss + 16 * t) * Din + f] : 0.f;
    w1s[i] = make_float4(v[0], v[1], v[2], v[3]);
  }
  __syncthreads();
  float b1[4], w2[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const int u = s + 16 * t;
    b1[t] = mlp.B[0] ? mlp.B[0][u] : 0.f;
    w2[t] = mlp.W[1][u];
  }
  const float b2 = mlp.B[1] ? mlp.B[1][0] : 0.f;
  // per-lane slot residues of the probes c*16 + s (the 64-bit deltas are recomputed on the rare slow path)
  uint32_t dr[kSdfHalfCpl];
#pragma unroll
  for (int c = 0; c < kSdfHalfCpl; ++c) dr[c] = (c * 16 + s < m.K) ? (uint32_t)np_slot_of(m, np_hash_delta(m, c * 16 + s)) : 0u;
  const uint32_t d32 = (uint32_t)m.buffer_size;
  const long long pairs = (N + 1) / 2;
  for (long long p = (long long)blockIdx.x * kSdfWarps + warp; p < pairs; p += (long long)gridDim.x * kSdfWarps) {
    const long long n = 2 * p + (hb >> 4);
    const bool live = n < N;
    const long long nq = live ? n : N - 1;     // the idle half of an odd tail repeats the last query, stores nothing
    const float qx = q[3 * nq], qy = q[3 * nq + 1], qz = q[3 * nq + 2];
    // ---- phase 1: the slot-table probes (6 independent loads per lane), optional time filter ----
    const long long h0 = np_hash_base(m, np_cell(qx, m.res), np_cell(qy, m.res), np_cell(qz, m.res));
    const bool fast = m.buffer_size < (1ll << 31) && h0 > -(1ll << 62) && h0 < (1ll << 62);
    const uint32_t b0 = fast ? (uint32_t)np_slot_of(m, h0) : 0u;
    int pidx[kSdfHalfCpl];
#pragma unroll
    for (int c = 0; c < kSdfHalfCpl; ++c) {
      pidx[c] = -1;
      if (c * 16 + s < m.K) {
        long long slot;
        if (fast) {
          uint32_t sl = b0 + dr[c];
          if (sl >= d32) sl -= d32;
          slot = (long long)sl;
        } else {
          slot = np_slot_of(m, (long long)((unsigned long long)h0 + (unsigned long long)np_hash_delta(m, c * 16 + s)));
        }
        pidx[c] = m.table32 ? m.table32[slot] : (int)m.table[slot];
      }
    }
    if (m.time_filtering) {
#pragma unroll
      for (int c = 0; c < kSdfHalfCpl; ++c) {
        if (pidx[c] >= 0) {
          const float diff = fabsf(m.travel[m.cur_ts] - m.travel[m.ts_create[pidx[c]]]);
          if (!(diff < m.diff_local)) pidx[c] = -1;
        }
      }
    }
    // ---- phase 2: compact the occupied slots of each half (about 8 of 81) into shared memory, in probe order ----
    uint32_t* hit_i = sh_hit + (size_t)(warp * 2 + (hb >> 4)) * kSdfHalfSlots;
    uint32_t* hit_p = hit_i + kSdfHalfSlots * kSdfWarps * 2;
    uint32_t nhit = 0;
    const uint32_t below = (1u << s) - 1u;
#pragma unroll
    for (int c = 0; c < kSdfHalfCpl; ++c) {
      const bool hit = pidx[c] >= 0;
      const uint32_t bm = (__ballot_sync(0xffffffffu, hit) >> hb) & 0xffffu;
      if (hit) {
        const uint32_t r = nhit + __popc(bm & below);
        hit_i[r] = (uint32_t)pidx[c];
        hit_p[r] = (uint32_t)(c * 16 + s);
      }
      nhit += __popc(bm);
    }
    __syncwarp();
    // ---- phase 3: distances only for the occupied slots (np_probe_slot's arithmetic), masks, local index; the
    //      surviving candidates are compacted again as (dist bits << 32 | probe position, index) ----
    unsigned long long* ck = sh_key + (size_t)(warp * 2 + (hb >> 4)) * kSdfHalfSlots;
    uint32_t* cv = sh_val + (size_t)(warp * 2 + (hb >> 4)) * kSdfHalfSlots;
    uint32_t ncand = 0;
    const uint32_t nhit_w = max(nhit, __shfl_xor_sync(0xffffffffu, nhit, 16));
    for (uint32_t t0 = 0; t0 < nhit_w; t0 += 16) {
      const uint32_t t = t0 + s;
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
      const uint32_t bm = (__ballot_sync(0xffffffffu, ok) >> hb) & 0xffffu;
      if (ok) {
        const uint32_t r = ncand + __popc(bm & below);
        ck[r] = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned long long)pos;
        cv[r] = (uint32_t)idx;
      }
      ncand += __popc(bm);
    }
    __syncwarp();
    const int cnt = (int)ncand;
    // ---- phase 4: the nn_k nearest by rank counting: rank = number of candidates with a smaller (distance, probe
    //      position) key - the stable order of torch.sort over the probe axis (NP:565); keys are unique ----
    float* od = sh_outd + (size_t)(warp * 2 + (hb >> 4)) * 16;
    int* oi = sh_outi + (size_t)(warp * 2 + (hb >> 4)) * 16;
    const uint32_t ncand_w = max(ncand, __shfl_xor_sync(0xffffffffu, ncand, 16));
    for (uint32_t t0 = 0; t0 < ncand_w; t0 += 16) {
      const uint32_t t = t0 + s;
      if (t < ncand) {
        const unsigned long long key = ck[t];
        uint32_t rank = 0;
        for (uint32_t k = 0; k < ncand; ++k) rank += ck[k] < key ? 1u : 0u;
        if (rank < (uint32_t)a.nn_k) {
          od[rank] = __uint_as_float((uint32_t)(key >> 32));
          oi[rank] = (int)cv[t];
        }
      }
    }
    __syncwarp();
    float sel_d = kNpInvalidDist;
    int sel_i = -1;
    if (s < min(a.nn_k, cnt)) {
      sel_d = od[s];
      sel_i = oi[s];
    }
    __syncwarp();   // the next query of this warp overwrites the shared arrays
    // ---- IDW weights (np_idw_weight on 16 lanes; xor offsets < 16 stay inside the half) ----
    float u = 0.f;
    if (s < a.nn_k) {
      u = sel_i >= 0 ? __fdiv_rn(1.0f, sel_d + kNpEps) : 0.f;
      if (cnt == 0) u = kNpEps;
    }
    float us = u;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) us += __shfl_xor_sync(0xffffffffu, us, o);
    float w = s < a.nn_k ? __fdiv_rn(u, us) : 0.f;
    if (sel_i < 0) w = 0.f;
    if (s == 0 && live && nn_counts) nn_counts[n] = cnt;
    // ---- weighted feature vector: x_f in lane f (f < F: feature, then the neighbour vector) ----
    float3 nv = make_float3(0.f, 0.f, 0.f);
    if (sel_i >= 0) nv = np_neighbor_vec(a, sel_i, qx, qy, qz);   // lane j: its own neighbour, once
    float x = 0.f;
    for (int j = 0; j < a.nn_k; ++j) {
      const int idx = __shfl_sync(0xffffffffu, sel_i, hb + j);
      const float wj = __shfl_sync(0xffffffffu, w, hb + j);
      const float vx = __shfl_sync(0xffffffffu, nv.x, hb + j);
      const float vy = __shfl_sync(0xffffffffu, nv.y, hb + j);
      const float vz = __shfl_sync(0xffffffffu, nv.z, hb + j);
      if (idx < 0 || s >= Fo) continue;
      const float v = s < F ? a.geo[(size_t)idx * F + s] : (s == F ? vx : (s == F + 1 ? vy : vz));
      x += v * wj;
    }
    // ---- decoder: 16 -> 64 -> 1 ----
    float h[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int f = 0; f < kSdfHalfIn; ++f) {
      const float xf = __shfl_sync(0xffffffffu, x, hb + f);
      const float4 wv = w1s[f * 16 + s];
      h[0] += xf * wv.x;
      h[1] += xf * wv.y;
      h[2] += xf * wv.z;
      h[3] += xf * wv.w;
    }
    float part = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) part += mlp_act(h[t] + b1[t], mlp.leaky) * w2[t];
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (s == 0 && live) sdf[n] = (part + b2) * sdf_scale;
  }
}

}  // namespace pg
