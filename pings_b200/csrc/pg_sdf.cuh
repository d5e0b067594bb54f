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

}  // namespace pg
