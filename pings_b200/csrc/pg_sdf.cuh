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
  np_lane_deltas<CPL>(m, lane, dh);
  for (long long n = (long long)blockIdx.x * kSdfWarps + warp; n < N; n += (long long)gridDim.x * kSdfWarps) {
    const float qx = q[3 * n], qy = q[3 * n + 1], qz = q[3 * n + 2];
    float sel_d;
    int sel_i;
    const int cnt = np_warp_knn<CPL>(m, a, dh, qx, qy, qz, lane, sel_d, sel_i);
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

}  // namespace pg
