// Decoder MLP (DEC = model/decoder.py:14-104): hidden_level x (Linear + ReLU|LeakyReLU) + Linear.
//
// fp32 SIMT path.  One CTA owns a tile of 64 rows; activations ping-pong between two shared
// buffers so that the whole layer chain runs without touching HBM in between (the reference
// launches one cuBLAS sgemm + one bias/ReLU kernel per layer and round-trips [N,H] each time).
// Each layer's weight matrix is staged transposed in shared memory; a thread computes a
// 4 row x 4 unit register tile (2 LDS.128-equivalent reads per 16 FMAs).
#pragma once
#include "pg_common.cuh"

namespace pg {

constexpr int kMlpRows = 64;      // rows per CTA tile
constexpr int kMlpThreads = 256;  // 16 row groups (4 rows each) x 16 unit lanes

struct MlpDev {
  int in_dim, hidden, levels, out_dim, leaky;
  const float* W[PG_MLP_MAX_LAYERS];
  const float* B[PG_MLP_MAX_LAYERS];
};

__host__ __device__ inline int mlp_layer_in(const MlpDev& m, int l) { return l == 0 ? m.in_dim : m.hidden; }
__host__ __device__ inline int mlp_layer_out(const MlpDev& m, int l) { return l == m.levels ? m.out_dim : m.hidden; }
__host__ __device__ inline int mlp_max_dim(const MlpDev& m) {
  int d = m.in_dim > m.hidden ? m.in_dim : m.hidden;
  return d > m.out_dim ? d : m.out_dim;
}
// shared floats: two activation buffers [64][Dmax+1] + transposed weights [Din][Dout_pad]
__host__ __device__ inline size_t mlp_smem_floats(const MlpDev& m) {
  const int D = mlp_max_dim(m);
  size_t wmax = 0;
  for (int l = 0; l <= m.levels; ++l) {
    size_t w = (size_t)mlp_layer_in(m, l) * ((mlp_layer_out(m, l) + 3) / 4 * 4);
    if (w > wmax) wmax = w;
  }
  return 2 * (size_t)kMlpRows * (D + 1) + wmax + 4 * 64;
}

__device__ __forceinline__ float mlp_act(float v, int leaky) { return v > 0.f ? v : (leaky ? 0.01f * v : 0.f); }

// One dense layer on the CTA's tile: out[r][u] = act(b[u] + sum_f in[r][f] * W[u][f]).
// `in`/`out` are [64][ld] shared buffers, Wt the staged transposed weights [Din][Dp].
__device__ __forceinline__ void mlp_layer_tile(const float* __restrict__ in, float* __restrict__ out, int ld,
                                               const float* __restrict__ Wg, const float* __restrict__ Bg,
                                               float* __restrict__ Wt, int Din, int Dout, bool act, int leaky) {
  const int Dp = (Dout + 3) / 4 * 4;
  for (int i = threadIdx.x; i < Din * Dp; i += kMlpThreads) {
    const int f = i / Dp, u = i - f * Dp;
    Wt[i] = u < Dout ? Wg[(size_t)u * Din + f] : 0.f;
  }
  __syncthreads();
  const int rg = threadIdx.x >> 4, ul = threadIdx.x & 15;  // rows rg + 16*i; units (ul + 16*t)*4 .. +3
  for (int u0 = ul * 4; u0 < Dp; u0 += 64) {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int f = 0; f < Din; ++f) {
      const float4 w = *reinterpret_cast<const float4*>(&Wt[f * Dp + u0]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float x = in[(rg + 16 * i) * ld + f];
        acc[i][0] += x * w.x; acc[i][1] += x * w.y; acc[i][2] += x * w.z; acc[i][3] += x * w.w;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int u = u0 + j;
      if (u < Dout) {
        const float b = Bg ? Bg[u] : 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float v = acc[i][j] + b;
          out[(rg + 16 * i) * ld + u] = act ? mlp_act(v, leaky) : v;
        }
      }
    }
  }
  __syncthreads();
}

// y = mlp(x); optional hidden_save[l][N][H] (post-activation) for the backward.
__global__ void __launch_bounds__(kMlpThreads) mlp_forward_kernel(MlpDev m, long long N, const float* __restrict__ x,
                                                                  float* __restrict__ y, float* __restrict__ hsave) {
  extern __shared__ float smem[];
  const int D = mlp_max_dim(m), ld = D + 1;
  float* bufA = smem;
  float* bufB = smem + kMlpRows * ld;
  float* Wt = bufB + kMlpRows * ld;
  const long long r0 = (long long)blockIdx.x * kMlpRows;
  for (int i = threadIdx.x; i < kMlpRows * m.in_dim; i += kMlpThreads) {
    const int r = i / m.in_dim, f = i - r * m.in_dim;
    bufA[r * ld + f] = (r0 + r < N) ? x[(r0 + r) * m.in_dim + f] : 0.f;
  }
  __syncthreads();
  float* cur = bufA;
  float* nxt = bufB;
  for (int l = 0; l <= m.levels; ++l) {
    const int Din = mlp_layer_in(m, l), Dout = mlp_layer_out(m, l);
    mlp_layer_tile(cur, nxt, ld, m.W[l], m.B[l], Wt, Din, Dout, l < m.levels, m.leaky);
    if (l < m.levels && hsave) {
      float* hs = hsave + (size_t)l * N * m.hidden;
      for (int i = threadIdx.x; i < kMlpRows * Dout; i += kMlpThreads) {
        const int r = i / Dout, u = i - r * Dout;
        if (r0 + r < N) hs[(r0 + r) * Dout + u] = nxt[r * ld + u];
      }
    }
    float* t = cur; cur = nxt; nxt = t;
  }
  for (int i = threadIdx.x; i < kMlpRows * m.out_dim; i += kMlpThreads) {
    const int r = i / m.out_dim, u = i - r * m.out_dim;
    if (r0 + r < N) y[(r0 + r) * m.out_dim + u] = cur[r * ld + u];
  }
}

// ---- backward ----------------------------------------------------------------------------------
// d_in[r][f] = sum_u g[r][u] * W[u][f]   (same register tile as the forward, W staged as is)
__device__ __forceinline__ void mlp_dgrad_tile(const float* __restrict__ g, float* __restrict__ out, int ld,
                                               const float* __restrict__ Wg, float* __restrict__ Ws, int Din, int Dout) {
  const int Dp = (Din + 3) / 4 * 4;
  // warp per row, lanes along the row: no per-element integer division (it was about half of this kernel's instructions)
  for (int u = threadIdx.x >> 5; u < Dout; u += kMlpThreads / 32)
    for (int f = threadIdx.x & 31; f < Dp; f += 32) Ws[u * Dp + f] = f < Din ? Wg[(size_t)u * Din + f] : 0.f;
  __syncthreads();
  const int rg = threadIdx.x >> 4, fl = threadIdx.x & 15;
  for (int f0 = fl * 4; f0 < Dp; f0 += 64) {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int u = 0; u < Dout; ++u) {
      const float4 w = *reinterpret_cast<const float4*>(&Ws[u * Dp + f0]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float x = g[(rg + 16 * i) * ld + u];
        acc[i][0] += x * w.x; acc[i][1] += x * w.y; acc[i][2] += x * w.z; acc[i][3] += x * w.w;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (f0 + j < Din) {
#pragma unroll
        for (int i = 0; i < 4; ++i) out[(rg + 16 * i) * ld + f0 + j] = acc[i][j];
      }
  }
  __syncthreads();
}

// Shared floats for the backward: G, IN, DIN buffers + weights staged [Dout][Din_pad].
__host__ __device__ inline size_t mlp_bwd_smem_floats(const MlpDev& m) {
  const int D = mlp_max_dim(m);
  size_t wmax = 0;
  for (int l = 0; l <= m.levels; ++l) {
    size_t w = (size_t)mlp_layer_out(m, l) * ((mlp_layer_in(m, l) + 3) / 4 * 4);
    if (w > wmax) wmax = w;
  }
  return 3 * (size_t)kMlpRows * (D + 1) + wmax + 4 * 64;
}

struct MlpGradDev {
  float* dW[PG_MLP_MAX_LAYERS];
  float* dB[PG_MLP_MAX_LAYERS];
};

__global__ void __launch_bounds__(kMlpThreads) mlp_backward_kernel(MlpDev m, MlpGradDev gr, long long N,
                                                                   const float* __restrict__ x,
                                                                   const float* __restrict__ hsave,
                                                                   const float* __restrict__ dY, float* __restrict__ dX) {
  extern __shared__ float smem[];
  const int D = mlp_max_dim(m), ld = D + 1;
  float* G = smem;                        // gradient w.r.t. the pre-activation output of layer l
  float* IN = G + kMlpRows * ld;          // input of layer l
  float* DIN = IN + kMlpRows * ld;        // gradient w.r.t. that input
  float* Ws = DIN + kMlpRows * ld;
  const long long r0 = (long long)blockIdx.x * kMlpRows;
  const int wrp = threadIdx.x >> 5, lan = threadIdx.x & 31;
  for (int r = wrp; r < kMlpRows; r += kMlpThreads / 32)
    for (int u = lan; u < m.out_dim; u += 32) G[r * ld + u] = (r0 + r < N) ? dY[(r0 + r) * m.out_dim + u] : 0.f;
  for (int l = m.levels; l >= 0; --l) {
    const int Din = mlp_layer_in(m, l), Dout = mlp_layer_out(m, l);
    const float* src = l == 0 ? x : hsave + (size_t)(l - 1) * N * m.hidden;
    for (int r = wrp; r < kMlpRows; r += kMlpThreads / 32)
      for (int f = lan; f < Din; f += 32) IN[r * ld + f] = (r0 + r < N) ? src[(r0 + r) * Din + f] : 0.f;
    __syncthreads();
    // weight / bias gradients of this tile -> global accumulation.  Register tile: thread (ul, fl) of a 16 x 16
    // arrangement owns outputs u = ul + 16 a, inputs f = fl + 16 b, four of each at a time: 8 shared loads per
    // 16 independent FMAs (the one-accumulator-per-entry loop it replaces issued 2 loads per dependent FMA and
    // ran at ~5 % of the FP32 pipe with the 8 warps per SM this kernel's shared memory allows).
    {
      const int ul = threadIdx.x >> 4, fl = threadIdx.x & 15;
      for (int a0 = 0; a0 * 16 < Dout; a0 += 4) {
        for (int b0 = 0; b0 * 16 < Din; b0 += 4) {
          int uo[4], fo[4];
          bool uv[4], fv[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int u = ul + 16 * (a0 + q), f = fl + 16 * (b0 + q);
            uv[q] = u < Dout; fv[q] = f < Din;
            uo[q] = uv[q] ? u : 0; fo[q] = fv[q] ? f : 0;   // clamped: in range, masked at the store
          }
          float acc[4][4];
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 4
          for (int r = 0; r < kMlpRows; ++r) {
            const float* gr_ = G + r * ld;
            const float* ir_ = IN + r * ld;
            float gv[4], iv[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) { gv[q] = gr_[uo[q]]; iv[q] = ir_[fo[q]]; }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
              for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(gv[a], iv[b], acc[a][b]);
          }
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b)
              if (uv[a] && fv[b]) atomicAdd(&gr.dW[l][(size_t)uo[a] * Din + fo[b]], acc[a][b]);
        }
      }
    }
    if (gr.dB[l]) {
      for (int u = threadIdx.x; u < Dout; u += kMlpThreads) {
        float s = 0.f;
        for (int r = 0; r < kMlpRows; ++r) s += G[r * ld + u];
        atomicAdd(&gr.dB[l][u], s);
      }
    }
    if (l == 0 && !dX) break;
    mlp_dgrad_tile(G, DIN, ld, m.W[l], Ws, Din, Dout);
    if (l > 0) {
      // through the activation of layer l-1 (its post-activation output is IN)
      for (int r = wrp; r < kMlpRows; r += kMlpThreads / 32)
        for (int f = lan; f < Din; f += 32) {
          const float h = IN[r * ld + f];
          G[r * ld + f] = DIN[r * ld + f] * (h > 0.f ? 1.f : (m.leaky ? 0.01f : 0.f));
        }
      __syncthreads();
    } else {
      for (int r = wrp; r < kMlpRows; r += kMlpThreads / 32)
        for (int f = lan; f < Din; f += 32)
          if (r0 + r < N) dX[(r0 + r) * Din + f] = DIN[r * ld + f];
    }
  }
}

}  // namespace pg
