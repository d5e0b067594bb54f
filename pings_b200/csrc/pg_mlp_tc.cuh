// Decoder MLP on the 5th-gen tensor cores (tcgen05 + TMEM), for the large-batch `mlp_batch` calls
// (DEC = model/decoder.py:84-98; spawn_gaussians runs 5 of them per neural point, GR:609-714).
//
// One CTA (128 threads) owns a 128-row tile and runs the whole layer chain on it:
//   * operands live in shared memory in the no-swizzle K-major canonical layout
//     [K/4 chunks][rows][4 floats] (a core matrix = 8 rows x 16 B contiguous; SBO = 128 B between
//     row groups, LBO = rows*16 B between the two 16-B K chunks of one K=8 MMA step);
//   * fp32 accuracy comes from the 3xTF32 split: every operand is stored as hi (low 13 mantissa
//     bits cleared, exactly what kind::tf32 reads) and lo = v - hi, and each K step issues
//     hi*hi + lo*hi + hi*lo into the same fp32 TMEM accumulator (lo*lo < 2^-22 relative is dropped);
//   * one elected thread issues tcgen05.mma (M=128, N = piece of <= 64 output units, K=8 per
//     instruction) and commits to an mbarrier; weights are streamed per layer in 64-unit pieces so
//     hidden = 128 fits (2 x 64 KB activations + 2 x 32 KB weights);
//   * epilogue: thread t reads accumulator row t with tcgen05.ld (32x32b: lane = row), adds the
//     bias, applies ReLU/LeakyReLU and writes the next layer's hi/lo operands straight back in the
//     canonical layout (consecutive threads -> consecutive 16 B: conflict free), so activations never
//     leave the SM between layers.
#pragma once
#include "pg_mlp.cuh"

namespace pg {

constexpr int kTcRows = 128;
constexpr int kTcThreads = 128;
constexpr int kTcPiece = 64;     // output units per weight piece / MMA N
constexpr int kTcMaxDim = 128;   // in_dim, hidden <= 128

__host__ __device__ inline int tc_round8(int v) { return (v + 7) & ~7; }
__host__ __device__ inline int tc_round16(int v) { return (v + 15) & ~15; }
__host__ __device__ inline int tc_kmax(const MlpDev& m) {
  const int a = tc_round8(m.in_dim), b = tc_round16(m.hidden);   // the epilogue writes 16 columns at a time
  return a > b ? a : b;
}
__host__ __device__ inline int tc_tmem_cols(const MlpDev& m) {
  int need = tc_round16(m.hidden);
  const int o = tc_round16(m.out_dim);
  if (o > need) need = o;
  int c = 32;
  while (c < need) c <<= 1;
  return c;
}
// bytes: A hi/lo [Kmax/4][128][16 B] x2, W hi/lo [Kmax/4][64][16 B] x2, mbarrier + tmem slot
__host__ __device__ inline size_t tc_smem_bytes(const MlpDev& m) {
  const size_t k = (size_t)tc_kmax(m);
  return 2 * k * kTcRows * 4 + 2 * k * kTcPiece * 4 + (size_t)PG_MLP_MAX_LAYERS * kTcMaxDim * 4 + 64;
}
inline bool tc_supported(const MlpDev& m) {
  return m.in_dim <= kTcMaxDim && m.hidden <= kTcMaxDim && m.out_dim <= kTcPiece && tc_smem_bytes(m) <= 227 * 1024;
}

__device__ __forceinline__ uint32_t tc_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_NONE shared-memory matrix descriptor (sm_100 version field = 1).
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fffu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// kind::tf32, fp32 accumulate, A and B K-major, M = 128
__device__ __forceinline__ uint32_t tc_idesc(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(a), "l"(b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t mbar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void tc_mbar_wait(uint32_t mbar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(ok)
        : "r"(mbar), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, "
      "[%16];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_split(float v, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(v) & 0xffffe000u);
  lo = v - hi;
}

template <int COLS>
__global__ void __launch_bounds__(kTcThreads) mlp_tc_forward_kernel(MlpDev m, long long N, const float* __restrict__ x,
                                                                    float* __restrict__ y, float* __restrict__ hsave) {
  extern __shared__ __align__(128) unsigned char tc_sm[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int kmax = tc_kmax(m);
  float* A_hi = reinterpret_cast<float*>(tc_sm);
  float* A_lo = A_hi + (size_t)kmax * kTcRows;
  float* W_hi = A_lo + (size_t)kmax * kTcRows;
  float* W_lo = W_hi + (size_t)kmax * kTcPiece;
  float* bias_s = W_lo + (size_t)kmax * kTcPiece;   // [layer][128], zero padded (also when the layer has no bias)
  uint64_t* mbar = reinterpret_cast<uint64_t*>(bias_s + PG_MLP_MAX_LAYERS * kTcMaxDim);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)),
                 "n"(COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < (m.levels + 1) * kTcMaxDim; i += kTcThreads) {
    const int l = i / kTcMaxDim, c = i % kTcMaxDim;
    bias_s[i] = (m.B[l] && c < mlp_layer_out(m, l)) ? __ldg(m.B[l] + c) : 0.f;
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tc_smem_u32(mbar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  const uint32_t mbar_a = tc_smem_u32(mbar);

  // stage the input tile: the tile's rows are one contiguous block of x, read fully coalesced
  const long long row0 = (long long)blockIdx.x * kTcRows;
  const long long row = row0 + tid;
  {
    const int Din = m.in_dim, kp = tc_round8(Din);
    const int total = kTcRows * Din;
    const long long left = (N - row0) * Din;
    const int valid = left < total ? (int)left : total;
    const float* xb = x + row0 * Din;
    const float inv_din = 1.0f / (float)Din;   // (i + 0.5) * inv floors exactly for i < 2^14, Din <= 128
    for (int i0 = 0; i0 < total; i0 += kTcThreads * 8) {
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i = i0 + j * kTcThreads + tid;
        v[j] = i < valid ? __ldg(xb + i) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i = i0 + j * kTcThreads + tid;
        if (i < total) {
          const int r = (int)(((float)i + 0.5f) * inv_din), k = i - r * Din;
          float h, lo;
          tc_split(v[j], h, lo);
          const size_t o = ((size_t)(k >> 2) * kTcRows + r) * 4 + (k & 3);
          A_hi[o] = h;
          A_lo[o] = lo;
        }
      }
    }
    for (int k = Din; k < kp; ++k) {   // K padding of the first layer
      const size_t o = ((size_t)(k >> 2) * kTcRows + tid) * 4 + (k & 3);
      A_hi[o] = 0.f;
      A_lo[o] = 0.f;
    }
  }
  uint32_t phase = 0;
  for (int l = 0; l <= m.levels; ++l) {
    const int Din = mlp_layer_in(m, l), Dout = mlp_layer_out(m, l);
    const int kp = tc_round8(Din);
    const float* Wg = m.W[l];
    for (int n0 = 0; n0 < Dout; n0 += kTcPiece) {
      const int np = min(kTcPiece, Dout - n0), NP = tc_round16(np);
      // weight piece -> [kp/4][NP][4] hi / lo   (W is [Dout][Din] row-major = K-major already);
      // loads are issued four chunks at a time so their latencies overlap
      const int chunks = NP * (kp >> 2);
      const float inv_np = 1.0f / (float)NP;
      const bool vec = (Din & 3) == 0 && (reinterpret_cast<uintptr_t>(Wg) & 15) == 0;
      for (int i0 = 0; i0 < chunks; i0 += kTcThreads * 4) {
        float4 v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int i = i0 + q * kTcThreads + tid;
          v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (i < chunks) {
            const int kc = (int)(((float)i + 0.5f) * inv_np), n = i - kc * NP;
            if (n < np) {
              const float* wr = Wg + (size_t)(n0 + n) * Din + kc * 4;
              if (vec) {
                if (kc * 4 < Din) v[q] = __ldg(reinterpret_cast<const float4*>(wr));
              } else {
                if (kc * 4 + 0 < Din) v[q].x = __ldg(wr + 0);
                if (kc * 4 + 1 < Din) v[q].y = __ldg(wr + 1);
                if (kc * 4 + 2 < Din) v[q].z = __ldg(wr + 2);
                if (kc * 4 + 3 < Din) v[q].w = __ldg(wr + 3);
              }
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int i = i0 + q * kTcThreads + tid;
          if (i < chunks) {
            const int kc = (int)(((float)i + 0.5f) * inv_np), n = i - kc * NP;
            float4 h, lo4;
            tc_split(v[q].x, h.x, lo4.x); tc_split(v[q].y, h.y, lo4.y);
            tc_split(v[q].z, h.z, lo4.z); tc_split(v[q].w, h.w, lo4.w);
            const size_t o = ((size_t)kc * NP + n) * 4;
            *reinterpret_cast<float4*>(W_hi + o) = h;
            *reinterpret_cast<float4*>(W_lo + o) = lo4;
          }
        }
      }
      // generic-proxy shared writes (A from the previous epilogue, W) -> visible to the tensor core
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncthreads();
      if (tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t idesc = tc_idesc(NP);
        const uint32_t a_hi = tc_smem_u32(A_hi), a_lo = tc_smem_u32(A_lo);
        const uint32_t w_hi = tc_smem_u32(W_hi), w_lo = tc_smem_u32(W_lo);
        const uint32_t a_lbo = kTcRows * 16, w_lbo = (uint32_t)NP * 16;
        for (int ks = 0; ks < (kp >> 3); ++ks) {
          const uint32_t ao = (uint32_t)ks * 2u * a_lbo, wo = (uint32_t)ks * 2u * w_lbo;
          const uint64_t dah = tc_desc(a_hi + ao, a_lbo, 128), dal = tc_desc(a_lo + ao, a_lbo, 128);
          const uint64_t dwh = tc_desc(w_hi + wo, w_lbo, 128), dwl = tc_desc(w_lo + wo, w_lbo, 128);
          tc_mma(tmem + (uint32_t)n0, dah, dwh, idesc, ks > 0 ? 1u : 0u);
          tc_mma(tmem + (uint32_t)n0, dal, dwh, idesc, 1u);
          tc_mma(tmem + (uint32_t)n0, dah, dwl, idesc, 1u);
        }
        tc_commit(mbar_a);
      }
      tc_mbar_wait(mbar_a, phase);
      phase ^= 1u;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    // epilogue: thread t <- accumulator row t (TMEM lane 32*warp + lane).  Padded columns need no
    // guard: their weights and bias are zero, so they come out as act(0) = 0.
    const bool last = l == m.levels;
    const int leaky = m.leaky;
    const int cols = tc_round16(Dout);
    const float* bl = bias_s + l * kTcMaxDim;
    float* hs = (!last && hsave && row < N) ? hsave + ((size_t)l * N + row) * Dout : nullptr;
    float* yr = (last && row < N) ? y + row * Dout : nullptr;
    for (int c0 = 0; c0 < cols; c0 += 16) {
      uint32_t r[16];
      tc_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
      float v[16];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 b4 = *reinterpret_cast<const float4*>(bl + c0 + 4 * q);
        v[4 * q + 0] = __uint_as_float(r[4 * q + 0]) + b4.x;
        v[4 * q + 1] = __uint_as_float(r[4 * q + 1]) + b4.y;
        v[4 * q + 2] = __uint_as_float(r[4 * q + 2]) + b4.z;
        v[4 * q + 3] = __uint_as_float(r[4 * q + 3]) + b4.w;
      }
      if (last) {
        if (yr) {
          if (c0 + 16 <= Dout) {
#pragma unroll
            for (int j = 0; j < 16; ++j) yr[c0 + j] = v[j];
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j < Dout) yr[c0 + j] = v[j];
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = mlp_act(v[j], leaky);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float4 h, lo4;
          tc_split(v[4 * q + 0], h.x, lo4.x); tc_split(v[4 * q + 1], h.y, lo4.y);
          tc_split(v[4 * q + 2], h.z, lo4.z); tc_split(v[4 * q + 3], h.w, lo4.w);
          const size_t o = ((size_t)((c0 >> 2) + q) * kTcRows + tid) * 4;
          *reinterpret_cast<float4*>(A_hi + o) = h;
          *reinterpret_cast<float4*>(A_lo + o) = lo4;
        }
        if (hs) {
          if (c0 + 16 <= Dout) {
#pragma unroll
            for (int j = 0; j < 16; ++j) hs[c0 + j] = v[j];
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j < Dout) hs[c0 + j] = v[j];
          }
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(COLS) : "memory");
  }
}

// ---- v2: persistent CTAs, resident weights, two tiles in flight ------------------------------------
// When all layers' split weights fit next to two activation tiles (hidden <= 64), a CTA of 256 threads
// stages the weights ONCE, then its two 4-warp groups each stride over 128-row tiles on their own
// (own activation buffers, own TMEM columns, own mbarrier, named barriers instead of __syncthreads):
// while one group runs its register epilogue the other group's MMA or input loads are in flight.
constexpr int kTc2Threads = 256;

// float offset of layer l's resident weights (hi; lo follows at +tc2_w_floats) and total per half
__host__ __device__ inline size_t tc2_w_offset(const MlpDev& m, int layer) {
  size_t o = 0;
  for (int l = 0; l < layer; ++l) o += (size_t)tc_round16(mlp_layer_out(m, l)) * tc_round8(mlp_layer_in(m, l));
  return o;
}
__host__ __device__ inline size_t tc2_w_floats(const MlpDev& m) { return tc2_w_offset(m, m.levels + 1); }
__host__ __device__ inline size_t tc2_smem_bytes(const MlpDev& m) {
  const size_t k = (size_t)tc_kmax(m);
  return 2 * tc2_w_floats(m) * 4 + 2 * 2 * k * kTcRows * 4 + (size_t)PG_MLP_MAX_LAYERS * kTcMaxDim * 4 + 64;
}
inline bool tc2_supported(const MlpDev& m) {
  return tc_supported(m) && m.hidden <= kTcPiece && tc_tmem_cols(m) <= 64 && tc2_smem_bytes(m) <= 227 * 1024;
}

__device__ __forceinline__ void tc2_group_sync(int g) {
  asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(kTcThreads) : "memory");
}

template <int COLS>
__global__ void __launch_bounds__(kTc2Threads) mlp_tc2_forward_kernel(MlpDev m, long long N, const float* __restrict__ x,
                                                                      float* __restrict__ y, float* __restrict__ hsave) {
  extern __shared__ __align__(128) unsigned char tc_sm[];
  const int tid = threadIdx.x, g = tid >> 7, t = tid & 127, wg = (tid >> 5) & 3;
  const int kmax = tc_kmax(m);
  const size_t wfl = tc2_w_floats(m);
  float* Wr_hi = reinterpret_cast<float*>(tc_sm);
  float* Wr_lo = Wr_hi + wfl;
  float* A_hi = Wr_lo + wfl + (size_t)g * 2 * kmax * kTcRows;
  float* A_lo = A_hi + (size_t)kmax * kTcRows;
  float* bias_s = Wr_lo + wfl + (size_t)4 * kmax * kTcRows;
  uint64_t* mbar = reinterpret_cast<uint64_t*>(bias_s + PG_MLP_MAX_LAYERS * kTcMaxDim);   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 2);

  if ((tid >> 5) == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc_smem_u32(tmem_slot)),
                 "n"(2 * COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = tid; i < (m.levels + 1) * kTcMaxDim; i += kTc2Threads) {
    const int l = i / kTcMaxDim, c = i % kTcMaxDim;
    bias_s[i] = (m.B[l] && c < mlp_layer_out(m, l)) ? __ldg(m.B[l] + c) : 0.f;
  }
  if (tid < 2) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tc_smem_u32(mbar + tid)) : "memory");
  if (tid == 0) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  // resident weights: every layer as [kp/4][NP][4] hi / lo (one piece per layer: Dout <= 64)
  for (int l = 0; l <= m.levels; ++l) {
    const int Din = mlp_layer_in(m, l), Dout = mlp_layer_out(m, l);
    const int kp = tc_round8(Din), NP = tc_round16(Dout);
    const float* Wg = m.W[l];
    float* wh = Wr_hi + tc2_w_offset(m, l);
    float* wl = Wr_lo + tc2_w_offset(m, l);
    const int chunks = NP * (kp >> 2);
    const float inv_np = 1.0f / (float)NP;
    for (int i = tid; i < chunks; i += kTc2Threads) {
      const int kc = (int)(((float)i + 0.5f) * inv_np), n = i - kc * NP;
      float v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = kc * 4 + j;
        v[j] = (n < Dout && k < Din) ? __ldg(Wg + (size_t)n * Din + k) : 0.f;
      }
      float4 h, lo4;
      tc_split(v[0], h.x, lo4.x); tc_split(v[1], h.y, lo4.y); tc_split(v[2], h.z, lo4.z); tc_split(v[3], h.w, lo4.w);
      *reinterpret_cast<float4*>(wh + (size_t)i * 4) = h;    // i == kc * NP + n
      *reinterpret_cast<float4*>(wl + (size_t)i * 4) = lo4;
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot + (uint32_t)(g * COLS);
  const uint32_t mbar_a = tc_smem_u32(mbar + g);
  const int leaky = m.leaky;
  const long long tiles = (N + kTcRows - 1) / kTcRows;
  uint32_t phase = 0;
  for (long long tile = (long long)blockIdx.x * 2 + g; tile < tiles; tile += (long long)gridDim.x * 2) {
    const long long row0 = tile * kTcRows, row = row0 + t;
    {
      const int Din = m.in_dim, kp = tc_round8(Din);
      const int total = kTcRows * Din;
      const long long left = (N - row0) * Din;
      const int valid = left < total ? (int)left : total;
      const float* xb = x + row0 * Din;
      const float inv_din = 1.0f / (float)Din;
      for (int i0 = 0; i0 < total; i0 += kTcThreads * 8) {
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int i = i0 + j * kTcThreads + t;
          v[j] = i < valid ? __ldg(xb + i) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int i = i0 + j * kTcThreads + t;
          if (i < total) {
            const int r = (int)(((float)i + 0.5f) * inv_din), k = i - r * Din;
            float h, lo;
            tc_split(v[j], h, lo);
            const size_t o = ((size_t)(k >> 2) * kTcRows + r) * 4 + (k & 3);
            A_hi[o] = h;
            A_lo[o] = lo;
          }
        }
      }
      for (int k = Din; k < kp; ++k) {
        const size_t o = ((size_t)(k >> 2) * kTcRows + t) * 4 + (k & 3);
        A_hi[o] = 0.f;
        A_lo[o] = 0.f;
      }
    }
    for (int l = 0; l <= m.levels; ++l) {
      const int Din = mlp_layer_in(m, l), Dout = mlp_layer_out(m, l);
      const int kp = tc_round8(Din), NP = tc_round16(Dout);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      tc2_group_sync(g);
      if (t == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t idesc = tc_idesc(NP);
        const uint32_t a_hi = tc_smem_u32(A_hi), a_lo = tc_smem_u32(A_lo);
        const uint32_t w_hi = tc_smem_u32(Wr_hi + tc2_w_offset(m, l)), w_lo = tc_smem_u32(Wr_lo + tc2_w_offset(m, l));
        const uint32_t a_lbo = kTcRows * 16, w_lbo = (uint32_t)NP * 16;
        for (int ks = 0; ks < (kp >> 3); ++ks) {
          const uint32_t ao = (uint32_t)ks * 2u * a_lbo, wo = (uint32_t)ks * 2u * w_lbo;
          const uint64_t dah = tc_desc(a_hi + ao, a_lbo, 128), dal = tc_desc(a_lo + ao, a_lbo, 128);
          const uint64_t dwh = tc_desc(w_hi + wo, w_lbo, 128), dwl = tc_desc(w_lo + wo, w_lbo, 128);
          tc_mma(tmem, dah, dwh, idesc, ks > 0 ? 1u : 0u);
          tc_mma(tmem, dal, dwh, idesc, 1u);
          tc_mma(tmem, dah, dwl, idesc, 1u);
        }
        tc_commit(mbar_a);
      }
      tc_mbar_wait(mbar_a, phase);
      phase ^= 1u;
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const bool last = l == m.levels;
      const int cols = NP;
      const float* bl = bias_s + l * kTcMaxDim;
      float* hs = (!last && hsave && row < N) ? hsave + ((size_t)l * N + row) * Dout : nullptr;
      float* yr = (last && row < N) ? y + row * Dout : nullptr;
      for (int c0 = 0; c0 < cols; c0 += 16) {
        uint32_t r[16];
        tc_ld16(tmem + ((uint32_t)(wg * 32) << 16) + (uint32_t)c0, r);
        float v[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 b4 = *reinterpret_cast<const float4*>(bl + c0 + 4 * q);
          v[4 * q + 0] = __uint_as_float(r[4 * q + 0]) + b4.x;
          v[4 * q + 1] = __uint_as_float(r[4 * q + 1]) + b4.y;
          v[4 * q + 2] = __uint_as_float(r[4 * q + 2]) + b4.z;
          v[4 * q + 3] = __uint_as_float(r[4 * q + 3]) + b4.w;
        }
        if (last) {
          if (yr) {
            if (c0 + 16 <= Dout) {
#pragma unroll
              for (int j = 0; j < 16; ++j) yr[c0 + j] = v[j];
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (c0 + j < Dout) yr[c0 + j] = v[j];
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = mlp_act(v[j], leaky);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float4 h, lo4;
            tc_split(v[4 * q + 0], h.x, lo4.x); tc_split(v[4 * q + 1], h.y, lo4.y);
            tc_split(v[4 * q + 2], h.z, lo4.z); tc_split(v[4 * q + 3], h.w, lo4.w);
            const size_t o = ((size_t)((c0 >> 2) + q) * kTcRows + t) * 4;
            *reinterpret_cast<float4*>(A_hi + o) = h;
            *reinterpret_cast<float4*>(A_lo + o) = lo4;
          }
          if (hs) {
            if (c0 + 16 <= Dout) {
#pragma unroll
              for (int j = 0; j < 16; ++j) hs[c0 + j] = v[j];
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (c0 + j < Dout) hs[c0 + j] = v[j];
            }
          }
        }
      }
    }
    // the next tile's staging overwrites A: every thread of the group must be done reading TMEM / writing A
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    tc2_group_sync(g);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if ((tid >> 5) == 0) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(*tmem_slot), "n"(2 * COLS) : "memory");
  }
}

inline cudaError_t mlp_tc_forward(const MlpDev& m, long long N, const float* x, float* y, float* hsave, cudaStream_t st) {
  if (tc2_supported(m) && N > 2 * kTcRows) {
    static int sms = 0;
    if (!sms) {
      int dev = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      if (sms <= 0) sms = 148;
    }
    const size_t smem2 = tc2_smem_bytes(m);
    const long long tiles = (N + kTcRows - 1) / kTcRows;
    const unsigned grid2 = (unsigned)((tiles + 1) / 2 < sms ? (tiles + 1) / 2 : sms);
    cudaError_t e2;
    if (tc_tmem_cols(m) <= 32) {
      e2 = cudaFuncSetAttribute(mlp_tc2_forward_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
      if (e2 != cudaSuccess) return e2;
      mlp_tc2_forward_kernel<32><<<grid2, kTc2Threads, smem2, st>>>(m, N, x, y, hsave);
    } else {
      e2 = cudaFuncSetAttribute(mlp_tc2_forward_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
      if (e2 != cudaSuccess) return e2;
      mlp_tc2_forward_kernel<64><<<grid2, kTc2Threads, smem2, st>>>(m, N, x, y, hsave);
    }
    return cudaGetLastError();
  }
  const size_t smem = tc_smem_bytes(m);
  const unsigned grid = (unsigned)((N + kTcRows - 1) / kTcRows);
  const int cols = tc_tmem_cols(m);
  cudaError_t e;
#define PG_TC_LAUNCH(C)                                                                                          \
  e = cudaFuncSetAttribute(mlp_tc_forward_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);    \
  if (e != cudaSuccess) return e;                                                                                \
  mlp_tc_forward_kernel<C><<<grid, kTcThreads, smem, st>>>(m, N, x, y, hsave);
  if (cols <= 32) { PG_TC_LAUNCH(32) }
  else if (cols <= 64) { PG_TC_LAUNCH(64) }
  else { PG_TC_LAUNCH(128) }
#undef PG_TC_LAUNCH
  return cudaGetLastError();
}

}  // namespace pg
