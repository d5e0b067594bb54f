// Fused SSIM forward / backward (reference: submodules/fused-ssim/ssim.cu:62-278, :286-427).
//
// Same math as the reference (separable 11-tap Gaussian, sigma 1.5, zero padding; per pixel
//   ssim = (2 mu1 mu2 + C1)(2 s12 + C2) / ((mu1^2 + mu2^2 + C1)(s1 + s2 + C2))
// and the three partials dm/dmu1, dm/dsigma1^2, dm/dsigma12 saved for the backward), but laid out
// for B200: a CTA owns a 32x32 output tile of ONE (batch, channel) plane (grid.z = B*C instead of a
// serial channel loop), the halo tile is 42x42 so the L2->SM read amplification is 1.72x instead of
// 2.64x for 16x16 tiles, global loads/stores are 32-lane row-contiguous, and all five (three)
// running sums of a pass stay in registers.  HBM-bound: 24 B/elem fwd (train), 28 B/elem bwd.
#pragma once
#include "pg_common.cuh"

namespace pg {

constexpr int kSsTile = 32;
constexpr int kSsHalo = 5;
constexpr int kSsIn = kSsTile + 2 * kSsHalo;  // 42
constexpr int kSsThreads = 256;

__constant__ float kGauss11[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f,
                                   0.10936068743467331f,  0.21300552785396576f,   0.26601171493530273f,
                                   0.21300552785396576f,  0.10936068743467331f,   0.036000773310661316f,
                                   0.0075987582094967365f, 0.001028380123898387f};

// Both 11-tap passes are register-blocked 4 wide: a thread produces 4 adjacent outputs of a row (horizontal pass,
// lanes along the rows: the 43-/33-float row strides are odd, so a warp's accesses are conflict free) or of a column
// (vertical pass, lanes along x) from a 14-value window held in registers - 3.5 shared loads per output and statistic
// instead of 11.  ncu on the previous one-output-per-thread version: shared-memory pipe 64-68 %, issue 62-81 %, DRAM
// 19 % - the separable passes, not HBM, bound it.
// sum_out (optional): adds the sum of the SSIM map over the tile to a double (the `.mean()` of fused_ssim), in which
// case ssim_map may be NULL and the 4-B/element map is never written.
__global__ void __launch_bounds__(kSsThreads) ssim_fwd_kernel(int H, int W, float C1, float C2,
                                                              const float* __restrict__ img1,
                                                              const float* __restrict__ img2, float* __restrict__ ssim_map,
                                                              float* __restrict__ dm_dmu1, float* __restrict__ dm_ds1,
                                                              float* __restrict__ dm_ds12, double* __restrict__ sum_out) {
  __shared__ float sX[kSsIn][kSsIn + 1];
  __shared__ float sY[kSsIn][kSsIn + 1];
  __shared__ float sH[5][kSsIn][kSsTile + 1];  // horizontal pass results: X, X^2, Y, Y^2, XY
  const size_t plane = (size_t)blockIdx.z * H * W;
  const int x0 = blockIdx.x * kSsTile, y0 = blockIdx.y * kSsTile;
  const int tid = threadIdx.x;
  for (int i = tid; i < kSsIn * kSsIn; i += kSsThreads) {
    const int ly = i / kSsIn, lx = i - ly * kSsIn;
    const int gy = y0 + ly - kSsHalo, gx = x0 + lx - kSsHalo;
    float a = 0.f, b = 0.f;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      a = img1[plane + (size_t)gy * W + gx];
      b = img2[plane + (size_t)gy * W + gx];
    }
    sX[ly][lx] = a;
    sY[ly][lx] = b;
  }
  __syncthreads();
  float g[11];
#pragma unroll
  for (int k = 0; k < 11; ++k) g[k] = kGauss11[k];
  // horizontal: item = (row ly, group of 4 outputs); 42 rows x 8 groups
  for (int i = tid; i < kSsIn * (kSsTile / 4); i += kSsThreads) {
    const int ly = i % kSsIn, ox = (i / kSsIn) * 4;
    float xs[14], ys[14];
#pragma unroll
    for (int k = 0; k < 14; ++k) {
      xs[k] = sX[ly][ox + k];
      ys[k] = sY[ly][ox + k];
    }
#pragma unroll
    for (int o = 0; o < 4; ++o) {
      float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
#pragma unroll
      for (int d = 1; d <= kSsHalo; ++d) {
        const float w = g[kSsHalo - d];
        const float xl = xs[o + kSsHalo - d], xr = xs[o + kSsHalo + d], yl = ys[o + kSsHalo - d], yr = ys[o + kSsHalo + d];
        s0 += (xl + xr) * w;
        s1 += ((xl * xl) + (xr * xr)) * w;
        s2 += (yl + yr) * w;
        s3 += ((yl * yl) + (yr * yr)) * w;
        s4 += ((xl * yl) + (xr * yr)) * w;
      }
      const float w = g[kSsHalo], xc = xs[o + kSsHalo], yc = ys[o + kSsHalo];
      s0 += xc * w; s1 += (xc * xc) * w; s2 += yc * w; s3 += (yc * yc) * w; s4 += (xc * yc) * w;
      sH[0][ly][ox + o] = s0; sH[1][ly][ox + o] = s1; sH[2][ly][ox + o] = s2; sH[3][ly][ox + o] = s3; sH[4][ly][ox + o] = s4;
    }
  }
  __syncthreads();
  // vertical: thread = (column ox, group of 4 rows); 32 columns x 8 groups = 256 threads
  float tsum = 0.f;
  {
    const int ox = tid & 31, oy0 = (tid >> 5) * 4;
    const int px = x0 + ox;
    float o4[5][4];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      float v[14];
#pragma unroll
      for (int r = 0; r < 14; ++r) v[r] = sH[k][oy0 + r][ox];
#pragma unroll
      for (int o = 0; o < 4; ++o) {
        float s = 0.f;
#pragma unroll
        for (int d = 1; d <= kSsHalo; ++d) s += (v[o + kSsHalo - d] + v[o + kSsHalo + d]) * g[kSsHalo - d];
        s += v[o + kSsHalo] * g[kSsHalo];
        o4[k][o] = s;
      }
    }
#pragma unroll
    for (int o = 0; o < 4; ++o) {
      const int py = y0 + oy0 + o;
      if (px >= W || py >= H) continue;
      const float mu1 = o4[0][o], mu2 = o4[2][o];
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2;
      const float sigma1_sq = o4[1][o] - mu1_sq, sigma2_sq = o4[3][o] - mu2_sq, sigma12 = o4[4][o] - mu1 * mu2;
      const float A = mu1_sq + mu2_sq + C1, B = sigma1_sq + sigma2_sq + C2;
      const float C_ = 2.f * mu1 * mu2 + C1, D_ = 2.f * sigma12 + C2;
      const size_t gi = plane + (size_t)py * W + px;
      const float val = (C_ * D_) / (A * B);
      tsum += val;
      if (ssim_map) ssim_map[gi] = val;
      if (dm_dmu1) {
        dm_dmu1[gi] = (mu2 * 2.f * D_) / (A * B) - (mu2 * 2.f * C_) / (A * B) - (mu1 * 2.f * C_ * D_) / (A * A * B) +
                      (mu1 * 2.f * C_ * D_) / (A * B * B);
        dm_ds1[gi] = (-C_ * D_) / (A * B * B);
        dm_ds12[gi] = (2.f * C_) / (A * B);
      }
    }
  }
  if (sum_out) {
    __shared__ float red[kSsThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tsum += __shfl_xor_sync(0xffffffffu, tsum, o);
    if ((tid & 31) == 0) red[tid >> 5] = tsum;
    __syncthreads();
    if (tid == 0) {
      double t = 0.0;
      for (int w = 0; w < kSsThreads / 32; ++w) t += (double)red[w];
      atomicAdd(sum_out, t);
    }
  }
}

// dL/dimg1 = conv(dL*dm_dmu1) + 2 img1 conv(dL*dm_ds1) + img2 conv(dL*dm_ds12)   (ssim.cu:286-427)
// dL_dmap may be NULL with dl_scalar != 0: the map gradient is then the constant dl_scalar (the
// backward of `.mean()`), which saves reading one full plane; dl_scalar_dev (optional) is a device scalar that
// multiplies it (the upstream gradient of the mean, without a host read).
__global__ void __launch_bounds__(kSsThreads) ssim_bwd_kernel(int H, int W, const float* __restrict__ img1,
                                                              const float* __restrict__ img2,
                                                              const float* __restrict__ dL_dmap, float dl_scalar,
                                                              const float* __restrict__ dl_scalar_dev,
                                                              const float* __restrict__ dm_dmu1,
                                                              const float* __restrict__ dm_ds1,
                                                              const float* __restrict__ dm_ds12,
                                                              float* __restrict__ dL_dimg1) {
  __shared__ float sD[3][kSsIn][kSsIn + 1];
  __shared__ float sH[3][kSsIn][kSsTile + 1];
  const size_t plane = (size_t)blockIdx.z * H * W;
  const int x0 = blockIdx.x * kSsTile, y0 = blockIdx.y * kSsTile;
  const int tid = threadIdx.x;
  if (dl_scalar_dev) dl_scalar *= dl_scalar_dev[0];
  for (int i = tid; i < kSsIn * kSsIn; i += kSsThreads) {
    const int ly = i / kSsIn, lx = i - ly * kSsIn;
    const int gy = y0 + ly - kSsHalo, gx = x0 + lx - kSsHalo;
    float a = 0.f, b = 0.f, c = 0.f;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      const size_t g = plane + (size_t)gy * W + gx;
      const float chain = dL_dmap ? dL_dmap[g] : dl_scalar;
      a = dm_dmu1[g] * chain;
      b = dm_ds1[g] * chain;
      c = dm_ds12[g] * chain;
    }
    sD[0][ly][lx] = a; sD[1][ly][lx] = b; sD[2][ly][lx] = c;
  }
  __syncthreads();
  float g[11];
#pragma unroll
  for (int k = 0; k < 11; ++k) g[k] = kGauss11[k];
  for (int i = tid; i < kSsIn * (kSsTile / 4); i += kSsThreads) {
    const int ly = i % kSsIn, ox = (i / kSsIn) * 4;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float v[14];
#pragma unroll
      for (int r = 0; r < 14; ++r) v[r] = sD[k][ly][ox + r];
#pragma unroll
      for (int o = 0; o < 4; ++o) {
        float s = 0.f;
#pragma unroll
        for (int d = 1; d <= kSsHalo; ++d) s += (v[o + kSsHalo - d] + v[o + kSsHalo + d]) * g[kSsHalo - d];
        s += v[o + kSsHalo] * g[kSsHalo];
        sH[k][ly][ox + o] = s;
      }
    }
  }
  __syncthreads();
  {
    const int ox = tid & 31, oy0 = (tid >> 5) * 4;
    const int px = x0 + ox;
    float o3[3][4];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float v[14];
#pragma unroll
      for (int r = 0; r < 14; ++r) v[r] = sH[k][oy0 + r][ox];
#pragma unroll
      for (int o = 0; o < 4; ++o) {
        float s = 0.f;
#pragma unroll
        for (int d = 1; d <= kSsHalo; ++d) s += (v[o + kSsHalo - d] + v[o + kSsHalo + d]) * g[kSsHalo - d];
        s += v[o + kSsHalo] * g[kSsHalo];
        o3[k][o] = s;
      }
    }
#pragma unroll
    for (int o = 0; o < 4; ++o) {
      const int py = y0 + oy0 + o;
      if (px >= W || py >= H) continue;
      const size_t gi = plane + (size_t)py * W + px;
      dL_dimg1[gi] = o3[0][o] + (2.f * img1[gi]) * o3[1][o] + img2[gi] * o3[2][o];
    }
  }
}

}  // namespace pg
