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

__global__ void __launch_bounds__(kSsThreads) ssim_fwd_kernel(int H, int W, float C1, float C2,
                                                              const float* __restrict__ img1,
                                                              const float* __restrict__ img2, float* __restrict__ ssim_map,
                                                              float* __restrict__ dm_dmu1, float* __restrict__ dm_ds1,
                                                              float* __restrict__ dm_ds12) {
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
  for (int i = tid; i < kSsIn * kSsTile; i += kSsThreads) {
    const int ly = i / kSsTile, ox = i - ly * kSsTile;
    const int lx = ox + kSsHalo;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f, s4 = 0.f;
#pragma unroll
    for (int d = 1; d <= kSsHalo; ++d) {
      const float w = kGauss11[kSsHalo - d];
      const float xl = sX[ly][lx - d], xr = sX[ly][lx + d], yl = sY[ly][lx - d], yr = sY[ly][lx + d];
      s0 += (xl + xr) * w;
      s1 += ((xl * xl) + (xr * xr)) * w;
      s2 += (yl + yr) * w;
      s3 += ((yl * yl) + (yr * yr)) * w;
      s4 += ((xl * yl) + (xr * yr)) * w;
    }
    {
      const float w = kGauss11[kSsHalo];
      const float xc = sX[ly][lx], yc = sY[ly][lx];
      s0 += xc * w; s1 += (xc * xc) * w; s2 += yc * w; s3 += (yc * yc) * w; s4 += (xc * yc) * w;
    }
    sH[0][ly][ox] = s0; sH[1][ly][ox] = s1; sH[2][ly][ox] = s2; sH[3][ly][ox] = s3; sH[4][ly][ox] = s4;
  }
  __syncthreads();
  for (int i = tid; i < kSsTile * kSsTile; i += kSsThreads) {
    const int oy = i / kSsTile, ox = i - oy * kSsTile;
    const int px = x0 + ox, py = y0 + oy;
    if (px >= W || py >= H) continue;
    const int ly = oy + kSsHalo;
    float o[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      float s = 0.f;
#pragma unroll
      for (int d = 1; d <= kSsHalo; ++d) s += (sH[k][ly - d][ox] + sH[k][ly + d][ox]) * kGauss11[kSsHalo - d];
      s += sH[k][ly][ox] * kGauss11[kSsHalo];
      o[k] = s;
    }
    const float mu1 = o[0], mu2 = o[2];
    const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2;
    const float sigma1_sq = o[1] - mu1_sq, sigma2_sq = o[3] - mu2_sq, sigma12 = o[4] - mu1 * mu2;
    const float A = mu1_sq + mu2_sq + C1, B = sigma1_sq + sigma2_sq + C2;
    const float C_ = 2.f * mu1 * mu2 + C1, D_ = 2.f * sigma12 + C2;
    const size_t g = plane + (size_t)py * W + px;
    ssim_map[g] = (C_ * D_) / (A * B);
    if (dm_dmu1) {
      dm_dmu1[g] = (mu2 * 2.f * D_) / (A * B) - (mu2 * 2.f * C_) / (A * B) - (mu1 * 2.f * C_ * D_) / (A * A * B) +
                   (mu1 * 2.f * C_ * D_) / (A * B * B);
      dm_ds1[g] = (-C_ * D_) / (A * B * B);
      dm_ds12[g] = (2.f * C_) / (A * B);
    }
  }
}

// dL/dimg1 = conv(dL*dm_dmu1) + 2 img1 conv(dL*dm_ds1) + img2 conv(dL*dm_ds12)   (ssim.cu:286-427)
// dL_dmap may be NULL with dl_scalar != 0: the map gradient is then the constant dl_scalar (the
// backward of `.mean()`), which saves reading one full plane.
__global__ void __launch_bounds__(kSsThreads) ssim_bwd_kernel(int H, int W, const float* __restrict__ img1,
                                                              const float* __restrict__ img2,
                                                              const float* __restrict__ dL_dmap, float dl_scalar,
                                                              const float* __restrict__ dm_dmu1,
                                                              const float* __restrict__ dm_ds1,
                                                              const float* __restrict__ dm_ds12,
                                                              float* __restrict__ dL_dimg1) {
  __shared__ float sD[3][kSsIn][kSsIn + 1];
  __shared__ float sH[3][kSsIn][kSsTile + 1];
  const size_t plane = (size_t)blockIdx.z * H * W;
  const int x0 = blockIdx.x * kSsTile, y0 = blockIdx.y * kSsTile;
  const int tid = threadIdx.x;
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
  for (int i = tid; i < kSsIn * kSsTile; i += kSsThreads) {
    const int ly = i / kSsTile, ox = i - ly * kSsTile;
    const int lx = ox + kSsHalo;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float s = 0.f;
#pragma unroll
      for (int d = 1; d <= kSsHalo; ++d) s += (sD[k][ly][lx - d] + sD[k][ly][lx + d]) * kGauss11[kSsHalo - d];
      s += sD[k][ly][lx] * kGauss11[kSsHalo];
      sH[k][ly][ox] = s;
    }
  }
  __syncthreads();
  for (int i = tid; i < kSsTile * kSsTile; i += kSsThreads) {
    const int oy = i / kSsTile, ox = i - oy * kSsTile;
    const int px = x0 + ox, py = y0 + oy;
    if (px >= W || py >= H) continue;
    const int ly = oy + kSsHalo;
    float o[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float s = 0.f;
#pragma unroll
      for (int d = 1; d <= kSsHalo; ++d) s += (sH[k][ly - d][ox] + sH[k][ly + d][ox]) * kGauss11[kSsHalo - d];
      s += sH[k][ly][ox] * kGauss11[kSsHalo];
      o[k] = s;
    }
    const size_t g = plane + (size_t)py * W + px;
    dL_dimg1[g] = o[0] + (2.f * img1[g]) * o[1] + img2[g] * o[2];
  }
}

}  // namespace pg
