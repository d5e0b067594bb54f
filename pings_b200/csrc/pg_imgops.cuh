// Post-raster loss assembly: the photometric L1 and the masked depth L1 of the
// GS mapping loss (utils/mapper.py:1239,1252-1267; l1_loss = mean |a - b|) in one pass over the
// rendered images, with their backward seeds.
//
// Forward: one streaming kernel reads rendered rgb / gt rgb / rendered depth / gt depth / alpha once,
// writes sign(.) seeds for rgb and the (masked, optionally inverse-depth) depth derivative, and
// reduces three scalars {sum |rgb diff|, sum |depth diff| over valid, valid count} with a block
// reduction + one atomic per block.  Backward: one kernel scales the seeds by upstream / count,
// reading the count on the device (no host sync).
#pragma once
#include "pg_common.cuh"

namespace pg {

struct PhotoDepthArgs {
  int HW;
  const float* rgb;        // [3*HW]
  const float* gt_rgb;     // [3*HW]
  const float* depth;      // [HW]  (nullable -> no depth term)
  const float* gt_depth;   // [HW]
  const float* alpha;      // [HW]  (nullable -> no accumulated-alpha mask)
  float depth_min, depth_max, min_alpha;
  int inverse_depth;
};

// sums[0] = sum |rgb - gt|, sums[1] = sum depth error over valid pixels, sums[2] = number of valid pixels
__global__ void __launch_bounds__(256) photo_depth_fwd_kernel(PhotoDepthArgs a, float* __restrict__ seed_rgb,
                                                              float* __restrict__ seed_depth, double* __restrict__ sums) {
  float s_rgb = 0.f, s_d = 0.f, s_n = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.HW; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float d = a.rgb[c * a.HW + i] - a.gt_rgb[c * a.HW + i];
      s_rgb += fabsf(d);
      seed_rgb[c * a.HW + i] = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);   // torch: sign(0) = 0
    }
    if (a.depth) {
      const float g = a.gt_depth[i], r = a.depth[i];
      bool valid = g > a.depth_min && g < a.depth_max;
      if (a.alpha) valid = valid && a.alpha[i] > a.min_alpha;
      float seed = 0.f;
      if (valid) {
        if (a.inverse_depth) {   // l1(1/gt, 1/rendered): d/dr |1/g - 1/r| = sign(1/g - 1/r) / r^2
          const float e = 1.f / g - 1.f / r;
          s_d += fabsf(e);
          seed = (e > 0.f ? 1.f : (e < 0.f ? -1.f : 0.f)) / (r * r);
        } else {                 // l1(gt, rendered): d/dr |g - r| = -sign(g - r)
          const float e = g - r;
          s_d += fabsf(e);
          seed = e > 0.f ? -1.f : (e < 0.f ? 1.f : 0.f);
        }
        s_n += 1.f;
      }
      seed_depth[i] = seed;
    }
  }
  __shared__ float red[3][8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s_rgb += __shfl_xor_sync(0xffffffffu, s_rgb, o);
    s_d += __shfl_xor_sync(0xffffffffu, s_d, o);
    s_n += __shfl_xor_sync(0xffffffffu, s_n, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = s_rgb; red[1][warp] = s_d; red[2][warp] = s_n; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += (double)red[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], t);
  }
}

// loss = sums[0] / (3 HW) + lambda_depth * sums[1] / max(sums[2], 1)   (an empty mask gives a zero depth term)
__global__ void photo_depth_finish_kernel(int HW, float lambda_depth, const double* __restrict__ sums,
                                          float* __restrict__ loss /* [3]: total, rgb l1, depth l1 */) {
  const double l_rgb = sums[0] / (3.0 * (double)HW);
  const double l_d = sums[2] > 0.0 ? sums[1] / sums[2] : 0.0;
  loss[0] = (float)(l_rgb + (double)lambda_depth * l_d);
  loss[1] = (float)l_rgb;
  loss[2] = (float)l_d;
}

__global__ void __launch_bounds__(256) photo_depth_bwd_kernel(int HW, float lambda_depth, const float* __restrict__ upstream,
                                                              const double* __restrict__ sums,
                                                              const float* __restrict__ seed_rgb,
                                                              const float* __restrict__ seed_depth,
                                                              float* __restrict__ d_rgb, float* __restrict__ d_depth) {
  const float up = upstream[0];
  const float k_rgb = up / (3.f * (float)HW);
  const float k_d = sums[2] > 0.0 ? up * lambda_depth / (float)sums[2] : 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += gridDim.x * blockDim.x) {
    d_rgb[i] = seed_rgb[i] * k_rgb;
    d_rgb[HW + i] = seed_rgb[HW + i] * k_rgb;
    d_rgb[2 * HW + i] = seed_rgb[2 * HW + i] * k_rgb;
    if (d_depth) d_depth[i] = seed_depth[i] * k_d;
  }
}

}  // namespace pg
