// C-ABI entry points of fused SSIM (see include/pings_b200.h).
#include "pg_ssim.cuh"
#include "pg_imgops.cuh"

using namespace pg;

extern "C" {

int pg_ssim_forward(int B, int C, int H, int W, float C1, float C2, const float* img1, const float* img2,
                    float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, void* stream) {
  if (B < 0 || C < 0 || H < 0 || W < 0) return PG_ERR_ARG;
  if ((size_t)B * C * H * W == 0) return PG_OK;
  if (!img1 || !img2 || !ssim_map) return PG_ERR_ARG;
  const bool train = dm_dmu1 != nullptr;
  if (train && (!dm_dsigma1_sq || !dm_dsigma12)) return PG_ERR_ARG;
  if ((long long)B * C > 65535) return PG_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(div_up(W, kSsTile), div_up(H, kSsTile), B * C);
  Phase ph("ssim_fwd", st);
  ssim_fwd_kernel<<<grid, kSsThreads, 0, st>>>(H, W, C1, C2, img1, img2, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12,
                                              nullptr);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_ssim_backward(int B, int C, int H, int W, float C1, float C2, const float* img1, const float* img2,
                     const float* dL_dmap, float dL_dmap_scalar, const float* dm_dmu1, const float* dm_dsigma1_sq,
                     const float* dm_dsigma12, float* dL_dimg1, void* stream) {
  (void)C1; (void)C2;
  if (B < 0 || C < 0 || H < 0 || W < 0) return PG_ERR_ARG;
  if ((size_t)B * C * H * W == 0) return PG_OK;
  if (!img1 || !img2 || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1) return PG_ERR_ARG;
  if ((long long)B * C > 65535) return PG_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(div_up(W, kSsTile), div_up(H, kSsTile), B * C);
  Phase ph("ssim_bwd", st);
  ssim_bwd_kernel<<<grid, kSsThreads, 0, st>>>(H, W, img1, img2, dL_dmap, dL_dmap_scalar, nullptr, dm_dmu1,
                                              dm_dsigma1_sq, dm_dsigma12, dL_dimg1);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

__global__ void ssim_mean_finish_kernel(const double* __restrict__ sum, double count, float* __restrict__ mean) {
  mean[0] = (float)(sum[0] / count);
}

int pg_ssim_mean_forward(int B, int C, int H, int W, float C1, float C2, const float* img1, const float* img2,
                         float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sum_scratch, float* mean,
                         void* stream) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !sum_scratch || !mean) return PG_ERR_ARG;
  const bool train = dm_dmu1 != nullptr;
  if (train && (!dm_dsigma1_sq || !dm_dsigma12)) return PG_ERR_ARG;
  if ((long long)B * C > 65535) return PG_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  PG_CUDA(cudaMemsetAsync(sum_scratch, 0, sizeof(double), st));
  dim3 grid(div_up(W, kSsTile), div_up(H, kSsTile), B * C);
  Phase ph("ssim_fwd", st);
  ssim_fwd_kernel<<<grid, kSsThreads, 0, st>>>(H, W, C1, C2, img1, img2, nullptr, dm_dmu1, dm_dsigma1_sq, dm_dsigma12,
                                              sum_scratch);
  ssim_mean_finish_kernel<<<1, 1, 0, st>>>(sum_scratch, (double)B * C * H * W, mean);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_ssim_mean_backward(int B, int C, int H, int W, const float* img1, const float* img2, const float* upstream,
                          const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1,
                          void* stream) {
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0) return PG_ERR_ARG;
  if (!img1 || !img2 || !upstream || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1) return PG_ERR_ARG;
  if ((long long)B * C > 65535) return PG_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(div_up(W, kSsTile), div_up(H, kSsTile), B * C);
  Phase ph("ssim_bwd", st);
  ssim_bwd_kernel<<<grid, kSsThreads, 0, st>>>(H, W, img1, img2, nullptr, (float)(1.0 / ((double)B * C * H * W)), upstream,
                                              dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dL_dimg1);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_photo_depth_loss_forward(int H, int W, const float* rgb, const float* gt_rgb, const float* depth,
                                const float* gt_depth, const float* alpha, float lambda_depth, float depth_min,
                                float depth_max, float min_alpha, int inverse_depth, float* seed_rgb, float* seed_depth,
                                double* sums, float* loss, void* stream) {
  if (H <= 0 || W <= 0 || !rgb || !gt_rgb || !seed_rgb || !sums || !loss) return PG_ERR_ARG;
  if (depth && (!gt_depth || !seed_depth)) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  PhotoDepthArgs a;
  a.HW = H * W; a.rgb = rgb; a.gt_rgb = gt_rgb; a.depth = depth; a.gt_depth = gt_depth; a.alpha = alpha;
  a.depth_min = depth_min; a.depth_max = depth_max; a.min_alpha = min_alpha; a.inverse_depth = inverse_depth;
  PG_CUDA(cudaMemsetAsync(sums, 0, 3 * sizeof(double), st));
  Phase ph("photo_depth_loss_fwd", st);
  const int blocks = min(div_up(a.HW, 256), 148 * 8);
  photo_depth_fwd_kernel<<<blocks, 256, 0, st>>>(a, seed_rgb, seed_depth, sums);
  photo_depth_finish_kernel<<<1, 1, 0, st>>>(a.HW, depth ? lambda_depth : 0.f, sums, loss);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_photo_depth_loss_backward(int H, int W, float lambda_depth, const float* upstream, const double* sums,
                                 const float* seed_rgb, const float* seed_depth, float* d_rgb, float* d_depth,
                                 void* stream) {
  if (H <= 0 || W <= 0 || !upstream || !sums || !seed_rgb || !d_rgb) return PG_ERR_ARG;
  if (d_depth && !seed_depth) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  Phase ph("photo_depth_loss_bwd", st);
  const int HW = H * W;
  photo_depth_bwd_kernel<<<min(div_up(HW, 256), 148 * 8), 256, 0, st>>>(HW, lambda_depth, upstream, sums, seed_rgb,
                                                                          seed_depth, d_rgb, d_depth);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_depth2normal_forward(int H, int W, float cx, float cy, float fx, float fy, const float* depth, const uint8_t* mask,
                            const float* alpha, float* normal, void* stream) {
  if (H <= 0 || W <= 0 || !depth || !mask || !normal || fx == 0.f || fy == 0.f) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  D2NArgs a{H, W, cx, cy, 1.f / fx, 1.f / fy, depth, mask, alpha};
  Phase ph("depth2normal_fwd", st);
  depth2normal_fwd_kernel<<<div_up(H * W, 256), 256, 0, st>>>(a, normal);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_depth2normal_backward(int H, int W, float cx, float cy, float fx, float fy, const float* depth, const uint8_t* mask,
                             const float* alpha, const float* d_normal, float* d_depth, void* stream) {
  if (H <= 0 || W <= 0 || !depth || !mask || !d_normal || !d_depth || fx == 0.f || fy == 0.f) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  D2NArgs a{H, W, cx, cy, 1.f / fx, 1.f / fy, depth, mask, alpha};
  PG_CUDA(cudaMemsetAsync(d_depth, 0, sizeof(float) * (size_t)H * W, st));
  Phase ph("depth2normal_bwd", st);
  depth2normal_bwd_kernel<<<div_up(H * W, 256), 256, 0, st>>>(a, d_normal, d_depth);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_exposure_affine_forward(int H, int W, const float* img, const float* mat, const float* offset, float* out,
                               void* stream) {
  if (H <= 0 || W <= 0 || !img || !mat || !offset || !out) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  exposure_fwd_kernel<<<min(div_up(H * W, 256), 148 * 8), 256, 0, st>>>(H * W, img, mat, offset, out);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_exposure_affine_backward(int H, int W, const float* img, const float* mat, const float* d_out, float* d_img,
                                float* d_mat_off, void* stream) {
  if (H <= 0 || W <= 0 || !img || !mat || !d_out || !d_mat_off) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  PG_CUDA(cudaMemsetAsync(d_mat_off, 0, 12 * sizeof(float), st));
  exposure_bwd_kernel<<<min(div_up(H * W, 256), 148 * 4), 256, 0, st>>>(H * W, img, mat, d_out, d_img, d_mat_off);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

}  // extern "C"
