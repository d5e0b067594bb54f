// C-ABI entry points of fused SSIM (see include/pings_b200.h).
#include "pg_ssim.cuh"

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
  ssim_fwd_kernel<<<grid, kSsThreads, 0, st>>>(H, W, C1, C2, img1, img2, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12);
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
  ssim_bwd_kernel<<<grid, kSsThreads, 0, st>>>(H, W, img1, img2, dL_dmap, dL_dmap_scalar, dm_dmu1, dm_dsigma1_sq,
                                              dm_dsigma12, dL_dimg1);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

}  // extern "C"
