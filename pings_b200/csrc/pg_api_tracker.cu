// Consumer of the SDF query path with tiny dense algebra: one Gauss-Newton / LM step of the
// point-to-implicit-model registration, utils/tracker.py:353-607 (registration_step: validity mask, Geman-McClure /
// normal-consistency weights) and :610-759 (implicit_reg, implicit_color_reg: J = [p x g, g], N = J^T W J, g = -J^T W r).
// The reference runs ~40 eager kernels, four boolean-mask compactions (each a host sync) and two [N,6] matmuls per
// iteration; here: one kernel for the per-point weights (no compaction: an invalid point gets weight 0), one for the
// 6x6 normal equations (27 sums + the weighted squared residual, accumulated in fp64), one 256-byte read-back, and the
// 6x6 solve on the host in fp64 as the reference does on the device.
#include "pg_common.cuh"

namespace pg {

constexpr int kRegThreads = 256;

struct RegWeightArgs {
  long long N;
  const float* sdf_pred;    // [N]
  const float* sdf_grad;    // [N,3]
  const float* sdf_labels;  // [N]
  const float* normals;     // [N,3] or NULL
  const uint8_t* mask;      // [N] nn_count >= mask_min_nn_count
  const float* sdf_std;     // [N]
  const float* w_extra;     // [N] or NULL (colour-consistency weight, tracker.py:498-503)
  float min_grad_norm, max_grad_norm, max_sdf_std, gm_dist, gm_grad;   // gm_* <= 0: that weight is 1
  int div_grad_norm;        // reg_dist_div_grad_norm
};

// residual / weight = 0 for invalid points; sums = {valid count, sum w, sum |residual|}
__global__ void __launch_bounds__(kRegThreads) reg_weights_kernel(RegWeightArgs a, float* __restrict__ residual,
                                                                  float* __restrict__ weight, uint8_t* __restrict__ valid,
                                                                  double* __restrict__ sums) {
  double s_cnt = 0.0, s_w = 0.0, s_abs = 0.0;
  for (long long i = (long long)blockIdx.x * kRegThreads + threadIdx.x; i < a.N; i += (long long)gridDim.x * kRegThreads) {
    const float gx = a.sdf_grad[3 * i], gy = a.sdf_grad[3 * i + 1], gz = a.sdf_grad[3 * i + 2];
    const float gn = sqrtf(gx * gx + gy * gy + gz * gz);
    const bool ok = a.mask[i] && gn < a.max_grad_norm && gn > a.min_grad_norm && a.sdf_std[i] < a.max_sdf_std;
    float r = 0.f, w = 0.f;
    if (ok) {
      float pred = a.sdf_pred[i];
      if (a.div_grad_norm) pred = pred / gn;
      r = pred - a.sdf_labels[i];
      const float an = gn - 1.0f;
      w = 1.f;
      if (a.gm_dist > 0.f) { const float t = a.gm_dist / (a.gm_dist + r * r); w *= t * t; }
      if (a.gm_grad > 0.f) { const float t = a.gm_grad / (a.gm_grad + an * an); w *= t * t; }
      if (a.normals) {
        const float inv = 1.f / (gn + 1e-7f);
        const float d = a.normals[3 * i] * (gx * inv) + a.normals[3 * i + 1] * (gy * inv) + a.normals[3 * i + 2] * (gz * inv);
        w *= 0.5f + fabsf(d);
      }
      if (a.w_extra) w *= a.w_extra[i];
      s_cnt += 1.0; s_w += (double)w; s_abs += (double)fabsf(r);
    }
    residual[i] = r;
    weight[i] = w;
    valid[i] = ok ? 1 : 0;
  }
  __shared__ double red[3][kRegThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s_cnt += __shfl_xor_sync(0xffffffffu, s_cnt, o);
    s_w += __shfl_xor_sync(0xffffffffu, s_w, o);
    s_abs += __shfl_xor_sync(0xffffffffu, s_abs, o);
  }
  if (lane == 0) { red[0][warp] = s_cnt; red[1][warp] = s_w; red[2][warp] = s_abs; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < kRegThreads / 32; ++w) t += red[threadIdx.x][w];
    atomicAdd(sums + threadIdx.x, t);
  }
}

// out[0..20] = sum w J_a J_b (a <= b, row-major upper triangle), out[21..26] = sum w J_a r, out[27] = sum w r^2,
// J = [p x g, g] (rotation first, then translation: tracker.py:640-644).
__global__ void __launch_bounds__(kRegThreads) reg_normal_eq_kernel(long long N, const float* __restrict__ points,
                                                                    const float* __restrict__ grad,
                                                                    const float* __restrict__ residual,
                                                                    const float* __restrict__ weight, double* __restrict__ out) {
  double acc[28];
#pragma unroll
  for (int k = 0; k < 28; ++k) acc[k] = 0.0;
  for (long long i = (long long)blockIdx.x * kRegThreads + threadIdx.x; i < N; i += (long long)gridDim.x * kRegThreads) {
    const float w = weight ? weight[i] : 1.f;
    if (w == 0.f) continue;
    const float px = points[3 * i], py = points[3 * i + 1], pz = points[3 * i + 2];
    const float gx = grad[3 * i], gy = grad[3 * i + 1], gz = grad[3 * i + 2];
    const float J[6] = {py * gz - pz * gy, pz * gx - px * gz, px * gy - py * gx, gx, gy, gz};
    const float r = residual[i];
    int k = 0;
#pragma unroll
    for (int p = 0; p < 6; ++p) {
      const float wj = w * J[p];
#pragma unroll
      for (int q = p; q < 6; ++q) acc[k++] += (double)(wj * J[q]);
    }
#pragma unroll
    for (int p = 0; p < 6; ++p) acc[21 + p] += (double)((J[p] * w) * r);
    acc[27] += (double)(w * (r * r));
  }
  __shared__ double red[28][kRegThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 28; ++k) {
    double v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[k][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x < 28) {
    double t = 0.0;
    for (int w = 0; w < kRegThreads / 32; ++w) t += red[threadIdx.x][w];
    atomicAdd(out + threadIdx.x, t);
  }
}

}  // namespace pg

using namespace pg;

extern "C" {

int pg_registration_weights(int64_t N, const float* sdf_pred, const float* sdf_grad, const float* sdf_labels,
                            const float* normals, const uint8_t* mask, const float* sdf_std, const float* w_extra,
                            float min_grad_norm, float max_grad_norm, float max_sdf_std, float gm_dist, float gm_grad,
                            int div_grad_norm, float* residual, float* weight, uint8_t* valid, double* sums, void* stream) {
  if (N < 0 || !sums) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  PG_CUDA(cudaMemsetAsync(sums, 0, 4 * sizeof(double), st));
  if (N == 0) return PG_OK;
  if (!sdf_pred || !sdf_grad || !sdf_labels || !mask || !sdf_std || !residual || !weight || !valid) return PG_ERR_ARG;
  RegWeightArgs a{N, sdf_pred, sdf_grad, sdf_labels, normals, mask, sdf_std, w_extra, min_grad_norm, max_grad_norm,
                  max_sdf_std, gm_dist, gm_grad, div_grad_norm};
  long long blocks = (N + kRegThreads - 1) / kRegThreads;
  if (blocks > 148 * 4) blocks = 148 * 4;
  reg_weights_kernel<<<(unsigned)blocks, kRegThreads, 0, st>>>(a, residual, weight, valid, sums);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_normal_equations(int64_t N, const float* points, const float* grad, const float* residual, const float* weight,
                        double* out28, void* stream) {
  if (N < 0 || !out28) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  PG_CUDA(cudaMemsetAsync(out28, 0, 28 * sizeof(double), st));
  if (N == 0) return PG_OK;
  if (!points || !grad || !residual) return PG_ERR_ARG;
  long long blocks = (N + kRegThreads - 1) / kRegThreads;
  if (blocks > 148 * 2) blocks = 148 * 2;
  reg_normal_eq_kernel<<<(unsigned)blocks, kRegThreads, 0, st>>>(N, points, grad, residual, weight, out28);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

}  // extern "C"
