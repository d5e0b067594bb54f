// Real spherical harmonics (degree <= 3) -> RGB and its backward.
// Behaviour follows R3 cuda_rasterizer/forward.cu:22-73 and backward.cu:21-145
// (unused by PINGS itself, which always passes colors_precomp, but part of the API).
#pragma once
#include <cuda_runtime.h>

namespace pg {

__device__ __constant__ float kShC0 = 0.28209479177387814f;
__device__ __constant__ float kShC1 = 0.4886025119029199f;
__device__ __constant__ float kShC2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                          -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float kShC3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                          0.3731763325901154f,  -0.4570457994644658f, 1.445305721320277f,
                                          -0.5900435899266435f};

struct F3 {
  float x, y, z;
};
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ F3 operator*(float s, F3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ F3 operator*(F3 a, float s) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ float dot3(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

__device__ inline F3 sh_to_rgb(int idx, int deg, int max_coeffs, F3 pos, F3 campos, const float* shs,
                               uint8_t* clamped) {
  F3 dir = pos - campos;
  float len = sqrtf(dot3(dir, dir));
  dir = {dir.x / len, dir.y / len, dir.z / len};
  const F3* sh = reinterpret_cast<const F3*>(shs) + (size_t)idx * max_coeffs;
  F3 res = kShC0 * sh[0];
  if (deg > 0) {
    float x = dir.x, y = dir.y, z = dir.z;
    res = res - kShC1 * y * sh[1] + kShC1 * z * sh[2] - kShC1 * x * sh[3];
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      res = res + kShC2[0] * xy * sh[4] + kShC2[1] * yz * sh[5] + kShC2[2] * (2.0f * zz - xx - yy) * sh[6] +
            kShC2[3] * xz * sh[7] + kShC2[4] * (xx - yy) * sh[8];
      if (deg > 2) {
        res = res + kShC3[0] * y * (3.0f * xx - yy) * sh[9] + kShC3[1] * xy * z * sh[10] +
              kShC3[2] * y * (4.0f * zz - xx - yy) * sh[11] +
              kShC3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
              kShC3[4] * x * (4.0f * zz - xx - yy) * sh[13] + kShC3[5] * z * (xx - yy) * sh[14] +
              kShC3[6] * x * (xx - 3.0f * yy) * sh[15];
      }
    }
  }
  res.x += 0.5f;
  res.y += 0.5f;
  res.z += 0.5f;
  clamped[3 * idx + 0] = res.x < 0;
  clamped[3 * idx + 1] = res.y < 0;
  clamped[3 * idx + 2] = res.z < 0;
  return {fmaxf(res.x, 0.f), fmaxf(res.y, 0.f), fmaxf(res.z, 0.f)};
}

// d(normalize(v))/dv applied to dv.
__device__ __forceinline__ F3 dnormvdv3(F3 v, F3 dv) {
  float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  F3 r;
  r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
  r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
  r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
  return r;
}

// Returns dL/dmean contribution of the view direction; writes dL_dsh.
__device__ inline F3 sh_to_rgb_backward(int idx, int deg, int max_coeffs, F3 pos, F3 campos, const float* shs,
                                        const uint8_t* clamped, F3 dL_dRGB, float* dL_dshs) {
  F3 dir_orig = pos - campos;
  float len = sqrtf(dot3(dir_orig, dir_orig));
  F3 dir = {dir_orig.x / len, dir_orig.y / len, dir_orig.z / len};
  const F3* sh = reinterpret_cast<const F3*>(shs) + (size_t)idx * max_coeffs;
  dL_dRGB.x *= clamped[3 * idx + 0] ? 0.f : 1.f;
  dL_dRGB.y *= clamped[3 * idx + 1] ? 0.f : 1.f;
  dL_dRGB.z *= clamped[3 * idx + 2] ? 0.f : 1.f;
  F3 dx = {0, 0, 0}, dy = {0, 0, 0}, dz = {0, 0, 0};
  float x = dir.x, y = dir.y, z = dir.z;
  F3* out = reinterpret_cast<F3*>(dL_dshs) + (size_t)idx * max_coeffs;
  // coefficients above the active degree get no gradient (the reference's dL_dsh is zero-filled,
  // R3 rasterize_points.cu:166; ours is not pre-zeroed)
  for (int k = (deg + 1) * (deg + 1); k < max_coeffs; ++k) out[k] = F3{0.f, 0.f, 0.f};
  out[0] = kShC0 * dL_dRGB;
  if (deg > 0) {
    out[1] = (-kShC1 * y) * dL_dRGB;
    out[2] = (kShC1 * z) * dL_dRGB;
    out[3] = (-kShC1 * x) * dL_dRGB;
    dx = -kShC1 * sh[3];
    dy = -kShC1 * sh[1];
    dz = kShC1 * sh[2];
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      out[4] = (kShC2[0] * xy) * dL_dRGB;
      out[5] = (kShC2[1] * yz) * dL_dRGB;
      out[6] = (kShC2[2] * (2.f * zz - xx - yy)) * dL_dRGB;
      out[7] = (kShC2[3] * xz) * dL_dRGB;
      out[8] = (kShC2[4] * (xx - yy)) * dL_dRGB;
      dx = dx + kShC2[0] * y * sh[4] + kShC2[2] * 2.f * -x * sh[6] + kShC2[3] * z * sh[7] + kShC2[4] * 2.f * x * sh[8];
      dy = dy + kShC2[0] * x * sh[4] + kShC2[1] * z * sh[5] + kShC2[2] * 2.f * -y * sh[6] + kShC2[4] * 2.f * -y * sh[8];
      dz = dz + kShC2[1] * y * sh[5] + kShC2[2] * 2.f * 2.f * z * sh[6] + kShC2[3] * x * sh[7];
      if (deg > 2) {
        out[9] = (kShC3[0] * y * (3.f * xx - yy)) * dL_dRGB;
        out[10] = (kShC3[1] * xy * z) * dL_dRGB;
        out[11] = (kShC3[2] * y * (4.f * zz - xx - yy)) * dL_dRGB;
        out[12] = (kShC3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy)) * dL_dRGB;
        out[13] = (kShC3[4] * x * (4.f * zz - xx - yy)) * dL_dRGB;
        out[14] = (kShC3[5] * z * (xx - yy)) * dL_dRGB;
        out[15] = (kShC3[6] * x * (xx - 3.f * yy)) * dL_dRGB;
        dx = dx + (kShC3[0] * sh[9] * 3.f * 2.f * xy + kShC3[1] * sh[10] * yz + kShC3[2] * sh[11] * -2.f * xy +
                   kShC3[3] * sh[12] * -3.f * 2.f * xz + kShC3[4] * sh[13] * (-3.f * xx + 4.f * zz - yy) +
                   kShC3[5] * sh[14] * 2.f * xz + kShC3[6] * sh[15] * 3.f * (xx - yy));
        dy = dy + (kShC3[0] * sh[9] * 3.f * (xx - yy) + kShC3[1] * sh[10] * xz +
                   kShC3[2] * sh[11] * (-3.f * yy + 4.f * zz - xx) + kShC3[3] * sh[12] * -3.f * 2.f * yz +
                   kShC3[4] * sh[13] * -2.f * xy + kShC3[5] * sh[14] * -2.f * yz + kShC3[6] * sh[15] * -3.f * 2.f * xy);
        dz = dz + (kShC3[1] * sh[10] * xy + kShC3[2] * sh[11] * 4.f * 2.f * yz +
                   kShC3[3] * sh[12] * 3.f * (2.f * zz - xx - yy) + kShC3[4] * sh[13] * 4.f * 2.f * xz +
                   kShC3[5] * sh[14] * (xx - yy));
      }
    }
  }
  F3 dL_ddir = {dot3(dx, dL_dRGB), dot3(dy, dL_dRGB), dot3(dz, dL_dRGB)};
  return dnormvdv3(dir_orig, dL_ddir);
}

}  // namespace pg
