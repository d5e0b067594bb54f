// Small fp32 linear-algebra helpers used by the rasterizer kernels.
//
// Parity note: tile rectangles, radii and depth keys are integer decisions taken on fp32
// results, so the expression *shapes* (association order of the 3-term dot products, which
// nvcc then contracts into FMAs) follow the reference's evaluation order:
// column-major 3x3 product with each element evaluated as (a*b + c*d) + e*f
// (R3 cuda_rasterizer/forward.cu:76-154 uses GLM's mat3 with exactly that association).
#pragma once
#include <cuda_runtime.h>

namespace pg {

// Column-major 3x3: c[col][row].
struct M3 {
  float c[3][3];
};

__host__ __device__ __forceinline__ M3 m3_mul(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
    }
  }
  return r;
}

__host__ __device__ __forceinline__ M3 m3_transpose(const M3& a) {
  M3 r;
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[i][j];
  return r;
}

// p' = M p + t for the 4x4 (stored transposed / column-major) matrices of the raster settings.
__host__ __device__ __forceinline__ float3 xform_point_4x3(const float3& p, const float* m) {
  float3 r;
  r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  return r;
}

__host__ __device__ __forceinline__ float4 xform_point_4x4(const float3& p, const float* m) {
  float4 r;
  r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  r.w = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
  return r;
}

// r = M^T v (upper-left 3x3 of the stored matrix).
__host__ __device__ __forceinline__ float3 xform_vec_4x3_T(const float3& p, const float* m) {
  float3 r;
  r.x = m[0] * p.x + m[1] * p.y + m[2] * p.z;
  r.y = m[4] * p.x + m[5] * p.y + m[6] * p.z;
  r.z = m[8] * p.x + m[9] * p.y + m[10] * p.z;
  return r;
}

// NDC -> pixel, evaluated in double like the reference (R3 auxiliary.h:41-44).
__host__ __device__ __forceinline__ float ndc_to_pix(float v, int S) {
  return (float)(((v + 1.0) * S - 1.0) * 0.5);
}

// Tile rectangle of a disc (R3 auxiliary.h:46-56), 16x16 tiles.
__host__ __device__ __forceinline__ void tile_rect(float px, float py, int radius, int gx, int gy,
                                                   int& x0, int& y0, int& x1, int& y1) {
  auto clampi = [](int v, int hi) { return v < 0 ? 0 : (v > hi ? hi : v); };
  x0 = clampi((int)((px - radius) / 16), gx);
  y0 = clampi((int)((py - radius) / 16), gy);
  x1 = clampi((int)((px + radius + 16 - 1) / 16), gx);
  y1 = clampi((int)((py + radius + 16 - 1) / 16), gy);
}

// Reciprocal for normal-range arguments: MUFU.RCP + one Newton step (3 instructions, no slow-path branch;
// <= 1 ulp).  __frcp_rn / IEEE division cost ~10 instructions and a divergent call each in the blend loops.
__device__ __forceinline__ float rcp_nr(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  const float e = fmaf(-x, r, 1.0f);
  return fmaf(r, e, r);
}

}  // namespace pg
