// 3DGS forward kernels: per-Gaussian preprocess and the per-tile alpha blend.
//
// Record layout written by preprocess (3 x float4 per Gaussian, 48 B, 16-B aligned so one
// LDGSTS/cp.async moves each third):
//   rec[3i+0] = (px, py, hx, hy)         pixel centre + conservative half-extent of the
//                                        alpha >= 1/255 ellipse (used for sub-tile culling)
//   rec[3i+1] = (conic.x, conic.y, conic.z, opacity)
//   rec[3i+2] = (depth, r, g, b)
// The reference keeps these in four separate arrays (means2D, conic_opacity, depths, rgb:
// R3 rasterizer_impl.cu:155-170) and reads rgb from global memory per pixel-Gaussian pair.
#pragma once
#include "pg_math.cuh"
#include "pg_raster_common.cuh"
#include "pg_sh.cuh"

namespace pg {

// Sigma = (S R)^T (S R) from scale + (un-normalised) quaternion (R3 forward.cu:120-154).
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float4 q,
                                                     float* cov6) {
  M3 S = {{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
  S.c[0][0] = mod * sx;
  S.c[1][1] = mod * sy;
  S.c[2][2] = mod * sz;
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 R;
  R.c[0][0] = 1.f - 2.f * (y * y + z * z);
  R.c[0][1] = 2.f * (x * y - r * z);
  R.c[0][2] = 2.f * (x * z + r * y);
  R.c[1][0] = 2.f * (x * y + r * z);
  R.c[1][1] = 1.f - 2.f * (x * x + z * z);
  R.c[1][2] = 2.f * (y * z - r * x);
  R.c[2][0] = 2.f * (x * z - r * y);
  R.c[2][1] = 2.f * (y * z + r * x);
  R.c[2][2] = 1.f - 2.f * (x * x + y * y);
  M3 Mm = m3_mul(S, R);
  M3 Sg = m3_mul(m3_transpose(Mm), Mm);
  cov6[0] = Sg.c[0][0];
  cov6[1] = Sg.c[0][1];
  cov6[2] = Sg.c[0][2];
  cov6[3] = Sg.c[1][1];
  cov6[4] = Sg.c[1][2];
  cov6[5] = Sg.c[2][2];
}

// EWA projection of the 3D covariance (R3 forward.cu:76-115). Returns (a, b, c) with the
// 0.3 px low-pass already added. `t_out`/`T_out` (optional) expose intermediates for backward.
struct Cov2DAux {
  float3 t;   // clamped view-space mean
  float txtz, tytz;
  M3 J, Wm, T, Vrk;
};
__device__ __forceinline__ float3 cov2d_ewa(const float3& mean, float fx, float fy, float tanx, float tany,
                                            const float* cov3D, const float* view, Cov2DAux* aux) {
  float3 t = xform_point_4x3(mean, view);
  const float limx = 1.3f * tanx, limy = 1.3f * tany;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  M3 J;
  J.c[0][0] = fx / t.z; J.c[0][1] = 0.f;       J.c[0][2] = -(fx * t.x) / (t.z * t.z);
  J.c[1][0] = 0.f;      J.c[1][1] = fy / t.z;  J.c[1][2] = -(fy * t.y) / (t.z * t.z);
  J.c[2][0] = 0.f;      J.c[2][1] = 0.f;       J.c[2][2] = 0.f;
  M3 Wm;
  Wm.c[0][0] = view[0]; Wm.c[0][1] = view[4]; Wm.c[0][2] = view[8];
  Wm.c[1][0] = view[1]; Wm.c[1][1] = view[5]; Wm.c[1][2] = view[9];
  Wm.c[2][0] = view[2]; Wm.c[2][1] = view[6]; Wm.c[2][2] = view[10];
  M3 T = m3_mul(Wm, J);
  M3 Vrk;
  Vrk.c[0][0] = cov3D[0]; Vrk.c[0][1] = cov3D[1]; Vrk.c[0][2] = cov3D[2];
  Vrk.c[1][0] = cov3D[1]; Vrk.c[1][1] = cov3D[3]; Vrk.c[1][2] = cov3D[4];
  Vrk.c[2][0] = cov3D[2]; Vrk.c[2][1] = cov3D[4]; Vrk.c[2][2] = cov3D[5];
  M3 cov = m3_mul(m3_mul(m3_transpose(T), m3_transpose(Vrk)), T);
  cov.c[0][0] += 0.3f;
  cov.c[1][1] += 0.3f;
  if (aux) {
    aux->t = t; aux->txtz = txtz; aux->tytz = tytz;
    aux->J = J; aux->Wm = Wm; aux->T = T; aux->Vrk = Vrk;
  }
  return make_float3(cov.c[0][0], cov.c[0][1], cov.c[1][1]);
}

// Conservative half-extents (pixels) of the region where opacity*exp(power) can reach 1/255.
// power = -0.5 d^T Q d  =>  d^T Q d <= 2 ln(255 o); the x/y extents of that ellipse are
// sqrt(2 ln(255 o) * cov_xx), sqrt(... * cov_yy).  Culling with these never removes a pair
// the exact per-pixel test would keep (margin covers fp32 rounding of power).
__device__ __forceinline__ float2 alpha_extent(float opacity, float cov_xx, float cov_yy) {
  if (!(opacity * 255.0f >= 1.0f)) return make_float2(-1.f, -1.f);
  const float s = 2.0f * logf(opacity * 255.0f);
  return make_float2(1.001f * sqrtf(s * cov_xx) + 0.5f, 1.001f * sqrtf(s * cov_yy) + 0.5f);
}

struct R3PreArgs {
  int P, D, M, W, H, gx, gy;
  const float *means, *scales, *rots, *opac, *shs, *cov3D_pre, *colors_pre, *view, *proj, *campos;
  float mod, tanx, tany, fx, fy;
  int prefiltered;
};

__global__ void __launch_bounds__(256) r3_preprocess_kernel(R3PreArgs a, int* __restrict__ radii,
                                                            float4* __restrict__ rec, uint32_t* __restrict__ dkey,
                                                            uint32_t* __restrict__ tiles, uint32_t* __restrict__ wts,
                                                            uint2* __restrict__ rect, uint8_t* __restrict__ clamped,
                                                            unsigned long long* __restrict__ total) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t my_tiles = 0, my_key = 0xffffffffu;
  if (idx < a.P) {
    int radius = 0;
    uint32_t key = 0xffffffffu;
    do {
      const float3 p = make_float3(a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]);
      const float3 pv = xform_point_4x3(p, a.view);
      if (pv.z <= 0.2f) {  // near cull (R3 auxiliary.h:154)
        if (a.prefiltered) {
          printf("Point is filtered although prefiltered is set. This shouldn't happen!");
          __trap();
        }
        break;
      }
      const float4 ph = xform_point_4x4(p, a.proj);
      const float pw = 1.0f / (ph.w + 0.0000001f);
      const float3 pp = make_float3(ph.x * pw, ph.y * pw, ph.z * pw);
      float cov6[6];
      if (a.cov3D_pre) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cov6[k] = a.cov3D_pre[6 * idx + k];
      } else {
        const float4 q = reinterpret_cast<const float4*>(a.rots)[idx];
        cov3d_from_scale_rot(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2], a.mod, q, cov6);
      }
      const float3 cov = cov2d_ewa(p, a.fx, a.fy, a.tanx, a.tany, cov6, a.view, nullptr);
      const float det = (cov.x * cov.z - cov.y * cov.y);
      if (det == 0.0f) break;
      const float det_inv = 1.f / det;
      const float3 conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);
      const float mid = 0.5f * (cov.x + cov.z);
      const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
      const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
      const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
      const float px = ndc_to_pix(pp.x, a.W), py = ndc_to_pix(pp.y, a.H);
      int x0, y0, x1, y1;
      tile_rect(px, py, (int)my_radius, a.gx, a.gy, x0, y0, x1, y1);
      if ((x1 - x0) * (y1 - y0) == 0) break;
      float cr, cg, cb;
      if (a.colors_pre) {
        cr = a.colors_pre[3 * idx]; cg = a.colors_pre[3 * idx + 1]; cb = a.colors_pre[3 * idx + 2];
      } else {
        F3 c = sh_to_rgb(idx, a.D, a.M, {p.x, p.y, p.z}, {a.campos[0], a.campos[1], a.campos[2]}, a.shs, clamped);
        cr = c.x; cg = c.y; cb = c.z;
      }
      const float o = a.opac[idx];
      const float2 h = alpha_extent(o, cov.x, cov.z);
      rec[3 * idx + 0] = make_float4(px, py, h.x, h.y);
      rec[3 * idx + 1] = make_float4(conic.x, conic.y, conic.z, o);
      rec[3 * idx + 2] = make_float4(pv.z, cr, cg, cb);
      radius = (int)my_radius;
      key = __float_as_uint(pv.z);
      my_tiles = (uint32_t)((y1 - y0) * (x1 - x0));
      rect[idx] = make_uint2((uint32_t)x0 | ((uint32_t)x1 << 16), (uint32_t)y0 | ((uint32_t)y1 << 16));
    } while (false);
    radii[idx] = radius;
    tiles[idx] = my_tiles;
    wts[idx] = bin_weight(my_tiles);
    dkey[idx] = key;
    my_key = key;
    if (!my_tiles) rect[idx] = make_uint2(0u, 0u);
  }
  // total instance count and key range: block reduce -> three atomics per CTA
  preprocess_block_tail(my_tiles, my_key, total);
}

// checkFrustum (R3 rasterizer_impl.cu:54-66 / auxiliary.h:139-164).
__global__ void __launch_bounds__(256) r3_mark_visible_kernel(int P, const float* __restrict__ means,
                                                              const float* __restrict__ view,
                                                              uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
  const float3 pv = xform_point_4x3(p, view);
  present[idx] = pv.z > 0.2f ? 1 : 0;
}

// ---- alpha blend forward -------------------------------------------------------------------
// CTA = one 16x16 tile, 2 warps; warp w owns rows [8w, 8w+8); each lane owns a 2x2 pixel quad.
// Gaussians are staged 64 at a time with cp.async (LDGSTS) into a double-buffered SoA slab;
// each warp ballots the staged quads' extents against its 16x8 region and only walks the hits
// (warp-ballot culling), and leaves the loop when all of its pixels have saturated.
constexpr int kRThreads = 64;
constexpr int kRBatch = 64;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gsrc));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

struct RSlab {
  float4 a[2][kRBatch];  // px,py,hx,hy
  float4 b[2][kRBatch];  // conic + opacity
  float4 c[2][kRBatch];  // depth,r,g,b
  uint32_t id[2][kRBatch];
};

__global__ void __launch_bounds__(kRThreads) r3_render_fwd_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H, int gx,
    const float4* __restrict__ rec, const float* __restrict__ bg, float* __restrict__ out_color,
    float* __restrict__ out_depth, float* __restrict__ out_opacity, float* __restrict__ final_T,
    uint32_t* __restrict__ n_contrib, int* __restrict__ n_touched) {
  __shared__ RSlab sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_wait();   // tail of the binning chain (programmatic dependent launch)
  const int tile = blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  // warp region and this lane's quad
  const int rx0 = tx * kTile, ry0 = ty * kTile + warp * 8;
  const int qx = rx0 + (lane & 7) * 2, qy = ry0 + (lane >> 3) * 2;
  const float fx0 = (float)qx, fx1 = (float)(qx + 1), fy0 = (float)qy, fy1 = (float)(qy + 1);
  const float wx0 = (float)rx0, wx1 = (float)min(rx0 + 15, W - 1);
  const float wy0 = (float)ry0, wy1 = (float)min(ry0 + 7, H - 1);
  bool in[4];
  in[0] = qx < W && qy < H;
  in[1] = qx + 1 < W && qy < H;
  in[2] = qx < W && qy + 1 < H;
  in[3] = qx + 1 < W && qy + 1 < H;
  bool done[4];
  float T[4], C0[4], C1[4], C2[4], Dp[4];
  uint32_t last[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    done[k] = !in[k];
    T[k] = 1.0f; C0[k] = C1[k] = C2[k] = Dp[k] = 0.f;
    last[k] = 0;
  }
  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + kRBatch - 1) / kRBatch;
  bool warp_done = __all_sync(0xffffffffu, done[0] && done[1] && done[2] && done[3]);

  // prologue: stage batch 0, prefetch ids of batch 1
  if (tid < n) {
    const uint32_t id = point_list[range.x + tid];
    sm.id[0][tid] = id;
    cp_async16(&sm.a[0][tid], &rec[3 * (size_t)id + 0]);
    cp_async16(&sm.b[0][tid], &rec[3 * (size_t)id + 1]);
    cp_async16(&sm.c[0][tid], &rec[3 * (size_t)id + 2]);
  }
  cp_async_commit();
  uint32_t nxt = 0;
  if (kRBatch + tid < n) nxt = point_list[range.x + kRBatch + tid];

  int cur = 0;
  for (int r = 0; r < rounds; ++r) {
    cp_async_wait_all();
    if (__syncthreads_and(warp_done)) break;
    // stage batch r+1 into the other buffer (free since everybody passed the barrier)
    if ((r + 1) * kRBatch + tid < n) {
      sm.id[cur ^ 1][tid] = nxt;
      cp_async16(&sm.a[cur ^ 1][tid], &rec[3 * (size_t)nxt + 0]);
      cp_async16(&sm.b[cur ^ 1][tid], &rec[3 * (size_t)nxt + 1]);
      cp_async16(&sm.c[cur ^ 1][tid], &rec[3 * (size_t)nxt + 2]);
    }
    cp_async_commit();
    if ((r + 2) * kRBatch + tid < n) nxt = point_list[range.x + (r + 2) * kRBatch + tid];

    const int cnt = min(kRBatch, n - r * kRBatch);
    if (!warp_done) {
#pragma unroll 1
      for (int g = 0; g < kRBatch / 32 && !warp_done; ++g) {
        const int j0 = g * 32 + lane;
        bool hit = false;
        if (j0 < cnt) {
          const float4 e = sm.a[cur][j0];
          hit = (e.x + e.z >= wx0) && (e.x - e.z <= wx1) && (e.y + e.w >= wy0) && (e.y - e.w <= wy1);
        }
        uint32_t m = __ballot_sync(0xffffffffu, hit);
        while (m) {
          const int j = g * 32 + (__ffs(m) - 1);
          m &= m - 1;
          const float4 e = sm.a[cur][j];
          const float4 co = sm.b[cur][j];
          const float dx0 = e.x - fx0, dx1 = e.x - fx1, dy0 = e.y - fy0, dy1 = e.y - fy1;
          float pw[4];
          pw[0] = -0.5f * (co.x * dx0 * dx0 + co.z * dy0 * dy0) - co.y * dx0 * dy0;
          pw[1] = -0.5f * (co.x * dx1 * dx1 + co.z * dy0 * dy0) - co.y * dx1 * dy0;
          pw[2] = -0.5f * (co.x * dx0 * dx0 + co.z * dy1 * dy1) - co.y * dx0 * dy1;
          pw[3] = -0.5f * (co.x * dx1 * dx1 + co.z * dy1 * dy1) - co.y * dx1 * dy1;
          float al[4];
          bool any = false;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            al[k] = fminf(0.99f, co.w * expf(pw[k]));
            const bool c = !done[k] && !(pw[k] > 0.0f) && !(al[k] < 1.0f / 255.0f);
            if (!c) al[k] = -1.f;
            any |= c;
          }
          int touched = 0;
          if (any) {
            const float4 cd = sm.c[cur][j];
            const uint32_t pos = (uint32_t)(r * kRBatch + j + 1);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (al[k] >= 0.f) {
                const float test_T = T[k] * (1 - al[k]);
                if (test_T < 0.0001f) {
                  done[k] = true;
                } else {
                  C0[k] += cd.y * al[k] * T[k];
                  C1[k] += cd.z * al[k] * T[k];
                  C2[k] += cd.w * al[k] * T[k];
                  Dp[k] += cd.x * al[k] * T[k];
                  if (test_T > 0.5f) ++touched;
                  T[k] = test_T;
                  last[k] = pos;
                }
              }
            }
          }
          const int tsum = __reduce_add_sync(0xffffffffu, touched);
          if (tsum && lane == 0) atomicAdd(&n_touched[sm.id[cur][j]], tsum);
        }
        warp_done = __all_sync(0xffffffffu, done[0] && done[1] && done[2] && done[3]);
      }
    }
    cur ^= 1;
  }
  cp_async_wait_all();

  const float b0 = bg[0], b1 = bg[1], b2 = bg[2];
  const size_t HW = (size_t)H * W;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (in[k]) {
      const size_t pix = (size_t)W * (qy + (k >> 1)) + qx + (k & 1);
      final_T[pix] = T[k];
      n_contrib[pix] = last[k];
      out_color[pix] = C0[k] + T[k] * b0;
      out_color[HW + pix] = C1[k] + T[k] * b1;
      out_color[2 * HW + pix] = C2[k] + T[k] * b2;
      out_depth[pix] = Dp[k];
      out_opacity[pix] = 1 - T[k];
    }
  }
}

}  // namespace pg
