// Gaussian-surfel forward kernels: per-surfel preprocess and the per-tile alpha blend that also
// renders the normal map, the per-pixel (local-homography) depth and `contributions`.
//
// Reference behaviour: RS = submodules/diff-gaussian-surfel-rasterization-w-pose,
// cuda_rasterizer/forward.cu:205-370 (preprocessCUDA), :375-624 (renderCUDA),
// auxiliary.h:146-245 (culls), :328-434 (local_homo / depth_differencing).
//
// Record layout (on top of the three float4 of the 3DGS record, pg_raster3_fwd.cuh):
//   ext[i] = (n_view.x, n_view.y, n_view.z, zx)      zyv[i] = zy
// where depth_differencing(d, Jinv).z == d.x * zx + d.y * zy with
//   zx = Jinv[0] * u0.z + Jinv[2] * u1.z,  zy = Jinv[1] * u0.z + Jinv[3] * u1.z
// (the reference stores the ten floats Jinv[0..3], u0, u1 per surfel and re-derives the same
// two products per pixel-surfel pair: RS auxiliary.h:427-434, backward.cu:894-897).
#pragma once
#include "pg_raster3_fwd.cuh"

namespace pg {

__host__ __device__ __forceinline__ float3 xform_vec_4x3(const float3& p, const float* m) {
  float3 r;
  r.x = m[0] * p.x + m[4] * p.y + m[8] * p.z;
  r.y = m[1] * p.x + m[5] * p.y + m[9] * p.z;
  r.z = m[2] * p.x + m[6] * p.y + m[10] * p.z;
  return r;
}

// GLM-layout rotation matrix of an un-normalised quaternion (RS forward.cu:141-156).
__device__ __forceinline__ M3 quat_to_rotmat(float4 q) {
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
  return R;
}

// Sigma = (S R)^T (S R) with the surfel z-scale rule `S[2][2] = mod * surface ? 0 : scale.z`
// (RS forward.cu:162-202: the ternary binds as (mod*surface) ? 0 : scale.z).
__device__ __forceinline__ void cov3d_surfel(float sx, float sy, float sz, float mod, bool surface, const M3& R,
                                             float* cov6) {
  M3 S = {{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
  S.c[0][0] = mod * sx;
  S.c[1][1] = mod * sy;
  S.c[2][2] = (mod * (surface ? 1.f : 0.f)) != 0.f ? 0.f : sz;
  M3 Mm = m3_mul(S, R);
  M3 Sg = m3_mul(m3_transpose(Mm), Mm);
  cov6[0] = Sg.c[0][0];
  cov6[1] = Sg.c[0][1];
  cov6[2] = Sg.c[0][2];
  cov6[3] = Sg.c[1][1];
  cov6[4] = Sg.c[1][2];
  cov6[5] = Sg.c[2][2];
}

__device__ __forceinline__ float normalize3(float* v) {
  const float mod = fmaxf(sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), (float)0.00000001);
  v[0] /= mod;
  v[1] /= mod;
  v[2] /= mod;
  return mod;
}

// Local homography between the screen plane and the surfel's tangent plane
// (RS auxiliary.h:328-425).  Returns true when the surfel is seen at a grazing angle (culled).
// On success zx/zy receive the two products the blend needs (see the header comment).
__device__ __forceinline__ bool local_homography(const float3& pv, const float3& n, float fx, float fy,
                                                 const float3& ax0, const float3& ax1, float& zx, float& zy) {
  const float px = pv.x / pv.z, py = pv.y / pv.z;
  const float S_fix = 1000, Svp = (fx + fy) / 2;
  float d0[3], d1[3];
  d0[0] = px + 1 / S_fix;
  d0[1] = py;
  d0[2] = 1;
  const float d0_mod = normalize3(d0);
  d1[0] = px;
  d1[1] = py + 1 / S_fix;
  d1[2] = 1;
  const float d1_mod = normalize3(d1);
  const float thrsh = 0.01;
  const float prj0 = d0[0] * n.x + d0[1] * n.y + d0[2] * n.z;
  const float prj1 = d1[0] * n.x + d1[1] * n.y + d1[2] * n.z;
  if ((fabsf(prj0 / d0_mod) < thrsh) || (fabsf(prj1 / d1_mod) < thrsh)) return true;
  const float t_temp = pv.x * n.x + pv.y * n.y + pv.z * n.z;
  const float t0 = t_temp / prj0, t1 = t_temp / prj1;
  float xu0[3], xu1[3];
  xu0[0] = d0[0] * t0 - pv.x;
  xu0[1] = d0[1] * t0 - pv.y;
  xu0[2] = d0[2] * t0 - pv.z;
  xu1[0] = d1[0] * t1 - pv.x;
  xu1[1] = d1[1] * t1 - pv.y;
  xu1[2] = d1[2] * t1 - pv.z;
  float J[4];
  J[0] = xu0[0] * ax0.x + xu0[1] * ax0.y + xu0[2] * ax0.z;
  J[1] = xu1[0] * ax0.x + xu1[1] * ax0.y + xu1[2] * ax0.z;
  J[2] = xu0[0] * ax1.x + xu0[1] * ax1.y + xu0[2] * ax1.z;
  J[3] = xu1[0] * ax1.x + xu1[1] * ax1.y + xu1[2] * ax1.z;
  const float sc = Svp / S_fix;
  J[0] /= sc;
  J[1] /= sc;
  J[2] /= sc;
  J[3] /= sc;
  // (Jinv[6]*Jinv[0] + Jinv[9]*Jinv[2], Jinv[6]*Jinv[1] + Jinv[9]*Jinv[3]) with Jinv[6]=u0.z, Jinv[9]=u1.z
  zx = ax0.z * J[0] + ax1.z * J[2];
  zy = ax0.z * J[1] + ax1.z * J[3];
  return false;
}

struct RSPreArgs {
  int P, D, M, W, H, gx, gy;
  const float *means, *scales, *rots, *opac, *shs, *cov3D_pre, *colors_pre, *view, *proj, *campos;
  const float *patch_bbox, *config;
  float mod, tanx, tany, fx, fy;
  int prefiltered;
};

__device__ __forceinline__ void rs_trap_if_prefiltered(int prefiltered) {
  if (prefiltered) {
    printf("Point is filtered although prefiltered is set. This shouldn't happen!");
    __trap();
  }
}

__global__ void __launch_bounds__(256) rs_preprocess_kernel(RSPreArgs a, int* __restrict__ radii,
                                                            float4* __restrict__ rec, float4* __restrict__ ext,
                                                            float* __restrict__ zyv, uint32_t* __restrict__ dkey,
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
      const float4 ph = xform_point_4x4(p, a.proj);
      const float pw = 1.0f / (ph.w + 0.0000001f);
      const float3 pp = make_float3(ph.x * pw, ph.y * pw, ph.z * pw);
      const float3 pv = xform_point_4x3(p, a.view);
      const float px = ndc_to_pix(pp.x, a.W), py = ndc_to_pix(pp.y, a.H);
      {  // in_frustum2 (RS auxiliary.h:146-175)
        const float x0 = a.patch_bbox[1], y0 = a.patch_bbox[0], x1 = a.patch_bbox[3], y1 = a.patch_bbox[2];
        const float w = x1 - x0, h = y1 - y0;
        const float expand = 0.05;
        if ((double)pv.z < 0.01 || px < x0 - w * expand || px >= x1 + w * expand || py < y0 - h * expand ||
            py >= y1 + h * expand) {
          rs_trap_if_prefiltered(a.prefiltered);
          break;
        }
      }
      const bool surface = a.config[0] > 0, pix_depth = a.config[2] > 0, front_only = a.config[4] > 0;
      const M3 R = quat_to_rotmat(reinterpret_cast<const float4*>(a.rots)[idx]);
      float3 nv = make_float3(0.f, 0.f, 0.f);
      float zx = 0.f, zy = 0.f;
      if (surface) {
        nv = xform_vec_4x3(make_float3(R.c[0][2], R.c[1][2], R.c[2][2]), a.view);
        const float3 ax0 = xform_vec_4x3(make_float3(R.c[0][0], R.c[1][0], R.c[2][0]), a.view);
        const float3 ax1 = xform_vec_4x3(make_float3(R.c[0][1], R.c[1][1], R.c[2][1]), a.view);
        if (front_only) {  // front_facing (RS auxiliary.h:210-245)
          const float dot = pv.x * nv.x + pv.y * nv.y + pv.z * nv.z;
          if ((double)dot > -0.01) {
            rs_trap_if_prefiltered(a.prefiltered);
            break;
          }
        }
        if (pix_depth) {
          if (local_homography(pv, nv, a.fx, a.fy, ax0, ax1, zx, zy)) break;
        }
      }
      float cov6[6];
      if (a.cov3D_pre) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cov6[k] = a.cov3D_pre[6 * idx + k];
      } else {
        cov3d_surfel(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2], a.mod, surface, R, cov6);
      }
      // RS computeCov2D takes the view-space mean (forward.cu:81); xform(p, view) is bit-identical to pv
      const float3 cov = cov2d_ewa(p, a.fx, a.fy, a.tanx, a.tany, cov6, a.view, nullptr);
      const float det = (cov.x * cov.z - cov.y * cov.y);
      if (det == 0.0f) break;
      const float det_inv = 1.f / det;
      const float3 conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);
      const float mid = 0.5f * (cov.x + cov.z);
      const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
      const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
      const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
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
      ext[idx] = make_float4(nv.x, nv.y, nv.z, zx);
      zyv[idx] = zy;
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
  preprocess_block_tail(my_tiles, my_key, total);
}

// checkFrustum of the surfel rasterizer (RS rasterizer_impl.cu:54-66 / auxiliary.h:177-208).
__global__ void __launch_bounds__(256) rs_mark_visible_kernel(int P, const float* __restrict__ means,
                                                              const float* __restrict__ view,
                                                              const float* __restrict__ proj,
                                                              uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float3 p = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
  const float4 ph = xform_point_4x4(p, proj);
  const float pw = 1.0f / (ph.w + 0.0000001f);
  const float ppx = ph.x * pw, ppy = ph.y * pw;
  const float3 pv = xform_point_4x3(p, view);
  const float wr = 1.05f;
  const bool out = (pv.z <= 0.1f) || (ppx < -wr) || (ppx > wr) || (ppy < -wr) || (ppy > wr);
  present[idx] = out ? 0 : 1;
}

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gsrc));
}

struct RSSlab {
  float4 a[2][kRBatch];  // px,py,hx,hy
  float4 b[2][kRBatch];  // conic + opacity
  float4 c[2][kRBatch];  // depth,r,g,b
  float4 e[2][kRBatch];  // normal, zx
  float zy[2][kRBatch];
  uint32_t id[2][kRBatch];
};

__device__ __forceinline__ void rs_stage(RSSlab& sm, int buf, int slot, uint32_t id, const float4* __restrict__ rec,
                                         const float4* __restrict__ ext, const float* __restrict__ zyv) {
  sm.id[buf][slot] = id;
  cp_async16(&sm.a[buf][slot], &rec[3 * (size_t)id + 0]);
  cp_async16(&sm.b[buf][slot], &rec[3 * (size_t)id + 1]);
  cp_async16(&sm.c[buf][slot], &rec[3 * (size_t)id + 2]);
  cp_async16(&sm.e[buf][slot], &ext[id]);
  cp_async4(&sm.zy[buf][slot], &zyv[id]);
}

// Same CTA shape as the 3DGS blend (pg_raster3_fwd.cuh): one 16x16 tile, 2 warps, 2x2 quad per lane.
__global__ void __launch_bounds__(kRThreads) rs_render_fwd_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H, int gx,
    const float4* __restrict__ rec, const float4* __restrict__ ext, const float* __restrict__ zyv,
    const float* __restrict__ bg, const float* __restrict__ config, float* __restrict__ out_color,
    float* __restrict__ out_normal, float* __restrict__ out_depth, float* __restrict__ out_opac,
    float* __restrict__ final_T, float* __restrict__ final_D, uint32_t* __restrict__ n_contrib,
    float* __restrict__ contributions) {
  __shared__ RSSlab sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_wait();   // tail of the binning chain (programmatic dependent launch)
  const int tile = blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  const int rx0 = tx * kTile, ry0 = ty * kTile + warp * 8;
  const int qx = rx0 + (lane & 7) * 2, qy = ry0 + (lane >> 3) * 2;
  const float fx0 = (float)qx, fx1 = (float)(qx + 1), fy0 = (float)qy, fy1 = (float)(qy + 1);
  const float wx0 = (float)rx0, wx1 = (float)min(rx0 + 15, W - 1);
  const float wy0 = (float)ry0, wy1 = (float)min(ry0 + 7, H - 1);
  const bool surface = config[0] > 0, normalize_depth = config[1] > 0;
  bool in[4];
  in[0] = qx < W && qy < H;
  in[1] = qx + 1 < W && qy < H;
  in[2] = qx < W && qy + 1 < H;
  in[3] = qx + 1 < W && qy + 1 < H;
  bool done[4];
  float T[4], C0[4], C1[4], C2[4], Dp[4], N0[4], N1[4], N2[4];
  uint32_t last[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    done[k] = !in[k];
    T[k] = 1.0f;
    C0[k] = C1[k] = C2[k] = Dp[k] = N0[k] = N1[k] = N2[k] = 0.f;
    last[k] = 0;
  }
  const uint2 range = ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + kRBatch - 1) / kRBatch;
  bool warp_done = __all_sync(0xffffffffu, done[0] && done[1] && done[2] && done[3]);

  if (tid < n) rs_stage(sm, 0, tid, point_list[range.x + tid], rec, ext, zyv);
  cp_async_commit();
  uint32_t nxt = 0;
  if (kRBatch + tid < n) nxt = point_list[range.x + kRBatch + tid];

  int cur = 0;
  for (int r = 0; r < rounds; ++r) {
    cp_async_wait_all();
    if (__syncthreads_and(warp_done)) break;
    if ((r + 1) * kRBatch + tid < n) rs_stage(sm, cur ^ 1, tid, nxt, rec, ext, zyv);
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
          const float dx[2] = {e.x - fx0, e.x - fx1};
          const float dy[2] = {e.y - fy0, e.y - fy1};
          float al[4];
          bool any = false;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float ddx = dx[k & 1], ddy = dy[k >> 1];
            const float dist = (co.x * ddx * ddx + co.z * ddy * ddy) + 2 * co.y * ddx * ddy;
            const float power = -0.5f * dist;
            al[k] = fminf(0.99f, co.w * expf(power));
            const bool c = !done[k] && !(power > 0.0f) && !(al[k] < 1.0f / 255.0f);
            if (!c) al[k] = -1.f;
            any |= c;
          }
          float wsum = 0.f;
          if (any) {
            const float4 cd = sm.c[cur][j];
            const float4 ne = sm.e[cur][j];
            const float zyj = sm.zy[cur][j];
            const uint32_t pos = (uint32_t)(r * kRBatch + j + 1);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (al[k] >= 0.f) {
                const float test_T = T[k] * (1 - al[k]);
                if (test_T < 0.0001f) {
                  done[k] = true;
                } else {
                  const float w = al[k] * T[k];
                  // depth_differencing(d, Jinv).z (zx = zy = 0 unless surface && per-pixel depth)
                  const float dpix = cd.x - (dx[k & 1] * ne.w + dy[k >> 1] * zyj);
                  Dp[k] += dpix * w;
                  C0[k] += cd.y * w;
                  C1[k] += cd.z * w;
                  C2[k] += cd.w * w;
                  if (surface) {
                    N0[k] += ne.x * w;
                    N1[k] += ne.y * w;
                    N2[k] += ne.z * w;
                  }
                  if (test_T > 0.5f) wsum += w;
                  T[k] = test_T;
                  last[k] = pos;
                }
              }
            }
          }
          // contributions (RS forward.cu:586-588): one float atomic per (warp, surfel)
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
          if (lane == 0 && wsum != 0.f) atomicAdd(&contributions[sm.id[cur][j]], wsum);
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
      const float Tk = fminf((float)(1 - 0.000001), T[k]);  // RS forward.cu:608
      final_T[pix] = Tk;
      n_contrib[pix] = last[k];
      out_color[pix] = C0[k] + Tk * b0;
      out_color[HW + pix] = C1[k] + Tk * b1;
      out_color[2 * HW + pix] = C2[k] + Tk * b2;
      out_normal[pix] = surface ? N0[k] : 0.f;
      out_normal[HW + pix] = surface ? N1[k] : 0.f;
      out_normal[2 * HW + pix] = surface ? N2[k] : 0.f;
      out_depth[pix] = normalize_depth ? Dp[k] / (1 - Tk) : Dp[k] + Tk * 10;
      out_opac[pix] = 1 - Tk;
      if (normalize_depth) final_D[pix] = Dp[k];
    }
  }
}

}  // namespace pg
