// 3DGS backward kernels: per-tile blend backward with warp-shuffle gradient reduction, and
// the fused per-Gaussian backward (conic -> cov3D -> scale/rot, mean2D -> mean3D, depth,
// SH, and the SE(3) pose gradient reduced in-kernel).
//
// Reference behaviour: R3 cuda_rasterizer/backward.cu:563-787 (blend), :150-346
// (computeCov2DCUDA), :418-539 (preprocessCUDA), :350-413 (computeCov3D) and the
// torch.sum over dL_dtau in R3 diff_gaussian_rasterization/__init__.py:152-154.
#pragma once
#include "pg_raster3_fwd.cuh"

#ifndef PG_BWD_MIN_CTAS
#define PG_BWD_MIN_CTAS 1   // 13 (72 registers, one wave at C2) measured slower: 0.157 vs 0.149 ms
#endif
namespace pg {

// Sum 12 per-lane values over the 32 lanes of a warp with a halving butterfly:
// 6+3+2+1+1 = 13 shuffles instead of 12*5 = 60.  The 12 totals end up in 12 distinct
// lanes, which store them to dst[0..11].
__device__ __forceinline__ void warp_reduce12_store(float (&v)[12], int lane, float* dst) {
  const unsigned full = 0xffffffffu;
  const bool u4 = lane & 16;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    const float send = u4 ? v[k] : v[k + 6];
    const float keep = u4 ? v[k + 6] : v[k];
    v[k] = keep + __shfl_xor_sync(full, send, 16);
  }
  const bool u3 = lane & 8;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float send = u3 ? v[k] : v[k + 3];
    const float keep = u3 ? v[k + 3] : v[k];
    v[k] = keep + __shfl_xor_sync(full, send, 8);
  }
  const bool u2 = lane & 4;
  {
    const float r1 = __shfl_xor_sync(full, u2 ? v[0] : v[2], 4);
    const float r2 = __shfl_xor_sync(full, v[1], 4);
    if (u2) {
      v[0] = v[2] + r1;
    } else {
      v[0] += r1;
      v[1] += r2;
    }
  }
  const bool u1 = lane & 2;
  {
    const float send = u2 ? v[0] : (u1 ? v[0] : v[1]);
    const float r = __shfl_xor_sync(full, send, 2);
    v[0] = (u2 ? v[0] : (u1 ? v[1] : v[0])) + r;
  }
  v[0] += __shfl_xor_sync(full, v[0], 1);
  const int index = (u4 ? 6 : 0) + (u3 ? 3 : 0) + (u2 ? 2 : (u1 ? 1 : 0));
  const bool writer = !(lane & 1) && (!u2 || !u1);
  if (writer) dst[index] = v[0];
}

struct RBwdSlab {
  float4 a[2][kRBatch];
  float4 b[2][kRBatch];
  float4 c[2][kRBatch];
  uint32_t id[2][kRBatch];
  float acc[2][kRBatch][12];   // per-warp partial sums of the current batch (plain stores)
  uint32_t flag[2][kRBatch];   // acc[w][j] valid
  uint32_t wmax[2];
};

// Gradient accumulation record per Gaussian (3 x float4, REDG.F32x4 targets):
//   gacc[3i+0] = (dL_dmean2D.x, dL_dmean2D.y, dL_dconic.x, dL_dconic.y)
//   gacc[3i+1] = (dL_dconic.w, dL_dopacity, dL_ddepth, 0)
//   gacc[3i+2] = (dL_dcolor.r, .g, .b, 0)
__global__ void __launch_bounds__(kRThreads, PG_BWD_MIN_CTAS) r3_render_bwd_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H, int gx,
    const float4* __restrict__ rec, const float* __restrict__ bg, const float* __restrict__ final_T,
    const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
    const float* __restrict__ dL_dpix_depth, float4* __restrict__ gacc) {
  __shared__ RBwdSlab sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  const int rx0 = tx * kTile, ry0 = ty * kTile + warp * 8;
  const int qx = rx0 + (lane & 7) * 2, qy = ry0 + (lane >> 3) * 2;
  const float fxs[2] = {(float)qx, (float)(qx + 1)};
  const float fys[2] = {(float)qy, (float)(qy + 1)};
  const float wx0 = (float)rx0, wx1 = (float)min(rx0 + 15, W - 1);
  const float wy0 = (float)ry0, wy1 = (float)min(ry0 + 7, H - 1);
  const size_t HW = (size_t)H * W;

  float T[4], Tfin[4], accA[4], lastB[4], lalpha[4], dpx[4][3], dpd[4], bgdot[4];
  uint32_t lastc[4];
  const float b0 = bg[0], b1 = bg[1], b2 = bg[2];
  uint32_t my_max = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int x = qx + (k & 1), y = qy + (k >> 1);
    const bool in = x < W && y < H;
    const size_t pix = (size_t)W * y + x;
    Tfin[k] = in ? final_T[pix] : 0.f;
    T[k] = Tfin[k];
    lastc[k] = in ? n_contrib[pix] : 0u;
    my_max = max(my_max, lastc[k]);
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      dpx[k][ch] = in ? dL_dpix[ch * HW + pix] : 0.f;
    }
    dpd[k] = in ? dL_dpix_depth[pix] : 0.f;
    accA[k] = 0.f; lastB[k] = 0.f; lalpha[k] = 0.f;
    bgdot[k] = b0 * dpx[k][0] + b1 * dpx[k][1] + b2 * dpx[k][2];
  }
  const uint32_t wmax = __reduce_max_sync(0xffffffffu, my_max);
  if (lane == 0) sm.wmax[warp] = wmax;
  sm.flag[0][tid] = 0;
  sm.flag[1][tid] = 0;
  __syncthreads();
  const uint2 range = ranges[tile];
  // Only the first n_eff list entries can contribute to any pixel of this tile.
  const int n_eff = (int)min(max(sm.wmax[0], sm.wmax[1]), range.y - range.x);
  const int rounds = (n_eff + kRBatch - 1) / kRBatch;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

  // slot t of batch r holds list position n_eff-1-(r*64+t) (back to front)
  if (tid < n_eff) {
    const uint32_t id = point_list[range.x + (n_eff - 1 - tid)];
    sm.id[0][tid] = id;
    cp_async16(&sm.a[0][tid], &rec[3 * (size_t)id + 0]);
    cp_async16(&sm.b[0][tid], &rec[3 * (size_t)id + 1]);
    cp_async16(&sm.c[0][tid], &rec[3 * (size_t)id + 2]);
  }
  cp_async_commit();
  uint32_t nxt = 0;
  if (kRBatch + tid < n_eff) nxt = point_list[range.x + (n_eff - 1 - kRBatch - tid)];

  int cur = 0;
  for (int r = 0; r < rounds; ++r) {
    cp_async_wait_all();
    __syncthreads();
    if ((r + 1) * kRBatch + tid < n_eff) {
      sm.id[cur ^ 1][tid] = nxt;
      cp_async16(&sm.a[cur ^ 1][tid], &rec[3 * (size_t)nxt + 0]);
      cp_async16(&sm.b[cur ^ 1][tid], &rec[3 * (size_t)nxt + 1]);
      cp_async16(&sm.c[cur ^ 1][tid], &rec[3 * (size_t)nxt + 2]);
    }
    cp_async_commit();
    if ((r + 2) * kRBatch + tid < n_eff) nxt = point_list[range.x + (n_eff - 1 - (r + 2) * kRBatch - tid)];

    const int cnt = min(kRBatch, n_eff - r * kRBatch);
#pragma unroll 1
    for (int g = 0; g < kRBatch / 32; ++g) {
      const int j0 = g * 32 + lane;
      bool hit = false;
      if (j0 < cnt) {
        const float4 e = sm.a[cur][j0];
        const uint32_t pos0 = (uint32_t)(n_eff - 1 - (r * kRBatch + j0));
        hit = (pos0 < wmax) && (e.x + e.z >= wx0) && (e.x - e.z <= wx1) && (e.y + e.w >= wy0) &&
              (e.y - e.w <= wy1);
      }
      uint32_t m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int j = g * 32 + (__ffs(m) - 1);
        m &= m - 1;
        const uint32_t pos = (uint32_t)(n_eff - 1 - (r * kRBatch + j));
        const float4 e = sm.a[cur][j];
        const float4 co = sm.b[cur][j];
        float G[4], al[4], ddx[4], ddy[4];
        bool act[4];
        bool any = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          ddx[k] = e.x - fxs[k & 1];
          ddy[k] = e.y - fys[k >> 1];
          const float power = -0.5f * (co.x * ddx[k] * ddx[k] + co.z * ddy[k] * ddy[k]) - co.y * ddx[k] * ddy[k];
          // ex2.approx(power * log2e): ~5 ulp on power in [-5.5, 0] (2 instructions instead of expf's 8).  The forward
          // keeps the accurate expf because n_contrib / the alpha >= 1/255 decisions are compared bit for bit; here a
          // decision can only differ for alpha within a few ulp of 1/255, a <= 0.4 % change of one pixel's terms.
          G[k] = __expf(power);
          al[k] = fminf(0.99f, co.w * G[k]);
          act[k] = (pos < lastc[k]) && !(power > 0.0f) && !(al[k] < 1.0f / 255.0f);
          any |= act[k];
        }
        if (!__any_sync(0xffffffffu, any)) continue;
        float v[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) v[q] = 0.f;
        if (any) {
          const float4 cd = sm.c[cur][j];
          const float col[3] = {cd.y, cd.z, cd.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (act[k]) {
              const float alpha = al[k];
              // one reciprocal serves both divisions by (1 - alpha) of R3 backward.cu:672,719
              // (<= 1.5 ulp from the quotients; far inside the gradient tolerance)
              const float inv1ma = rcp_nr(1.f - alpha);
              T[k] = T[k] * inv1ma;
              const float dchannel_dcolor = alpha * T[k];
              // R3 backward.cu:683-716 keeps, per channel, the colour accumulated behind this Gaussian (accum_rec)
              // and adds (c_ch - accum_rec_ch) * dL_dpixel_ch to dL_dalpha.  Only the dot product with the pixel
              // cotangent is ever used, and the recurrence is linear, so one scalar per pixel carries it:
              // A <- lalpha * B_last + (1 - lalpha) * A with B = sum_ch c_ch * dL_dpixel_ch (depth is a 4th channel).
              const float B = fmaf(cd.x, dpd[k], fmaf(col[2], dpx[k][2], fmaf(col[1], dpx[k][1], col[0] * dpx[k][0])));
              accA[k] = fmaf(lalpha[k], lastB[k] - accA[k], accA[k]);
              lastB[k] = B;
#pragma unroll
              for (int ch = 0; ch < 3; ++ch) v[8 + ch] += dchannel_dcolor * dpx[k][ch];
              v[6] += dchannel_dcolor * dpd[k];
              float dL_dalpha = (B - accA[k]) * T[k];
              lalpha[k] = alpha;
              dL_dalpha += (-Tfin[k] * inv1ma) * bgdot[k];
              // v[0..4] are accumulated without their constant factors (-ddelx_dx, -ddely_dy, -0.5), which the
              // flush applies once per staged Gaussian instead of once per pixel pair.
              const float dL_dG = co.w * dL_dalpha;
              const float hx = dL_dG * (G[k] * ddx[k]), hy = dL_dG * (G[k] * ddy[k]);
              v[0] = fmaf(hy, co.y, fmaf(hx, co.x, v[0]));
              v[1] = fmaf(hx, co.y, fmaf(hy, co.z, v[1]));
              v[2] = fmaf(hx, ddx[k], v[2]);
              v[3] = fmaf(hx, ddy[k], v[3]);
              v[4] = fmaf(hy, ddy[k], v[4]);
              v[5] += G[k] * dL_dalpha;
            }
          }
        }
        warp_reduce12_store(v, lane, &sm.acc[warp][j][0]);
        if (lane == 0) sm.flag[warp][j] = 1;
      }
    }
    __syncthreads();
    // flush slot tid: one owner thread per staged Gaussian, 3 vector reductions to HBM/L2
    {
      const uint32_t f0 = sm.flag[0][tid], f1 = sm.flag[1][tid];
      if (f0 | f1) {
        float s[12];
#pragma unroll
        for (int q = 0; q < 12; ++q) s[q] = (f0 ? sm.acc[0][tid][q] : 0.f) + (f1 ? sm.acc[1][tid][q] : 0.f);
        float4* dst = gacc + 3 * (size_t)sm.id[cur][tid];
        atomicAdd(dst + 0, make_float4(-ddelx_dx * s[0], -ddely_dy * s[1], -0.5f * s[2], -0.5f * s[3]));
        atomicAdd(dst + 1, make_float4(-0.5f * s[4], s[5], s[6], 0.f));
        atomicAdd(dst + 2, make_float4(s[8], s[9], s[10], 0.f));
        sm.flag[0][tid] = 0;
        sm.flag[1][tid] = 0;
      }
    }
    cur ^= 1;
  }
  cp_async_wait_all();
}

// ---- fused per-Gaussian backward ----------------------------------------------------------
// Surfel covariance with the `(mod*surface) ? 0 : scale.z` rule (RS forward.cu:162-202), from a quaternion.
__device__ __forceinline__ void cov3d_surfel_q(float sx, float sy, float sz, float mod, bool surface, float4 q,
                                               float* cov6) {
  const float zs = (mod * (surface ? 1.f : 0.f)) != 0.f ? 0.f : sz;
  // (mod*zs) would rescale: build S explicitly instead of reusing cov3d_from_scale_rot's mod*sz
  M3 S = {{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
  S.c[0][0] = mod * sx;
  S.c[1][1] = mod * sy;
  S.c[2][2] = zs;
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 R;
  R.c[0][0] = 1.f - 2.f * (y * y + z * z); R.c[0][1] = 2.f * (x * y - r * z); R.c[0][2] = 2.f * (x * z + r * y);
  R.c[1][0] = 2.f * (x * y + r * z); R.c[1][1] = 1.f - 2.f * (x * x + z * z); R.c[1][2] = 2.f * (y * z - r * x);
  R.c[2][0] = 2.f * (x * z - r * y); R.c[2][1] = 2.f * (y * z + r * x); R.c[2][2] = 1.f - 2.f * (x * x + y * y);
  M3 Mm = m3_mul(S, R);
  M3 Sg = m3_mul(m3_transpose(Mm), Mm);
  cov6[0] = Sg.c[0][0]; cov6[1] = Sg.c[0][1]; cov6[2] = Sg.c[0][2];
  cov6[3] = Sg.c[1][1]; cov6[4] = Sg.c[1][2]; cov6[5] = Sg.c[2][2];
}

struct R3BwdPreArgs {
  int P, D, M, W, H;
  const float *means, *scales, *rots, *shs, *cov3D_pre, *view, *proj, *proj_raw, *campos;
  const int* radii;
  const uint8_t* clamped;
  const float4* gacc;
  float mod, tanx, tany, fx, fy;
  float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales, *dL_drots;
  float* dL_dtau_full;  // [P,6] or null
  double* tau_acc;      // [6] + counter at [7] (as unsigned long long)
  float* dL_dtau;       // [6]
  // surfel variant only (pg_rasters_fwd.cuh): normal/zx record, zy, config flags
  const float4* ext;
  const float* zyv;
  const float* config;
};

__device__ __forceinline__ float3 cross3(const float3& a, const float3& b) {
  return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// SURFEL = true adds the differences of RS cuda_rasterizer/backward.cu:362-452,457-582: 4-float4
// gradient records, the z-scale rule in the recomputed covariance, the per-pixel-depth term of
// dL_dmean2D, dL_dnormal -> last row of dL_dRt, dL_dscale.z = 0 when `surface`.
template <bool SURFEL>
__global__ void __launch_bounds__(256) r3_preprocess_bwd_kernel(R3BwdPreArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  float tau[6] = {0, 0, 0, 0, 0, 0};
  if (idx < a.P) {
    float dmean3[3] = {0, 0, 0}, dcov[6] = {0, 0, 0, 0, 0, 0}, dscale[3] = {0, 0, 0}, drot[4] = {0, 0, 0, 0};
    float dmean2[2] = {0, 0}, dcol[3] = {0, 0, 0}, dopac = 0.f;
    const bool vis = a.radii[idx] > 0;
    if (vis) {
      constexpr int kStride = SURFEL ? 4 : 3;
      const float4 g0 = a.gacc[kStride * (size_t)idx + 0];
      const float4 g1 = a.gacc[kStride * (size_t)idx + 1];
      const float4 g2 = a.gacc[kStride * (size_t)idx + 2];
      dmean2[0] = g0.x; dmean2[1] = g0.y;
      bool surface = false;
      float3 dL_dcamN = make_float3(0.f, 0.f, 0.f);
      if (SURFEL) {
        surface = a.config[0] > 0;
        // per-pixel-depth term (RS backward.cu:894-897): -sum(dL_dpixD) * (zx, zy)
        dmean2[0] += -g1.w * a.ext[idx].w;
        dmean2[1] += -g1.w * a.zyv[idx];
        const float4 g3 = a.gacc[kStride * (size_t)idx + 3];
        dL_dcamN = make_float3(g3.x, g3.y, g3.z);
      }
      const float3 dL_dconic = make_float3(g0.z, g0.w, g1.x);
      dopac = g1.y;
      const float dL_ddepth = g1.z;
      dcol[0] = g2.x; dcol[1] = g2.y; dcol[2] = g2.z;

      const float3 mean = make_float3(a.means[3 * idx], a.means[3 * idx + 1], a.means[3 * idx + 2]);
      float cov6[6];
      float4 q = make_float4(0, 0, 0, 0);
      float3 sc = make_float3(0, 0, 0);
      if (a.cov3D_pre) {
#pragma unroll
        for (int k = 0; k < 6; ++k) cov6[k] = a.cov3D_pre[6 * idx + k];
      } else {
        q = reinterpret_cast<const float4*>(a.rots)[idx];
        sc = make_float3(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2]);
        if (SURFEL) cov3d_surfel_q(sc.x, sc.y, sc.z, a.mod, surface, q, cov6);
        else cov3d_from_scale_rot(sc.x, sc.y, sc.z, a.mod, q, cov6);
      }
      // ---- conic -> cov2D -> cov3D, T, J, t (computeCov2DCUDA) ----
      Cov2DAux ax;
      const float3 cv = cov2d_ewa(mean, a.fx, a.fy, a.tanx, a.tany, cov6, a.view, &ax);
      const float limx = 1.3f * a.tanx, limy = 1.3f * a.tany;
      const float x_grad_mul = (ax.txtz < -limx || ax.txtz > limx) ? 0.f : 1.f;
      const float y_grad_mul = (ax.tytz < -limy || ax.tytz > limy) ? 0.f : 1.f;
      const float ca = cv.x, cb = cv.y, cc = cv.z;
      const float denom = ca * cc - cb * cb;
      float dL_da = 0, dL_db = 0, dL_dc = 0;
      const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
      const M3& T = ax.T;
      const M3& V = ax.Vrk;
      const M3& Wm = ax.Wm;
      const M3& J = ax.J;
      if (denom2inv != 0) {
        dL_da = denom2inv * (-cc * cc * dL_dconic.x + 2 * cb * cc * dL_dconic.y + (denom - ca * cc) * dL_dconic.z);
        dL_dc = denom2inv * (-ca * ca * dL_dconic.z + 2 * ca * cb * dL_dconic.y + (denom - ca * cc) * dL_dconic.x);
        dL_db = denom2inv * 2 * (cb * cc * dL_dconic.x - (denom + 2 * cb * cb) * dL_dconic.y + ca * cb * dL_dconic.z);
        dcov[0] = (T.c[0][0] * T.c[0][0] * dL_da + T.c[0][0] * T.c[1][0] * dL_db + T.c[1][0] * T.c[1][0] * dL_dc);
        dcov[3] = (T.c[0][1] * T.c[0][1] * dL_da + T.c[0][1] * T.c[1][1] * dL_db + T.c[1][1] * T.c[1][1] * dL_dc);
        dcov[5] = (T.c[0][2] * T.c[0][2] * dL_da + T.c[0][2] * T.c[1][2] * dL_db + T.c[1][2] * T.c[1][2] * dL_dc);
        dcov[1] = 2 * T.c[0][0] * T.c[0][1] * dL_da + (T.c[0][0] * T.c[1][1] + T.c[0][1] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][1] * dL_dc;
        dcov[2] = 2 * T.c[0][0] * T.c[0][2] * dL_da + (T.c[0][0] * T.c[1][2] + T.c[0][2] * T.c[1][0]) * dL_db + 2 * T.c[1][0] * T.c[1][2] * dL_dc;
        dcov[4] = 2 * T.c[0][2] * T.c[0][1] * dL_da + (T.c[0][1] * T.c[1][2] + T.c[0][2] * T.c[1][1]) * dL_db + 2 * T.c[1][1] * T.c[1][2] * dL_dc;
      }
      // rows of (T Vrk): tv0[k] = sum_m T[0][m] V[k][m], tv1[k] = sum_m T[1][m] V[k][m]
      float tv0[3], tv1[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        tv0[k] = T.c[0][0] * V.c[k][0] + T.c[0][1] * V.c[k][1] + T.c[0][2] * V.c[k][2];
        tv1[k] = T.c[1][0] * V.c[k][0] + T.c[1][1] * V.c[k][1] + T.c[1][2] * V.c[k][2];
      }
      float dT0[3], dT1[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        dT0[k] = 2 * tv0[k] * dL_da + tv1[k] * dL_db;
        dT1[k] = 2 * tv1[k] * dL_dc + tv0[k] * dL_db;
      }
      const float dL_dJ00 = Wm.c[0][0] * dT0[0] + Wm.c[0][1] * dT0[1] + Wm.c[0][2] * dT0[2];
      const float dL_dJ02 = Wm.c[2][0] * dT0[0] + Wm.c[2][1] * dT0[1] + Wm.c[2][2] * dT0[2];
      const float dL_dJ11 = Wm.c[1][0] * dT1[0] + Wm.c[1][1] * dT1[1] + Wm.c[1][2] * dT1[2];
      const float dL_dJ12 = Wm.c[2][0] * dT1[0] + Wm.c[2][1] * dT1[1] + Wm.c[2][2] * dT1[2];
      const float3 t = ax.t;
      const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
      const float dL_dtx = x_grad_mul * -a.fx * tz2 * dL_dJ02;
      const float dL_dty = y_grad_mul * -a.fy * tz2 * dL_dJ12;
      const float dL_dtz = -a.fx * tz2 * dL_dJ00 - a.fy * tz2 * dL_dJ11 + (2 * a.fx * t.x) * tz3 * dL_dJ02 +
                           (2 * a.fy * t.y) * tz3 * dL_dJ12;
      const float3 dLt = make_float3(dL_dtx, dL_dty, dL_dtz);
      // pose, part 1: rho via identity, theta via -skew(t) (clamped t), plus rotation part of W
      tau[0] += dLt.x; tau[1] += dLt.y; tau[2] += dLt.z;
      {
        const float3 th = cross3(t, dLt);
        tau[3] += th.x; tau[4] += th.y; tau[5] += th.z;
      }
      {
        const float3 dm = xform_vec_4x3_T(dLt, a.view);
        dmean3[0] = dm.x; dmean3[1] = dm.y; dmean3[2] = dm.z;
      }
      {
        const float3 dW1 = make_float3(J.c[0][0] * dT0[0], J.c[1][1] * dT1[0], J.c[0][2] * dT0[0] + J.c[1][2] * dT1[0]);
        const float3 dW2 = make_float3(J.c[0][0] * dT0[1], J.c[1][1] * dT1[1], J.c[0][2] * dT0[1] + J.c[1][2] * dT1[1]);
        const float3 dW3 = make_float3(J.c[0][0] * dT0[2], J.c[1][1] * dT1[2], J.c[0][2] * dT0[2] + J.c[1][2] * dT1[2]);
        const float3 c1 = make_float3(a.view[0], a.view[1], a.view[2]);
        const float3 c2 = make_float3(a.view[4], a.view[5], a.view[6]);
        const float3 c3 = make_float3(a.view[8], a.view[9], a.view[10]);
        const float3 x1 = cross3(c1, dW1), x2 = cross3(c2, dW2), x3 = cross3(c3, dW3);
        tau[3] += x1.x + x2.x + x3.x;
        tau[4] += x1.y + x2.y + x3.y;
        tau[5] += x1.z + x2.z + x3.z;
      }
      // ---- mean2D -> mean3D and pose, part 2 (preprocessCUDA bwd) ----
      const float* proj = a.proj;
      const float4 m_hom = xform_point_4x4(mean, proj);
      const float m_w = 1.0f / (m_hom.w + 0.0000001f);
      const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
      const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
      dmean3[0] += (proj[0] * m_w - proj[3] * mul1) * dmean2[0] + (proj[1] * m_w - proj[3] * mul2) * dmean2[1];
      dmean3[1] += (proj[4] * m_w - proj[7] * mul1) * dmean2[0] + (proj[5] * m_w - proj[7] * mul2) * dmean2[1];
      dmean3[2] += (proj[8] * m_w - proj[11] * mul1) * dmean2[0] + (proj[9] * m_w - proj[11] * mul2) * dmean2[1];
      {
        const float alpha = 1.0f * m_w;
        const float beta = -m_hom.x * m_w * m_w;
        const float gamma = -m_hom.y * m_w * m_w;
        const float pa = a.proj_raw[0], pb = a.proj_raw[5], pe = a.proj_raw[11];
        const float3 pC = xform_point_4x3(mean, a.view);
        const float3 d1 = make_float3(alpha * pa, 0.f, beta * pe);
        const float3 d2 = make_float3(0.f, alpha * pb, gamma * pe);
        const float3 d1t = cross3(pC, d1), d2t = cross3(pC, d2);
        tau[0] += dmean2[0] * d1.x + dmean2[1] * d2.x;
        tau[1] += dmean2[0] * d1.y + dmean2[1] * d2.y;
        tau[2] += dmean2[0] * d1.z + dmean2[1] * d2.z;
        tau[3] += dmean2[0] * d1t.x + dmean2[1] * d2t.x;
        tau[4] += dmean2[0] * d1t.y + dmean2[1] * d2t.y;
        tau[5] += dmean2[0] * d1t.z + dmean2[1] * d2t.z;
        // depth = p_view.z
        dmean3[0] += dL_ddepth * a.view[2];
        dmean3[1] += dL_ddepth * a.view[6];
        dmean3[2] += dL_ddepth * a.view[10];
        tau[2] += dL_ddepth;
        tau[3] += dL_ddepth * pC.y;
        tau[4] += dL_ddepth * -pC.x;
      }
      if (a.shs) {
        F3 dm = sh_to_rgb_backward(idx, a.D, a.M, {mean.x, mean.y, mean.z},
                                   {a.campos[0], a.campos[1], a.campos[2]}, a.shs, a.clamped,
                                   {dcol[0], dcol[1], dcol[2]}, a.dL_dsh);
        dmean3[0] += dm.x; dmean3[1] += dm.y; dmean3[2] += dm.z;
        tau[0] += -dm.x; tau[1] += -dm.y; tau[2] += -dm.z;
      }
      // ---- cov3D -> scale, rotation (computeCov3D bwd) ----
      if (a.scales) {
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        M3 R;
        R.c[0][0] = 1.f - 2.f * (y * y + z * z); R.c[0][1] = 2.f * (x * y - r * z); R.c[0][2] = 2.f * (x * z + r * y);
        R.c[1][0] = 2.f * (x * y + r * z); R.c[1][1] = 1.f - 2.f * (x * x + z * z); R.c[1][2] = 2.f * (y * z - r * x);
        R.c[2][0] = 2.f * (x * z - r * y); R.c[2][1] = 2.f * (y * z + r * x); R.c[2][2] = 1.f - 2.f * (x * x + y * y);
        const float s0 = a.mod * sc.x, s1 = a.mod * sc.y, s2 = a.mod * sc.z;
        M3 S = {{{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}};
        S.c[0][0] = s0; S.c[1][1] = s1; S.c[2][2] = s2;
        const M3 Mm = m3_mul(S, R);
        M3 dSig;
        dSig.c[0][0] = dcov[0];        dSig.c[0][1] = 0.5f * dcov[1]; dSig.c[0][2] = 0.5f * dcov[2];
        dSig.c[1][0] = 0.5f * dcov[1]; dSig.c[1][1] = dcov[3];        dSig.c[1][2] = 0.5f * dcov[4];
        dSig.c[2][0] = 0.5f * dcov[2]; dSig.c[2][1] = 0.5f * dcov[4]; dSig.c[2][2] = dcov[5];
        M3 M2;
#pragma unroll
        for (int j = 0; j < 3; ++j)
#pragma unroll
          for (int i = 0; i < 3; ++i) M2.c[j][i] = 2.0f * Mm.c[j][i];
        const M3 dM = m3_mul(M2, dSig);
        const M3 Rt = m3_transpose(R);
        M3 dMt = m3_transpose(dM);
        dscale[0] = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
        dscale[1] = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
        dscale[2] = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
        if (SURFEL && surface) dscale[2] = 0.f;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          dMt.c[0][i] *= s0;
          dMt.c[1][i] *= s1;
          dMt.c[2][i] *= s2;
        }
        if (SURFEL) {  // loss gradient of the normal output w.r.t. R (RS backward.cu:427-436)
          const float3 dwn = xform_vec_4x3_T(dL_dcamN, a.view);
          dMt.c[2][0] += dwn.x;
          dMt.c[2][1] += dwn.y;
          dMt.c[2][2] += dwn.z;
        }
        drot[0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
        drot[1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
        drot[2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
        drot[3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
      }
    } else if (a.dL_dsh) {
      for (int k = 0; k < a.M * 3; ++k) a.dL_dsh[(size_t)idx * a.M * 3 + k] = 0.f;
    }
    // every output row is written (callers need not pre-zero)
    a.dL_dmeans2D[3 * idx + 0] = dmean2[0]; a.dL_dmeans2D[3 * idx + 1] = dmean2[1]; a.dL_dmeans2D[3 * idx + 2] = 0.f;
    a.dL_dcolors[3 * idx + 0] = dcol[0]; a.dL_dcolors[3 * idx + 1] = dcol[1]; a.dL_dcolors[3 * idx + 2] = dcol[2];
    a.dL_dopacity[idx] = dopac;
#pragma unroll
    for (int k = 0; k < 3; ++k) a.dL_dmeans3D[3 * idx + k] = dmean3[k];
#pragma unroll
    for (int k = 0; k < 6; ++k) a.dL_dcov3D[6 * idx + k] = dcov[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) a.dL_dscales[3 * idx + k] = dscale[k];
    reinterpret_cast<float4*>(a.dL_drots)[idx] = make_float4(drot[0], drot[1], drot[2], drot[3]);
    if (a.dL_dtau_full) {
#pragma unroll
      for (int k = 0; k < 6; ++k) a.dL_dtau_full[6 * (size_t)idx + k] = tau[k];
    }
  }
  // ---- pose gradient: warp shuffle -> block -> 6 fp64 atomics; last block publishes ----
  __shared__ float wsum[8][6];
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float s = tau[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5][k] = s;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += wsum[w][threadIdx.x];
    atomicAdd(&a.tau_acc[threadIdx.x], (double)s);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long* counter = reinterpret_cast<unsigned long long*>(a.tau_acc + 7);
    const unsigned long long done = atomicAdd(counter, 1ull);
    is_last = (done == (unsigned long long)gridDim.x - 1);
  }
  __syncthreads();
  if (is_last && threadIdx.x < 6) {
    __threadfence();
    a.dL_dtau[threadIdx.x] = (float)atomicAdd(&a.tau_acc[threadIdx.x], 0.0);
  }
}

}  // namespace pg
