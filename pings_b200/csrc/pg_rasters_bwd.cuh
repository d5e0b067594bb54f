// Gaussian-surfel backward kernels: per-tile blend backward (colour, normal, depth, opacity
// cotangents) with warp-shuffle gradient reduction.  The per-surfel backward (conic -> cov3D ->
// scale/rot, mean2D -> mean3D, normal -> rotation, pose gradient) is the templated kernel in
// pg_raster3_bwd.cuh.
//
// Reference behaviour: RS cuda_rasterizer/backward.cu:585-914 (renderCUDA; per-thread atomics
// there), :160-358 (computeCov2DCUDA), :362-452 (computeCov3D), :457-582 (preprocessCUDA).
#pragma once
#include "pg_raster3_bwd.cuh"
#include "pg_rasters_fwd.cuh"

namespace pg {

// Sum 16 per-lane values over the 32 lanes of a warp with a halving butterfly (8+4+2+1+1 = 16
// shuffles); total k ends up in lane 2*bitrev-ish index below and is stored to dst[k].
__device__ __forceinline__ void warp_reduce16_store(float (&v)[16], int lane, float* dst) {
  const unsigned full = 0xffffffffu;
  const bool u4 = lane & 16;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float send = u4 ? v[k] : v[k + 8];
    const float keep = u4 ? v[k + 8] : v[k];
    v[k] = keep + __shfl_xor_sync(full, send, 16);
  }
  const bool u3 = lane & 8;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float send = u3 ? v[k] : v[k + 4];
    const float keep = u3 ? v[k + 4] : v[k];
    v[k] = keep + __shfl_xor_sync(full, send, 8);
  }
  const bool u2 = lane & 4;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const float send = u2 ? v[k] : v[k + 2];
    const float keep = u2 ? v[k + 2] : v[k];
    v[k] = keep + __shfl_xor_sync(full, send, 4);
  }
  const bool u1 = lane & 2;
  {
    const float send = u1 ? v[0] : v[1];
    const float keep = u1 ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(full, send, 2);
  }
  v[0] += __shfl_xor_sync(full, v[0], 1);
  const int index = (u4 ? 8 : 0) + (u3 ? 4 : 0) + (u2 ? 2 : 0) + (u1 ? 1 : 0);
  if (!(lane & 1)) dst[index] = v[0];
}

template <int NW>
struct RSBwdSlab {
  RSSlab in;
  float acc[NW][kRBatch][16];  // per-warp partial sums of the current batch
  uint32_t flag[NW][kRBatch];
  uint32_t wmax[NW];
};

// Gradient accumulation record per surfel (4 x float4):
//   gacc[4i+0] = (dL_dmean2D.x, dL_dmean2D.y, dL_dconic.x, dL_dconic.y)
//   gacc[4i+1] = (dL_dconic.w, dL_dopacity, dL_ddepth, S = sum of dL_dpixD over blended pairs)
//   gacc[4i+2] = (dL_dcolor.r, .g, .b, 0)
//   gacc[4i+3] = (dL_dnormal.x, .y, .z, 0)      (already x10, RS backward.cu:783)
// The per-pixel-depth term of dL_dmean2D (RS backward.cu:894-897) is -S * (zx, zy): it does not
// depend on the pixel, so it is applied once per surfel in the per-surfel kernel.
// NW warps per 16x16 tile: a warp owns a 16 x (16/NW) strip and a lane 8/NW pixels of it (2x2 quad for
// NW = 2, 2x1 pair for NW = 4).  The surfel backward keeps ~36 floats of state per pixel, so the quad
// layout needs 205 registers (4 CTAs = 8 warps per SM); pairs halve the per-thread state and double the
// resident warps at the price of one more partial-sum reduction per (warp, surfel).
template <int NW>
__global__ void __launch_bounds__(NW * 32, NW == 2 ? 8 : 5) rs_render_bwd_kernel(
    const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H, int gx,
    const float4* __restrict__ rec, const float4* __restrict__ ext, const float* __restrict__ zyv,
    const float* __restrict__ bg, const float* __restrict__ config, const float* __restrict__ final_T,
    const float* __restrict__ final_D, const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix,
    const float* __restrict__ dL_dpix_normal, const float* __restrict__ dL_dpix_depth,
    const float* __restrict__ dL_dpix_opac, float4* __restrict__ gacc) {
  constexpr int PX = 8 / NW;          // pixels per lane
  constexpr int RH = kTile / NW;      // rows per warp
  static_assert(NW == 2 || NW == 4, "2 or 4 warps per tile");
  __shared__ RSBwdSlab<NW> sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  const int tx = tile % gx, ty = tile / gx;
  const int rx0 = tx * kTile, ry0 = ty * kTile + warp * RH;
  const int qx = rx0 + (lane & 7) * 2, qy = ry0 + (lane >> 3) * (PX / 2);
  const float fxs[2] = {(float)qx, (float)(qx + 1)};
  const float fys[2] = {(float)qy, (float)(qy + 1)};
  const float wx0 = (float)rx0, wx1 = (float)min(rx0 + 15, W - 1);
  const float wy0 = (float)ry0, wy1 = (float)min(ry0 + RH - 1, H - 1);
  const size_t HW = (size_t)H * W;
  const bool surface = config[0] > 0, normalize_depth = config[1] > 0, ppd = surface && (config[2] > 0);

  float T[PX], Tfin[PX], Dfin[PX], accA[PX], lastB[PX], lalpha[PX];
  float dpx[PX][3], dpn[PX][3], dpd[PX], dpo[PX], bgdot[PX], dLc[PX], cN[PX], dpoT[PX], bgx[PX];
  uint32_t lastc[PX];
  const float b0 = bg[0], b1 = bg[1], b2 = bg[2];
  uint32_t my_max = 0;
#pragma unroll
  for (int k = 0; k < PX; ++k) {
    const int x = qx + (k & 1), y = qy + (k >> 1);
    const bool in = x < W && y < H;
    const size_t pix = (size_t)W * y + x;
    Tfin[k] = in ? final_T[pix] : 0.f;
    Dfin[k] = (in && normalize_depth) ? final_D[pix] : 0.f;
    T[k] = Tfin[k];
    lastc[k] = in ? n_contrib[pix] : 0u;
    my_max = max(my_max, lastc[k]);
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      dpx[k][ch] = in ? dL_dpix[ch * HW + pix] : 0.f;
      dpn[k][ch] = in ? dL_dpix_normal[ch * HW + pix] : 0.f;
    }
    dpd[k] = in ? dL_dpix_depth[pix] : 0.f;
    dpo[k] = in ? dL_dpix_opac[pix] : 0.f;
    accA[k] = 0.f; lastB[k] = 0.f; lalpha[k] = 0.f;
    bgdot[k] = b0 * dpx[k][0] + b1 * dpx[k][1] + b2 * dpx[k][2];
    // Per-pixel constants of RS backward.cu:822-870 hoisted out of the list walk (same operand order as the
    // reference's in-loop expressions, so the hoisted parts are bit-identical): the walk keeps no IEEE division.
    const float omT = 1.f - Tfin[k];   // >= 1e-6 for in-image pixels (RS forward.cu:608 clamps T_final)
    const bool nd = normalize_depth && in;
    dLc[k] = nd ? dpd[k] / omT : dpd[k];
    cN[k] = nd ? dpd[k] * Dfin[k] / omT / omT * -Tfin[k] : 0.f;
    dpoT[k] = dpo[k] * Tfin[k];
    bgx[k] = normalize_depth ? bgdot[k] : bgdot[k] + 10 * dpd[k];
  }
  const uint32_t wmax = __reduce_max_sync(0xffffffffu, my_max);
  if (lane == 0) sm.wmax[warp] = wmax;
  if (tid < kRBatch) {
#pragma unroll
    for (int w = 0; w < NW; ++w) sm.flag[w][tid] = 0;
  }
  __syncthreads();
  const uint2 range = ranges[tile];
  uint32_t tmax = sm.wmax[0];
#pragma unroll
  for (int w = 1; w < NW; ++w) tmax = max(tmax, sm.wmax[w]);
  const int n_eff = (int)min(tmax, range.y - range.x);
  const int rounds = (n_eff + kRBatch - 1) / kRBatch;
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

  if (tid < kRBatch && tid < n_eff) rs_stage(sm.in, 0, tid, point_list[range.x + (n_eff - 1 - tid)], rec, ext, zyv);
  cp_async_commit();
  uint32_t nxt = 0;
  if (tid < kRBatch && kRBatch + tid < n_eff) nxt = point_list[range.x + (n_eff - 1 - kRBatch - tid)];

  int cur = 0;
  for (int r = 0; r < rounds; ++r) {
    cp_async_wait_all();
    __syncthreads();
    if (tid < kRBatch && (r + 1) * kRBatch + tid < n_eff) rs_stage(sm.in, cur ^ 1, tid, nxt, rec, ext, zyv);
    cp_async_commit();
    if (tid < kRBatch && (r + 2) * kRBatch + tid < n_eff) nxt = point_list[range.x + (n_eff - 1 - (r + 2) * kRBatch - tid)];

    const int cnt = min(kRBatch, n_eff - r * kRBatch);
#pragma unroll 1
    for (int g = 0; g < kRBatch / 32; ++g) {
      const int j0 = g * 32 + lane;
      bool hit = false;
      if (j0 < cnt) {
        const float4 e = sm.in.a[cur][j0];
        const uint32_t pos0 = (uint32_t)(n_eff - 1 - (r * kRBatch + j0));
        hit = (pos0 < wmax) && (e.x + e.z >= wx0) && (e.x - e.z <= wx1) && (e.y + e.w >= wy0) &&
              (e.y - e.w <= wy1);
      }
      uint32_t m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int j = g * 32 + (__ffs(m) - 1);
        m &= m - 1;
        const uint32_t pos = (uint32_t)(n_eff - 1 - (r * kRBatch + j));
        const float4 e = sm.in.a[cur][j];
        const float4 co = sm.in.b[cur][j];
        float G[PX], al[PX], ddx[PX], ddy[PX];
        bool act[PX];
        bool any = false;
#pragma unroll
        for (int k = 0; k < PX; ++k) {
          ddx[k] = e.x - fxs[k & 1];
          ddy[k] = e.y - fys[k >> 1];
          const float dist = (co.x * ddx[k] * ddx[k] + co.z * ddy[k] * ddy[k]) + 2 * co.y * ddx[k] * ddy[k];
          const float power = -0.5f * dist;
          // ex2.approx(power * log2e): ~5 ulp on power in [-5.5, 0] (2 instructions instead of expf's 8).  The forward
          // keeps the accurate expf because n_contrib / the alpha >= 1/255 decisions are compared bit for bit; here a
          // decision can only differ for alpha within a few ulp of 1/255, a <= 0.4 % change of one pixel's terms.
          G[k] = __expf(power);
          al[k] = fminf(0.99f, co.w * G[k]);
          act[k] = (pos < lastc[k]) && !(power > 0.0f) && !(al[k] < 1.0f / 255.0f);
          any |= act[k];
        }
        if (!__any_sync(0xffffffffu, any)) continue;
        float v[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) v[q] = 0.f;
        if (any) {
          const float4 cd = sm.in.c[cur][j];
          const float4 ne = sm.in.e[cur][j];
          const float zyj = sm.in.zy[cur][j];
          const float col[3] = {cd.y, cd.z, cd.w};
          const float nrm[3] = {ne.x, ne.y, ne.z};
#pragma unroll
          for (int k = 0; k < PX; ++k) {
            if (act[k]) {
              const float alpha = al[k];
              const float inv1ma = rcp_nr(1.f - alpha);   // serves every division by (1 - alpha) (see the R3 kernel)
              T[k] = T[k] * inv1ma;
              const float dchannel_dcolor = alpha * T[k];
              // Scalar form of the accum_rec recurrences of RS backward.cu:797-850 (see the R3 kernel): colour,
              // normal and depth are 7 channels of one dot product with the pixel cotangents.
              const float d_cur = cd.x - (ddx[k] * ne.w + ddy[k] * zyj);
              float B = fmaf(d_cur, dLc[k], fmaf(col[2], dpx[k][2], fmaf(col[1], dpx[k][1], col[0] * dpx[k][0])));
#pragma unroll
              for (int ch = 0; ch < 3; ++ch) v[8 + ch] += dchannel_dcolor * dpx[k][ch];
              if (surface) {
                B = fmaf(nrm[2], dpn[k][2], fmaf(nrm[1], dpn[k][1], fmaf(nrm[0], dpn[k][0], B)));
                const float dcn = dchannel_dcolor * 10;
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) v[11 + ch] += dcn * dpn[k][ch];
              }
              v[6] += dchannel_dcolor * dLc[k];
              accA[k] = fmaf(lalpha[k], lastB[k] - accA[k], accA[k]);
              lastB[k] = B;
              float dL_dalpha = B - accA[k];
              // RS backward.cu:842-850: dpd*D/(1-Tf)^2 * -Tf / (1-alpha) / T, the first four factors hoisted (cN)
              if (normalize_depth) dL_dalpha = fmaf(cN[k] * inv1ma, rcp_nr(T[k]), dL_dalpha);
              dL_dalpha *= T[k];
              dL_dalpha += dpoT[k] * inv1ma;
              lalpha[k] = alpha;
              // background colour term, plus the +10*T depth padding when depth is not normalised (bgx)
              dL_dalpha += (-Tfin[k] * inv1ma) * bgx[k];
              // dL_ddist = -0.5 * g; v[0..4] are accumulated without their constant factors (-ddelx_dx, -ddely_dy,
              // -0.5), which the flush applies once per staged surfel (RS backward.cu:872-880).
              const float g = (dL_dalpha * co.w) * G[k];
              const float hx = g * ddx[k], hy = g * ddy[k];
              v[0] = fmaf(hy, co.y, fmaf(hx, co.x, v[0]));
              v[1] = fmaf(hx, co.y, fmaf(hy, co.z, v[1]));
              v[2] = fmaf(hx, ddx[k], v[2]);
              v[3] = fmaf(hx, ddy[k], v[3]);
              v[4] = fmaf(hy, ddy[k], v[4]);
              v[5] += G[k] * dL_dalpha;
              if (ppd) v[7] += dpd[k];
            }
          }
        }
        warp_reduce16_store(v, lane, &sm.acc[warp][j][0]);
        if (lane == 0) sm.flag[warp][j] = 1;
      }
    }
    __syncthreads();
    if (tid < kRBatch) {
      uint32_t f[NW], fany = 0;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        f[w] = sm.flag[w][tid];
        fany |= f[w];
      }
      if (fany) {
        float s[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) s[q] = 0.f;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
          if (f[w]) {
            const float4* ap = reinterpret_cast<const float4*>(&sm.acc[w][tid][0]);
#pragma unroll
            for (int q4 = 0; q4 < 4; ++q4) {
              const float4 t = ap[q4];
              s[4 * q4 + 0] += t.x; s[4 * q4 + 1] += t.y; s[4 * q4 + 2] += t.z; s[4 * q4 + 3] += t.w;
            }
          }
        }
        float4* dst = gacc + 4 * (size_t)sm.in.id[cur][tid];
        atomicAdd(dst + 0, make_float4(-ddelx_dx * s[0], -ddely_dy * s[1], -0.5f * s[2], -0.5f * s[3]));
        atomicAdd(dst + 1, make_float4(-0.5f * s[4], s[5], s[6], s[7]));
        atomicAdd(dst + 2, make_float4(s[8], s[9], s[10], 0.f));
        if (surface) atomicAdd(dst + 3, make_float4(s[11], s[12], s[13], 0.f));
#pragma unroll
        for (int w = 0; w < NW; ++w) sm.flag[w][tid] = 0;
      }
    }
    cur ^= 1;
  }
  cp_async_wait_all();
}

}  // namespace pg
