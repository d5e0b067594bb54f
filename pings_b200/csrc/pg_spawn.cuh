// Fused epilogue of spawn_gaussians: everything the reference does between the five
// decoder heads and the rasterizer - ~60 eager torch kernels in gaussian_splatting/gaussian_renderer/__init__.py
// :551-764 with the quaternion helpers of utils/tools.py:743-751,811-844 - as three kernels:
//   spawn_prepare_kernel   gathers the feature rows of the spawning points into the MLP inputs and appends the
//                          view direction rotated into the point frame (GR:563-569, 587-597, 677-698)
//   spawn_epilogue_kernel  activations + quaternion algebra per Gaussian and the alpha > 0 (and optional scale)
//                          stream compaction with a device-side count (single pass, decoupled look-back)
//   spawn_epilogue_bwd_kernel   the adjoint, dense over the N*K raw decoder outputs
// Layout of the raw head outputs: [n, out*K] row-major, i.e. Gaussian r = i*K + k reads out consecutive floats
// starting at out*r (the reference's `.view(NK, -1)`).
#pragma once
#include "pg_common.cuh"

namespace pg {

using SpawnArgs = PgSpawnArgs;
using SpawnOut = PgSpawnOut;
using SpawnGrads = PgSpawnGrads;

struct F3s { float x, y, z; };
__device__ __forceinline__ F3s cross3s(F3s a, F3s b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// utils/tools.py:743-751 apply_quaternion_rotation: p' = p + w t + qv x t, qv = -q.xyz, t = 2 qv x p.
__device__ __forceinline__ F3s quat_rotate(float w, float x, float y, float z, F3s p) {
  const F3s qv = {-x, -y, -z};
  F3s t = cross3s(qv, p);
  t = {2.f * t.x, 2.f * t.y, 2.f * t.z};
  const F3s u = cross3s(qv, t);
  return {p.x + w * t.x + u.x, p.y + w * t.y + u.y, p.z + w * t.z + u.z};
}

// ---- prepare: gather + view direction -----------------------------------------------------------------------
// geo_in [n, Fg (+1 if dist_concat)], color_in [n, Fc (+3 if view_concat)]; one warp per point.
__global__ void __launch_bounds__(256) spawn_prepare_kernel(
    int64_t n, const int64_t* __restrict__ sidx, const float* __restrict__ position, const float* __restrict__ orientation,
    const float* __restrict__ cam_origin, const float* __restrict__ geo_feature, int Fg, const float* __restrict__ color_feature,
    int Fc, int dist_concat, int view_concat, int xy_only, float* __restrict__ geo_in, float* __restrict__ color_in) {
  const int lane = threadIdx.x & 31;
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  const int64_t src = sidx ? sidx[i] : i;
  const int gw = Fg + (dist_concat ? 1 : 0), cw = Fc + (view_concat ? 3 : 0);
  for (int c = lane; c < Fg; c += 32) geo_in[i * gw + c] = geo_feature[src * Fg + c];
  if (color_in)
    for (int c = lane; c < Fc; c += 32) color_in[i * cw + c] = color_feature[src * Fc + c];
  if (lane == 0 && cam_origin && (dist_concat || (view_concat && color_in))) {
    F3s v = {position[3 * src] - cam_origin[0], position[3 * src + 1] - cam_origin[1], position[3 * src + 2] - cam_origin[2]};
    if (xy_only) v.z = 0.f;
    const float dist = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    if (dist_concat) geo_in[i * gw + Fg] = dist;
    if (view_concat && color_in) {
      v = {v.x / dist, v.y / dist, v.z / dist};
      const float4 q = reinterpret_cast<const float4*>(orientation)[src];
      const F3s r = quat_rotate(q.x, -q.y, -q.z, -q.w, v);   // quat_inverse, utils/tools.py:811-815
      color_in[i * cw + Fc] = r.x;
      color_in[i * cw + Fc + 1] = r.y;
      color_in[i * cw + Fc + 2] = r.z;
    }
  }
}

// d_feature[sidx[i], :F] += d_in[i, :F]  (rows of sidx are distinct: plain read-modify-write)
__global__ void __launch_bounds__(256) spawn_prepare_bwd_kernel(int64_t n, const int64_t* __restrict__ sidx,
                                                                const float* __restrict__ d_in, int in_w, int F,
                                                                float* __restrict__ d_feature) {
  const int lane = threadIdx.x & 31;
  const int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (i >= n) return;
  const int64_t dst = sidx ? sidx[i] : i;
  for (int c = lane; c < F; c += 32) d_feature[dst * F + c] += d_in[i * in_w + c];
}

// ---- epilogue forward -----------------------------------------------------------------------------------------
constexpr uint32_t kSpawnAgg = 1u << 30, kSpawnPrefix = 2u << 30, kSpawnVal = (1u << 30) - 1u;

__global__ void __launch_bounds__(256) spawn_epilogue_kernel(SpawnArgs a, SpawnOut o, uint32_t nblocks) {
  __shared__ uint32_t s_bid, s_base, wcount[8];
  if (threadIdx.x == 0) s_bid = atomicAdd(o.state + nblocks, 1u);   // blocks take tiles in start order: look-back cannot deadlock
  __syncthreads();
  const uint32_t bid = s_bid;
  const int64_t NK = a.n * a.K;
  const int64_t r = (int64_t)bid * 256 + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  bool keep = false;
  float ox[3], os[3], oq[4], oa = 0.f, oc[3];
  uint8_t fm = 0;
  if (r < NK) {
    const int64_t i = r / a.K;
    const int64_t src = a.sidx ? a.sidx[i] : i;
    const float4 q = reinterpret_cast<const float4*>(a.orientation)[src];
    // xyz = p + q (x) (range * tanh(.))                                       GR:605-639
    const F3s d = {a.disp_range * tanhf(a.xyz_raw[3 * r]), a.disp_range * tanhf(a.xyz_raw[3 * r + 1]),
                   a.disp_range * tanhf(a.xyz_raw[3 * r + 2])};
    const F3s dr = quat_rotate(q.x, q.y, q.z, q.w, d);
    ox[0] = a.position[3 * src] + dr.x;
    ox[1] = a.position[3 * src + 1] + dr.y;
    ox[2] = a.position[3 * src + 2] + dr.z;
    // rot = q (*) nan_to_num(normalize(.))                                    GR:644-649
    float v[4] = {a.rot_raw[4 * r], a.rot_raw[4 * r + 1], a.rot_raw[4 * r + 2], a.rot_raw[4 * r + 3]};
    const float nrm = fmaxf(sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]), 1e-12f);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      v[c] = v[c] / nrm;
      if (!isfinite(v[c])) v[c] = 0.f;
    }
    oq[0] = q.x * v[0] - q.y * v[1] - q.z * v[2] - q.w * v[3];
    oq[1] = q.x * v[1] + q.y * v[0] + q.z * v[3] - q.w * v[2];
    oq[2] = q.x * v[2] - q.y * v[3] + q.z * v[0] + q.w * v[1];
    oq[3] = q.x * v[3] + q.y * v[2] - q.z * v[1] + q.w * v[0];
    if (!(oq[0] == oq[0]) || !(oq[1] == oq[1]) || !(oq[2] == oq[2]) || !(oq[3] == oq[3])) atomicOr(o.state + nblocks + 2, 1u);
    // scale = min(unit * exp(. [+ dist / z_far]), max); surfel: third = 1e-7  GR:654-670
    float dist_ratio = 0.f;
    if (a.z_far > 0.f && a.cam_origin) {
      F3s vd = {a.position[3 * src] - a.cam_origin[0], a.position[3 * src + 1] - a.cam_origin[1],
                a.position[3 * src + 2] - a.cam_origin[2]};
      if (a.xy_only) vd.z = 0.f;
      dist_ratio = sqrtf(vd.x * vd.x + vd.y * vd.y + vd.z * vd.z) / a.z_far;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) os[c] = fminf(a.unit_scale * expf(a.scale_raw[3 * r + c] + dist_ratio), a.max_scale);
    if (a.surfel) os[2] = 1e-7f;
    // alpha = tanh(.)                                                         GR:685
    oa = tanhf(a.alpha_raw[r]);
    o.alpha_all[r] = oa;
    // colour                                                                  GR:706-714
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float u = a.color_raw[3 * r + c];
      if (a.residual && a.color) oc[c] = fminf(fmaxf(a.color[3 * src + c] + 0.1f * tanhf(u), 0.f), 1.f);
      else oc[c] = 1.f / (1.f + expf(-u));
    }
    keep = !a.alpha_filter || oa > 0.f;                                        // GR:727-737
    if (a.scale_filter >= 0.f) keep = keep && (os[0] > a.scale_filter || os[1] > a.scale_filter || os[2] > a.scale_filter);
    // GR:719-720: `free_mask.repeat(1, K).view(-1)` on the 1-D per-point mask TILES it (entry r <- point r mod n),
    // it does not repeat each point K times; reproduced as is (parity quirk)
    if (o.free_mask) {
      const int64_t j = r % a.n;
      fm = o.free_mask[a.sidx ? a.sidx[j] : j];
    }
  }
  // ordered compaction: warp ballot -> block scan -> decoupled look-back over the blocks
  const uint32_t bal = __ballot_sync(0xffffffffu, keep);
  if (lane == 0) wcount[warp] = __popc(bal);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int w = 0; w < 8; ++w) {
      const uint32_t c = wcount[w];
      wcount[w] = run;
      run += c;
    }
    volatile uint32_t* st = o.state;
    uint32_t excl = 0;
    if (bid == 0) {
      st[0] = kSpawnPrefix | run;
    } else {
      st[bid] = kSpawnAgg | run;
      __threadfence();
      for (int64_t p = (int64_t)bid - 1; p >= 0; --p) {
        uint32_t sv;
        while (((sv = st[p]) >> 30) == 0u) {}
        excl += sv & kSpawnVal;
        if ((sv >> 30) == 2u) break;
      }
      st[bid] = kSpawnPrefix | (excl + run);
    }
    __threadfence();
    s_base = excl;
    if (bid == nblocks - 1) o.state[nblocks + 1] = excl + run;   // total kept
  }
  __syncthreads();
  if (r < NK) {
    int32_t m = -1;
    if (keep) {
      m = (int32_t)(s_base + wcount[warp] + __popc(bal & ((1u << lane) - 1u)));
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        o.xyz[3 * (int64_t)m + c] = ox[c];
        o.scale[3 * (int64_t)m + c] = os[c];
        o.color[3 * (int64_t)m + c] = oc[c];
      }
      reinterpret_cast<float4*>(o.rot)[m] = make_float4(oq[0], oq[1], oq[2], oq[3]);
      o.alpha[m] = oa;
      if (o.free_out) o.free_out[m] = fm;
    }
    o.cpos[r] = m;
  }
}

// ---- epilogue backward ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) spawn_epilogue_bwd_kernel(SpawnArgs a, SpawnGrads g) {
  const int64_t NK = a.n * a.K;
  const int64_t r = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (r >= NK) return;
  const int32_t m = g.cpos[r];
  float dx[3] = {0, 0, 0}, dq[4] = {0, 0, 0, 0}, ds[3] = {0, 0, 0}, da = 0.f, dc[3] = {0, 0, 0};
  const float ta = tanhf(a.alpha_raw[r]);
  if (g.d_alpha_all) da = g.d_alpha_all[r];
  if (m >= 0) {
    const int64_t i = r / a.K;
    const int64_t src = a.sidx ? a.sidx[i] : i;
    const float4 q = reinterpret_cast<const float4*>(a.orientation)[src];
    if (g.d_xyz) {   // d/du of q (x) (range tanh(u)): rotate the cotangent with the conjugate, then tanh'
      const F3s c = {g.d_xyz[3 * (int64_t)m], g.d_xyz[3 * (int64_t)m + 1], g.d_xyz[3 * (int64_t)m + 2]};
      // transpose of p -> p + w t + qv x t (t = 2 qv x p): the same map with w -> -w
      const F3s qv = {-q.y, -q.z, -q.w};
      F3s t = cross3s(qv, c);
      t = {2.f * t.x, 2.f * t.y, 2.f * t.z};
      const F3s u = cross3s(qv, t);
      const float bk[3] = {c.x - q.x * t.x + u.x, c.y - q.x * t.y + u.y, c.z - q.x * t.z + u.z};
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float th = tanhf(a.xyz_raw[3 * r + k]);
        dx[k] = a.disp_range * (1.f - th * th) * bk[k];
      }
    }
    if (g.d_rot) {
      const float4 c = reinterpret_cast<const float4*>(g.d_rot)[m];
      // adjoint of y -> q (*) y
      float dy[4];
      dy[0] = q.x * c.x + q.y * c.y + q.z * c.z + q.w * c.w;
      dy[1] = -q.y * c.x + q.x * c.y + q.w * c.z - q.z * c.w;
      dy[2] = -q.z * c.x - q.w * c.y + q.x * c.z + q.y * c.w;
      dy[3] = -q.w * c.x + q.z * c.y - q.y * c.z + q.x * c.w;
      float v[4] = {a.rot_raw[4 * r], a.rot_raw[4 * r + 1], a.rot_raw[4 * r + 2], a.rot_raw[4 * r + 3]};
      const float n2 = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3]);
      if (n2 > 1e-12f) {   // y = v / |v|: d_v = (d_y - y (y . d_y)) / |v|
        const float inv = 1.f / n2;
        float y[4], dot = 0.f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          y[k] = v[k] * inv;
          dot += y[k] * dy[k];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float gk = (dy[k] - y[k] * dot) * inv;
          dq[k] = isfinite(y[k]) ? gk : 0.f;
        }
      } else {             // clamped norm: y = v / eps
#pragma unroll
        for (int k = 0; k < 4; ++k) dq[k] = dy[k] / 1e-12f;
      }
    }
    if (g.d_scale) {
      float dist_ratio = 0.f;
      if (a.z_far > 0.f && a.cam_origin) {
        F3s vd = {a.position[3 * src] - a.cam_origin[0], a.position[3 * src + 1] - a.cam_origin[1],
                  a.position[3 * src + 2] - a.cam_origin[2]};
        if (a.xy_only) vd.z = 0.f;
        dist_ratio = sqrtf(vd.x * vd.x + vd.y * vd.y + vd.z * vd.z) / a.z_far;
      }
      const int dims = a.surfel ? 2 : 3;
      for (int k = 0; k < dims; ++k) {
        const float s = a.unit_scale * expf(a.scale_raw[3 * r + k] + dist_ratio);
        ds[k] = s <= a.max_scale ? g.d_scale[3 * (int64_t)m + k] * s : 0.f;   // clamp(max=): gradient passes where s <= max
      }
    }
    if (g.d_alpha) da += g.d_alpha[m];
    if (g.d_color) {
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const float u = a.color_raw[3 * r + k];
        const float c = g.d_color[3 * (int64_t)m + k];
        if (a.residual && a.color) {
          const float th = tanhf(u);
          const float val = a.color[3 * src + k] + 0.1f * th;
          dc[k] = (val >= 0.f && val <= 1.f) ? c * 0.1f * (1.f - th * th) : 0.f;
        } else {
          const float sg = 1.f / (1.f + expf(-u));
          dc[k] = c * sg * (1.f - sg);
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    g.d_xyz_raw[3 * r + k] = dx[k];
    g.d_scale_raw[3 * r + k] = ds[k];
    g.d_color_raw[3 * r + k] = dc[k];
  }
  reinterpret_cast<float4*>(g.d_rot_raw)[r] = make_float4(dq[0], dq[1], dq[2], dq[3]);
  g.d_alpha_raw[r] = da * (1.f - ta * ta);
}

}  // namespace pg
