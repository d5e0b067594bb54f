// Post-raster loss assembly: the photometric L1 and the masked depth L1 of the
// GS mapping loss (utils/mapper.py:1239,1252-1267; l1_loss = mean |a - b|) in one pass over the
// rendered images, with their backward seeds.
//
// Forward: one streaming kernel reads rendered rgb / gt rgb / rendered depth / gt depth / alpha once,
// writes sign(.) seeds for rgb and the (masked, optionally inverse-depth) depth derivative, and
// reduces three scalars {sum |rgb diff|, sum |depth diff| over valid, valid count} with a block
// reduction + one atomic per block.  Backward: one kernel scales the seeds by upstream / count,
// reading the count on the device (no host sync).
#pragma once
#include "pg_common.cuh"

namespace pg {

struct PhotoDepthArgs {
  int HW;
  const float* rgb;        // [3*HW]
  const float* gt_rgb;     // [3*HW]
  const float* depth;      // [HW]  (nullable -> no depth term)
  const float* gt_depth;   // [HW]
  const float* alpha;      // [HW]  (nullable -> no accumulated-alpha mask)
  float depth_min, depth_max, min_alpha;
  int inverse_depth;
};

// sums[0] = sum |rgb - gt|, sums[1] = sum depth error over valid pixels, sums[2] = number of valid pixels
__global__ void __launch_bounds__(256) photo_depth_fwd_kernel(PhotoDepthArgs a, float* __restrict__ seed_rgb,
                                                              float* __restrict__ seed_depth, double* __restrict__ sums) {
  float s_rgb = 0.f, s_d = 0.f, s_n = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.HW; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float d = a.rgb[c * a.HW + i] - a.gt_rgb[c * a.HW + i];
      s_rgb += fabsf(d);
      seed_rgb[c * a.HW + i] = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);   // torch: sign(0) = 0
    }
    if (a.depth) {
      const float g = a.gt_depth[i], r = a.depth[i];
      bool valid = g > a.depth_min && g < a.depth_max;
      if (a.alpha) valid = valid && a.alpha[i] > a.min_alpha;
      float seed = 0.f;
      if (valid) {
        if (a.inverse_depth) {   // l1(1/gt, 1/rendered): d/dr |1/g - 1/r| = sign(1/g - 1/r) / r^2
          const float e = 1.f / g - 1.f / r;
          s_d += fabsf(e);
          seed = (e > 0.f ? 1.f : (e < 0.f ? -1.f : 0.f)) / (r * r);
        } else {                 // l1(gt, rendered): d/dr |g - r| = -sign(g - r)
          const float e = g - r;
          s_d += fabsf(e);
          seed = e > 0.f ? -1.f : (e < 0.f ? 1.f : 0.f);
        }
        s_n += 1.f;
      }
      seed_depth[i] = seed;
    }
  }
  __shared__ float red[3][8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s_rgb += __shfl_xor_sync(0xffffffffu, s_rgb, o);
    s_d += __shfl_xor_sync(0xffffffffu, s_d, o);
    s_n += __shfl_xor_sync(0xffffffffu, s_n, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = s_rgb; red[1][warp] = s_d; red[2][warp] = s_n; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += (double)red[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], t);
  }
}

// loss = sums[0] / (3 HW) + lambda_depth * sums[1] / max(sums[2], 1)   (an empty mask gives a zero depth term)
__global__ void photo_depth_finish_kernel(int HW, float lambda_depth, const double* __restrict__ sums,
                                          float* __restrict__ loss /* [3]: total, rgb l1, depth l1 */) {
  const double l_rgb = sums[0] / (3.0 * (double)HW);
  const double l_d = sums[2] > 0.0 ? sums[1] / sums[2] : 0.0;
  loss[0] = (float)(l_rgb + (double)lambda_depth * l_d);
  loss[1] = (float)l_rgb;
  loss[2] = (float)l_d;
}

__global__ void __launch_bounds__(256) photo_depth_bwd_kernel(int HW, float lambda_depth, const float* __restrict__ upstream,
                                                              const double* __restrict__ sums,
                                                              const float* __restrict__ seed_rgb,
                                                              const float* __restrict__ seed_depth,
                                                              float* __restrict__ d_rgb, float* __restrict__ d_depth) {
  const float up = upstream[0];
  const float k_rgb = up / (3.f * (float)HW);
  const float k_d = sums[2] > 0.0 ? up * lambda_depth / (float)sums[2] : 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += gridDim.x * blockDim.x) {
    d_rgb[i] = seed_rgb[i] * k_rgb;
    d_rgb[HW + i] = seed_rgb[HW + i] * k_rgb;
    d_rgb[2 * HW + i] = seed_rgb[2 * HW + i] * k_rgb;
    if (d_depth) d_depth[i] = seed_depth[i] * k_d;
  }
}


// ---- depth2normal (gaussian_splatting/utils/point_utils.py:83-148) ------------------------------------------------
// Normal map in the camera frame from a depth map: back-project every pixel, difference to the four neighbours
// (replicate padding, masked), sum of the four cross products, normalise, mask [, times the detached alpha as
// GR:381-382 does right after].  The reference runs ~25 eager full-image kernels for this.
struct D2NArgs {
  int H, W;
  float cx, cy, inv_fx, inv_fy;   // principal point and 1/focal in pixels of the (down-scaled) image
  const float* depth;             // [H*W]
  const uint8_t* mask;            // [H*W] visibility mask (alpha > min_alpha)
  const float* alpha;             // [H*W] or NULL: the output is multiplied by it (no gradient: detached)
};

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 v3add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 v3mul(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ V3 v3cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ float v3dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct D2NStencil {
  int iu, il, ib, ir;     // linear indices of the (clamped) neighbours
  V3 ac, au, al, ab, ar;  // back-projection rays (x - cx)/fx, (y - cy)/fy, 1 at the five pixels
  float mc, mu, ml, mb, mr;
  V3 pu, pl, pb, pr;      // masked differences
};
__device__ __forceinline__ D2NStencil d2n_stencil(const D2NArgs& a, int x, int y) {
  D2NStencil s;
  const int yu = max(y - 1, 0), yb = min(y + 1, a.H - 1), xl = max(x - 1, 0), xr = min(x + 1, a.W - 1);
  const int ic = y * a.W + x;
  s.iu = yu * a.W + x; s.il = y * a.W + xl; s.ib = yb * a.W + x; s.ir = y * a.W + xr;
  auto ray = [&](int xx, int yy) { return V3{((float)xx - a.cx) * a.inv_fx, ((float)yy - a.cy) * a.inv_fy, 1.f}; };
  s.ac = ray(x, y); s.au = ray(x, yu); s.al = ray(xl, y); s.ab = ray(x, yb); s.ar = ray(xr, y);
  s.mc = a.mask[ic] ? 1.f : 0.f; s.mu = a.mask[s.iu] ? 1.f : 0.f; s.ml = a.mask[s.il] ? 1.f : 0.f;
  s.mb = a.mask[s.ib] ? 1.f : 0.f; s.mr = a.mask[s.ir] ? 1.f : 0.f;
  // Back-projected points with explicitly rounded products, in the reference's order ((x - cx) * d) / fx: at an image
  // border the replicated neighbour IS the centre pixel and their difference must be exactly zero (an FMA-contracted
  // a*d - round(a*d) leaves a 1e-7 residue that the normalisation blows up to a unit vector).
  auto point = [&](int xx, int yy, int idx) {
    const float d = a.depth[idx];
    return V3{__fmul_rn(__fmul_rn((float)xx - a.cx, d), a.inv_fx), __fmul_rn(__fmul_rn((float)yy - a.cy, d), a.inv_fy), d};
  };
  const V3 pc = v3mul(point(x, y, ic), s.mc);
  s.pu = v3mul(v3sub(point(x, yu, s.iu), pc), s.mu);
  s.pl = v3mul(v3sub(point(xl, y, s.il), pc), s.ml);
  s.pb = v3mul(v3sub(point(x, yb, s.ib), pc), s.mb);
  s.pr = v3mul(v3sub(point(xr, y, s.ir), pc), s.mr);
  return s;
}

__global__ void __launch_bounds__(256) depth2normal_fwd_kernel(D2NArgs a, float* __restrict__ out /* [3,H*W] */) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = a.H * a.W;
  if (i >= HW) return;
  const D2NStencil s = d2n_stencil(a, i % a.W, i / a.W);
  V3 n = v3add(v3add(v3cross(s.pu, s.pl), v3cross(s.pr, s.pu)), v3add(v3cross(s.pb, s.pr), v3cross(s.pl, s.pb)));
  const float len = fmaxf(sqrtf(v3dot(n, n)), 1e-12f);
  const float k = s.mc * (a.alpha ? a.alpha[i] : 1.f) / len;
  out[i] = n.x * k;
  out[HW + i] = n.y * k;
  out[2 * HW + i] = n.z * k;
}

// d_depth (zero-initialised by the caller) += adjoint; every pixel scatters into itself and its four neighbours.
__global__ void __launch_bounds__(256) depth2normal_bwd_kernel(D2NArgs a, const float* __restrict__ g /* [3,H*W] */,
                                                               float* __restrict__ d_depth) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = a.H * a.W;
  if (i >= HW) return;
  const D2NStencil s = d2n_stencil(a, i % a.W, i / a.W);
  if (s.mc == 0.f) return;
  const V3 nr = v3add(v3add(v3cross(s.pu, s.pl), v3cross(s.pr, s.pu)), v3add(v3cross(s.pb, s.pr), v3cross(s.pl, s.pb)));
  const float len = sqrtf(v3dot(nr, nr));
  const float sc = s.mc * (a.alpha ? a.alpha[i] : 1.f);
  V3 dn = v3mul(V3{g[i], g[HW + i], g[2 * HW + i]}, sc);
  V3 G;   // cotangent of the un-normalised sum
  if (len > 1e-12f) {
    const V3 n = v3mul(nr, 1.f / len);
    G = v3mul(v3sub(dn, v3mul(n, v3dot(n, dn))), 1.f / len);
  } else {
    G = v3mul(dn, 1.f / 1e-12f);
  }
  // n = pu x pl + pr x pu + pb x pr + pl x pb;  d(a x b).G / da = b x G,  / db = G x a
  const V3 dPu = v3add(v3cross(s.pl, G), v3cross(G, s.pr));
  const V3 dPl = v3add(v3cross(G, s.pu), v3cross(s.pb, G));
  const V3 dPr = v3add(v3cross(s.pu, G), v3cross(G, s.pb));
  const V3 dPb = v3add(v3cross(s.pr, G), v3cross(G, s.pl));
  // p_k = (P_k - P_c m_c) m_k
  const V3 du = v3mul(dPu, s.mu), dl = v3mul(dPl, s.ml), dr = v3mul(dPr, s.mr), db = v3mul(dPb, s.mb);
  const V3 dc = v3mul(v3add(v3add(du, dl), v3add(dr, db)), -s.mc);
  // At an image border the replicated neighbour IS the centre pixel, so its cotangent lands on this pixel too and
  // cancels the same term inside dc.  The two are combined HERE, in registers, and reach memory as one atomic: when
  // the normal degenerates (len = 0: border + a masked neighbour; G = dn / 1e-12 as F.normalize's eps gives) they
  // are +-1e12, and scattered separately the order of the atomics decided how much of the neighbours' contributions
  // survived (an intermittent 0.007 at one pixel of the parity test).  Inside dc the huge term still absorbs the
  // small ones exactly as it does in the eager reference's fp32 sum (tests/golden/render_post.npz contains that).
  float self = v3dot(dc, s.ac);
  if (s.iu == i) self += v3dot(du, s.au); else atomicAdd(d_depth + s.iu, v3dot(du, s.au));
  if (s.il == i) self += v3dot(dl, s.al); else atomicAdd(d_depth + s.il, v3dot(dl, s.al));
  if (s.ib == i) self += v3dot(db, s.ab); else atomicAdd(d_depth + s.ib, v3dot(db, s.ab));
  if (s.ir == i) self += v3dot(dr, s.ar); else atomicAdd(d_depth + s.ir, v3dot(dr, s.ar));
  atomicAdd(d_depth + i, self);
}

// ---- affine exposure correction (GR:449-458): out = M rgb + offset per pixel ------------------------------------------
__global__ void __launch_bounds__(256) exposure_fwd_kernel(int HW, const float* __restrict__ img, const float* __restrict__ mat,
                                                           const float* __restrict__ off, float* __restrict__ out) {
  float m[9], o[3];
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] = mat[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) o[k] = off[k];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += gridDim.x * blockDim.x) {
    const float r = img[i], gg = img[HW + i], b = img[2 * HW + i];
    out[i] = m[0] * r + m[1] * gg + m[2] * b + o[0];
    out[HW + i] = m[3] * r + m[4] * gg + m[5] * b + o[1];
    out[2 * HW + i] = m[6] * r + m[7] * gg + m[8] * b + o[2];
  }
}

// d_img = M^T g; d_mat[c][k] += sum g_c img_k; d_off[c] += sum g_c  (d_mat / d_off zero-initialised, 12 floats)
__global__ void __launch_bounds__(256) exposure_bwd_kernel(int HW, const float* __restrict__ img, const float* __restrict__ mat,
                                                           const float* __restrict__ g, float* __restrict__ d_img,
                                                           float* __restrict__ d_mat_off /* [12] */) {
  float m[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] = mat[k];
  float acc[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) acc[k] = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += gridDim.x * blockDim.x) {
    const float x[3] = {img[i], img[HW + i], img[2 * HW + i]};
    const float gc[3] = {g[i], g[HW + i], g[2 * HW + i]};
    if (d_img) {
#pragma unroll
      for (int k = 0; k < 3; ++k) d_img[k * HW + i] = m[k] * gc[0] + m[3 + k] * gc[1] + m[6 + k] * gc[2];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
      for (int k = 0; k < 3; ++k) acc[3 * c + k] += gc[c] * x[k];
      acc[9 + c] += gc[c];
    }
  }
  __shared__ float red[12][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 12; ++k) {
    float v = acc[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[k][warp] = v;
  }
  __syncthreads();
  if (threadIdx.x < 12) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[threadIdx.x][w];
    atomicAdd(d_mat_off + threadIdx.x, t);
  }
}

}  // namespace pg
