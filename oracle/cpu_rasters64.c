/* TEST INFRASTRUCTURE — not part of the product.
 *
 * fp64 restatement ("arbiter") of the reference Gaussian-SURFEL rasterizer with camera-pose gradients,
 * submodules/diff-gaussian-surfel-rasterization-w-pose (RS), forward blend + backward blend + the
 * per-surfel backward chain.  Every formula follows the reference statement cited next to it, evaluated in
 * double precision from the fp32 inputs.  The reference accumulates its backward with per-thread float
 * atomics and is ill-conditioned on thin edge-on surfels (1/det^2 of the projected covariance), so two runs
 * of the reference itself differ by percents on single rows; this restatement has no such noise and is used
 * to decide, row by row, whether the CUDA implementation or the reference is closer to the exact value
 * (tests/test_rasters_gpu.py).
 *
 * Discrete decisions that the fp32 pipelines agree on bit for bit are INPUTS here, not recomputed: the set
 * of visible surfels (radii > 0) and the depth-sorted per-tile lists (point_list, ranges).  Decisions inside
 * the blend (alpha >= 1/255, T termination) are taken in fp64.
 *
 * Parity: UNPINNED by reference golden vectors (the reference has none for this path); pinned (a) on the CPU by
 * central finite differences of its own forward (tests/test_oracle_rasters64.py) and (b) on the GPU box
 * against the reference CUDA build (tests/test_rasters_gpu.py).
 *
 * Only the colors_precomp + scale/rotation path (the one PINGS uses, GR:110,321) is restated.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { double x, y, z; } d3;
typedef struct { double c[3][3]; } m3;   /* column-major like glm::mat3: c[col][row] */

static d3 xf_point43(d3 p, const double* m) {   /* RS auxiliary.h transformPoint4x3 */
  d3 r = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
  return r;
}
static void xf_point44(d3 p, const double* m, double* o) {
  o[0] = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  o[1] = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  o[2] = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  o[3] = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
}
static d3 xf_vec43(d3 p, const double* m) {     /* transformVec4x3 */
  d3 r = {m[0] * p.x + m[4] * p.y + m[8] * p.z, m[1] * p.x + m[5] * p.y + m[9] * p.z,
          m[2] * p.x + m[6] * p.y + m[10] * p.z};
  return r;
}
static d3 xf_vec43T(d3 p, const double* m) {    /* transformVec4x3Transpose */
  d3 r = {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
          m[8] * p.x + m[9] * p.y + m[10] * p.z};
  return r;
}
static d3 cross(d3 a, d3 b) {
  d3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
  return r;
}
static m3 m3mul(m3 a, m3 b) {
  m3 r;
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
  return r;
}
static m3 m3t(m3 a) {
  m3 r;
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[i][j];
  return r;
}
static m3 quat_R(const float* q) { /* RS forward.cu:141-156 quaternion2rotmat (un-normalised) */
  double r = q[0], x = q[1], y = q[2], z = q[3];
  m3 R = {{{1. - 2. * (y * y + z * z), 2. * (x * y - r * z), 2. * (x * z + r * y)},
           {2. * (x * y + r * z), 1. - 2. * (x * x + z * z), 2. * (y * z - r * x)},
           {2. * (x * z - r * y), 2. * (y * z + r * x), 1. - 2. * (x * x + y * y)}}};
  return R;
}
/* RS forward.cu:162-202 computeCov3D: S[2][2] = (mod * surface) ? 0 : scale.z  (no scale_modifier on z) */
static void cov3d(const float* s, double mod, m3 R, int surface, double* c6) {
  const double sz = (mod * (surface ? 1.0 : 0.0)) != 0.0 ? 0.0 : (double)s[2];
  m3 S = {{{mod * s[0], 0, 0}, {0, mod * s[1], 0}, {0, 0, sz}}};
  m3 M = m3mul(S, R);
  m3 Sg = m3mul(m3t(M), M);
  c6[0] = Sg.c[0][0]; c6[1] = Sg.c[0][1]; c6[2] = Sg.c[0][2];
  c6[3] = Sg.c[1][1]; c6[4] = Sg.c[1][2]; c6[5] = Sg.c[2][2];
}
typedef struct { d3 t; double txtz, tytz; m3 J, W, T, V; } cov2d_aux;
/* RS forward.cu:74-139 computeCov2D (view-space mean in, same clamp) */
static void cov2d(d3 t, double fx, double fy, double tanx, double tany, const double* c6, const double* view,
                  double* abc, cov2d_aux* ax) {
  const double limx = 1.3 * tanx, limy = 1.3 * tany;
  const double txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fmin(limx, fmax(-limx, txtz)) * t.z;
  t.y = fmin(limy, fmax(-limy, tytz)) * t.z;
  m3 J = {{{fx / t.z, 0, -(fx * t.x) / (t.z * t.z)}, {0, fy / t.z, -(fy * t.y) / (t.z * t.z)}, {0, 0, 0}}};
  m3 W = {{{view[0], view[4], view[8]}, {view[1], view[5], view[9]}, {view[2], view[6], view[10]}}};
  m3 T = m3mul(W, J);
  m3 V = {{{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}}};
  m3 cov = m3mul(m3mul(m3t(T), m3t(V)), T);
  abc[0] = cov.c[0][0] + 0.3;
  abc[1] = cov.c[0][1];
  abc[2] = cov.c[1][1] + 0.3;
  if (ax) { ax->t = t; ax->txtz = txtz; ax->tytz = tytz; ax->J = J; ax->W = W; ax->T = T; ax->V = V; }
}
static double norm3(double* v) {   /* RS auxiliary.h normalize() */
  double m = fmax(sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), 0.00000001);
  v[0] /= m; v[1] /= m; v[2] /= m;
  return m;
}
/* RS auxiliary.h:328-425 local_homo, reduced to the two products depth_differencing().z needs
 * (auxiliary.h:427-434): zx = J0*u0.z + J2*u1.z, zy = J1*u0.z + J3*u1.z. */
static void local_homo(d3 pv, d3 n, double fx, double fy, d3 ax0, d3 ax1, double* zx, double* zy) {
  const double px = pv.x / pv.z, py = pv.y / pv.z, S_fix = 1000, Svp = (fx + fy) / 2;
  double d0[3] = {px + 1 / S_fix, py, 1}, d1[3] = {px, py + 1 / S_fix, 1};
  norm3(d0);
  norm3(d1);
  const double prj0 = d0[0] * n.x + d0[1] * n.y + d0[2] * n.z, prj1 = d1[0] * n.x + d1[1] * n.y + d1[2] * n.z;
  const double tt = pv.x * n.x + pv.y * n.y + pv.z * n.z, t0 = tt / prj0, t1 = tt / prj1;
  const double xu0[3] = {d0[0] * t0 - pv.x, d0[1] * t0 - pv.y, d0[2] * t0 - pv.z};
  const double xu1[3] = {d1[0] * t1 - pv.x, d1[1] * t1 - pv.y, d1[2] * t1 - pv.z};
  double J[4] = {xu0[0] * ax0.x + xu0[1] * ax0.y + xu0[2] * ax0.z, xu1[0] * ax0.x + xu1[1] * ax0.y + xu1[2] * ax0.z,
                 xu0[0] * ax1.x + xu0[1] * ax1.y + xu0[2] * ax1.z, xu1[0] * ax1.x + xu1[1] * ax1.y + xu1[2] * ax1.z};
  for (int k = 0; k < 4; ++k) J[k] /= (Svp / S_fix);
  *zx = ax0.z * J[0] + ax1.z * J[2];
  *zy = ax0.z * J[1] + ax1.z * J[3];
}

typedef struct {
  double px, py, ca, cb, cc, o, depth, nx, ny, nz, zx, zy;
} srec;

/* Blend thresholds (RS forward.cu:519,524 / backward.cu:712): alpha < 1/255 skip, T(1-alpha) < 1e-4 stop.
 * Settable ONLY so that tests/test_oracle_rasters64.py can switch the two discontinuities off and check the
 * restated calculus against central finite differences of the fp64 forward; every parity use keeps the
 * reference values. */
static double g_alpha_min = 1.0 / 255.0, g_t_min = 0.0001;
void rs64_set_thresholds(double alpha_min, double t_min) { g_alpha_min = alpha_min; g_t_min = t_min; }

/* Blend forward + backward in fp64 on the given per-tile lists.
 *   RS forward.cu:375-624 (renderCUDA fwd), RS backward.cu:585-914 (renderCUDA bwd).
 * Images out: color/normal [3,H,W], depth/opac [H,W] (may be NULL).  Blend sums out (zero-filled here):
 * dL_dmean2D [P,2] (including the per-pixel-depth term, backward.cu:894-897), dL_dconic [P,3] (x, y, w),
 * dL_dopacity [P], dL_dcolor [P,3], dL_dnormal [P,3] (x10, backward.cu:783), dL_ddepth [P].
 * n_contrib_out [H*W] (optional) receives the fp64 last-contributor counts. */
void rs64_blend(int P, int W, int H, const float* bg, const float* means, const float* colors, const float* opac,
                const float* scales, float mod_f, const float* rots, const float* view_f, const float* proj_f,
                float tanx_f, float tany_f, const float* config, const int* radii, const uint32_t* point_list,
                const uint32_t* ranges, const float* dL_dpixC, const float* dL_dpixN, const float* dL_dpixD,
                const float* dL_dpixO, double* out_color, double* out_normal, double* out_depth, double* out_opac,
                uint32_t* n_contrib_out, double* dL_dmean2D, double* dL_dconic, double* dL_dopacity,
                double* dL_dcolor, double* dL_dnormal, double* dL_ddepth) {
  const int gx = (W + 15) / 16, gy = (H + 15) / 16, T = gx * gy;
  const double mod = mod_f, tanx = tanx_f, tany = tany_f;
  const double fx = W / (2.0 * tanx), fy = H / (2.0 * tany);
  const int surface = config[0] > 0, normalize_depth = config[1] > 0, ppd = surface && config[2] > 0;
  double view[16], proj[16];
  for (int k = 0; k < 16; ++k) { view[k] = view_f[k]; proj[k] = proj_f[k]; }
  srec* rec = (srec*)calloc((size_t)P + 1, sizeof(srec));
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {   /* RS forward.cu:205-370 preprocessCUDA on the visible set */
    if (!(radii[i] > 0)) continue;
    d3 p = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    double ph[4];
    xf_point44(p, proj, ph);
    const double pw = 1.0 / (ph[3] + 0.0000001);
    const d3 pv = xf_point43(p, view);
    const m3 R = quat_R(rots + 4 * i);
    srec r;
    memset(&r, 0, sizeof(r));
    r.px = ((ph[0] * pw + 1.0) * W - 1.0) * 0.5;   /* ndc2Pix, RS auxiliary.h:42-46 */
    r.py = ((ph[1] * pw + 1.0) * H - 1.0) * 0.5;
    if (surface) {
      d3 c2 = {R.c[0][2], R.c[1][2], R.c[2][2]}, c0 = {R.c[0][0], R.c[1][0], R.c[2][0]}, c1 = {R.c[0][1], R.c[1][1], R.c[2][1]};
      const d3 nv = xf_vec43(c2, view), ax0 = xf_vec43(c0, view), ax1 = xf_vec43(c1, view);
      r.nx = nv.x; r.ny = nv.y; r.nz = nv.z;
      if (ppd) local_homo(pv, nv, fx, fy, ax0, ax1, &r.zx, &r.zy);
    }
    double c6[6], abc[3];
    cov3d(scales + 3 * i, mod, R, surface, c6);
    cov2d(pv, fx, fy, tanx, tany, c6, view, abc, 0);
    const double det = abc[0] * abc[2] - abc[1] * abc[1], di = 1.0 / det;
    r.ca = abc[2] * di; r.cb = -abc[1] * di; r.cc = abc[0] * di;
    r.o = opac[i];
    r.depth = pv.z;
    rec[i] = r;
  }
  const size_t HW = (size_t)W * H;
  memset(dL_dmean2D, 0, sizeof(double) * 2 * (size_t)P);
  memset(dL_dconic, 0, sizeof(double) * 3 * (size_t)P);
  memset(dL_dopacity, 0, sizeof(double) * (size_t)P);
  memset(dL_dcolor, 0, sizeof(double) * 3 * (size_t)P);
  memset(dL_dnormal, 0, sizeof(double) * 3 * (size_t)P);
  memset(dL_ddepth, 0, sizeof(double) * (size_t)P);
  const double ddelx_dx = 0.5 * W, ddely_dy = 0.5 * H;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < T; ++tile) {
    const uint32_t b = ranges[2 * tile], e = ranges[2 * tile + 1];
    const int tx = tile % gx, ty = tile / gx;
    for (int ly = 0; ly < 16; ++ly)
      for (int lx = 0; lx < 16; ++lx) {
        const int x = tx * 16 + lx, y = ty * 16 + ly;
        if (x >= W || y >= H) continue;
        const size_t pix = (size_t)W * y + x;
        /* ---- forward ---- */
        double Tt = 1.0, C[3] = {0, 0, 0}, N[3] = {0, 0, 0}, D = 0.0;
        uint32_t contributor = 0, last = 0;
        for (uint32_t k = b; k < e; ++k) {
          ++contributor;
          const uint32_t gi = point_list[k];
          const srec* g = rec + gi;
          const double dx = g->px - (double)x, dy = g->py - (double)y;
          const double dist = (g->ca * dx * dx + g->cc * dy * dy) + 2 * g->cb * dx * dy;
          const double power = -0.5 * dist;
          if (power > 0.0) continue;
          const double alpha = fmin(0.99, g->o * exp(power));
          if (alpha < g_alpha_min) continue;
          const double test_T = Tt * (1 - alpha);
          if (test_T < g_t_min) break;
          const double w = alpha * Tt;
          for (int ch = 0; ch < 3; ++ch) C[ch] += (double)colors[3 * gi + ch] * w;
          if (surface) { N[0] += g->nx * w; N[1] += g->ny * w; N[2] += g->nz * w; }
          D += (g->depth - (dx * g->zx + dy * g->zy)) * w;
          Tt = test_T;
          last = contributor;
        }
        const double T_final = fmin(1.0 - 0.000001, Tt);   /* RS forward.cu:608 */
        const double D_final = D;
        if (out_color) for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = C[ch] + T_final * bg[ch];
        if (out_normal) for (int ch = 0; ch < 3; ++ch) out_normal[ch * HW + pix] = N[ch];
        if (out_depth) out_depth[pix] = normalize_depth ? D / (1 - T_final) : D + T_final * 10;
        if (out_opac) out_opac[pix] = 1 - T_final;
        if (n_contrib_out) n_contrib_out[pix] = last;
        /* ---- backward (back to front) ---- */
        Tt = T_final;
        double accum_rec[3] = {0, 0, 0}, accum_rec_n[3] = {0, 0, 0}, accum_rec_d = 0;
        double last_alpha = 0, last_color[3] = {0, 0, 0}, last_normal[3] = {0, 0, 0}, last_depth = 0;
        const double dC[3] = {dL_dpixC[pix], dL_dpixC[HW + pix], dL_dpixC[2 * HW + pix]};
        const double dN[3] = {dL_dpixN[pix], dL_dpixN[HW + pix], dL_dpixN[2 * HW + pix]};
        const double dD = dL_dpixD[pix], dO = dL_dpixO[pix];
        const double bg_dot = bg[0] * dC[0] + bg[1] * dC[1] + bg[2] * dC[2];
        contributor = e - b;
        for (uint32_t k = e; k-- > b;) {
          contributor--;
          if (contributor >= last) continue;
          const uint32_t gi = point_list[k];
          const srec* g = rec + gi;
          const double dx = g->px - (double)x, dy = g->py - (double)y;
          const double dist = (g->ca * dx * dx + g->cc * dy * dy) + 2 * g->cb * dx * dy;
          const double power = -0.5 * dist;
          if (power > 0.0) continue;
          const double G = exp(power);
          const double alpha = fmin(0.99, g->o * G);
          if (alpha < g_alpha_min) continue;
          Tt = Tt / (1. - alpha);
          const double dchannel_dcolor = alpha * Tt;
          double dL_dalpha = 0.0;
          double add_col[3], add_n[3] = {0, 0, 0};
          for (int ch = 0; ch < 3; ++ch) {
            const double c_cur = colors[3 * gi + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1. - last_alpha) * accum_rec[ch];
            last_color[ch] = c_cur;
            dL_dalpha += (c_cur - accum_rec[ch]) * dC[ch];
            add_col[ch] = dchannel_dcolor * dC[ch];
          }
          if (surface) {
            const double nn[3] = {g->nx, g->ny, g->nz};
            for (int ch = 0; ch < 3; ++ch) {
              accum_rec_n[ch] = last_alpha * last_normal[ch] + (1. - last_alpha) * accum_rec_n[ch];
              last_normal[ch] = nn[ch];
              dL_dalpha += (nn[ch] - accum_rec_n[ch]) * dN[ch];
              add_n[ch] = dchannel_dcolor * dN[ch] * 10;   /* RS backward.cu:783 */
            }
          }
          double add_depth;
          {
            const double d_cur = g->depth - (dx * g->zx + dy * g->zy);
            accum_rec_d = last_alpha * last_depth + (1. - last_alpha) * accum_rec_d;
            last_depth = d_cur;
            double dL_dchannel = dD, dL_dalpha_depth = 0;
            if (normalize_depth) {   /* RS backward.cu:827-833 */
              dL_dchannel /= (1. - T_final);
              dL_dalpha_depth += dD * D_final / (1. - T_final) / (1. - T_final) * -T_final / (1 - alpha) / Tt;
            }
            dL_dalpha_depth += (d_cur - accum_rec_d) * dL_dchannel;
            add_depth = dchannel_dcolor * dL_dchannel;
            dL_dalpha += dL_dalpha_depth;
          }
          dL_dalpha *= Tt;
          dL_dalpha += dO * T_final / (1 - alpha);   /* RS backward.cu:847 */
          last_alpha = alpha;
          dL_dalpha += (-T_final / (1. - alpha)) * bg_dot;
          if (!normalize_depth) dL_dalpha += (-T_final / (1. - alpha)) * (10 * dD);   /* :858 */
          const double dL_ddist = dL_dalpha * g->o * -0.5 * G;
          double ndc_x = dL_ddist * 2 * (g->ca * dx + g->cb * dy) * ddelx_dx;
          double ndc_y = dL_ddist * 2 * (g->cc * dy + g->cb * dx) * ddely_dy;
          if (ppd) {   /* RS backward.cu:894-897: unweighted, raw dL_dpixD */
            ndc_x += -dD * g->zx;
            ndc_y += -dD * g->zy;
          }
          const double v_con[3] = {dL_ddist * (dx * dx), dL_ddist * (dx * dy), dL_ddist * (dy * dy)};
          const double v_op = G * dL_dalpha;
#pragma omp atomic
          dL_dmean2D[2 * gi] += ndc_x;
#pragma omp atomic
          dL_dmean2D[2 * gi + 1] += ndc_y;
          for (int ch = 0; ch < 3; ++ch) {
#pragma omp atomic
            dL_dconic[3 * gi + ch] += v_con[ch];
#pragma omp atomic
            dL_dcolor[3 * gi + ch] += add_col[ch];
#pragma omp atomic
            dL_dnormal[3 * gi + ch] += add_n[ch];
          }
#pragma omp atomic
          dL_dopacity[gi] += v_op;
#pragma omp atomic
          dL_ddepth[gi] += add_depth;
        }
      }
  }
  free(rec);
}

/* Per-surfel backward chain in fp64 from given blend sums (any source: rs64_blend, the CUDA implementation's
 * accumulation records, or the reference's dL_dconic / dL_dmean2D / dL_dnormal / dL_ddepth buffers).
 *   RS backward.cu:160-358 (computeCov2DCUDA), :457-582 (preprocessCUDA), :362-452 (computeCov3D).
 * Outputs (rows of invisible surfels are zero): dL_dmean3D [P,3], dL_dcov3D [P,6], dL_dscale [P,3],
 * dL_drot [P,4], dL_dtau [P,6]. */
void rs64_chain(int P, int W, int H, const float* means, const float* scales, float mod_f, const float* rots,
                const float* view_f, const float* proj_f, const float* proj_raw_f, float tanx_f, float tany_f,
                const float* config, const int* radii, const double* dL_dmean2D, const double* dL_dconic,
                const double* dL_dnormal, const double* dL_ddepth, double* dL_dmean3D, double* dL_dcov3D,
                double* dL_dscale, double* dL_drot, double* dL_dtau) {
  const double mod = mod_f, tanx = tanx_f, tany = tany_f;
  const double fx = W / (2.0 * tanx), fy = H / (2.0 * tany);
  const int surface = config[0] > 0;
  double view[16], proj[16], proj_raw[16];
  for (int k = 0; k < 16; ++k) { view[k] = view_f[k]; proj[k] = proj_f[k]; proj_raw[k] = proj_raw_f[k]; }
  memset(dL_dmean3D, 0, sizeof(double) * 3 * (size_t)P);
  memset(dL_dcov3D, 0, sizeof(double) * 6 * (size_t)P);
  memset(dL_dscale, 0, sizeof(double) * 3 * (size_t)P);
  memset(dL_drot, 0, sizeof(double) * 4 * (size_t)P);
  memset(dL_dtau, 0, sizeof(double) * 6 * (size_t)P);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    if (!(radii[i] > 0)) continue;
    const d3 mean = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    const m3 R = quat_R(rots + 4 * i);
    double c6[6], abc[3];
    cov2d_aux ax;
    cov3d(scales + 3 * i, mod, R, surface, c6);   /* the forward's cov3D (surfel z rule) is what :178 reads */
    cov2d(xf_point43(mean, view), fx, fy, tanx, tany, c6, view, abc, &ax);
    const double limx = 1.3 * tanx, limy = 1.3 * tany;
    const double xg = (ax.txtz < -limx || ax.txtz > limx) ? 0. : 1., yg = (ax.tytz < -limy || ax.tytz > limy) ? 0. : 1.;
    const double a = abc[0], bb = abc[1], c = abc[2];
    const double dcx = dL_dconic[3 * i], dcy = dL_dconic[3 * i + 1], dcz = dL_dconic[3 * i + 2];
    const double denom = a * c - bb * bb;
    const double d2 = 1.0 / ((denom * denom) + 0.0000001);
    double da = 0, db = 0, dc = 0, dcov[6] = {0, 0, 0, 0, 0, 0};
    const m3 Tm = ax.T, V = ax.V, Wm = ax.W, J = ax.J;
    if (d2 != 0) {   /* RS backward.cu:223-246 */
      da = d2 * (-c * c * dcx + 2 * bb * c * dcy + (denom - a * c) * dcz);
      dc = d2 * (-a * a * dcz + 2 * a * bb * dcy + (denom - a * c) * dcx);
      db = d2 * 2 * (bb * c * dcx - (denom + 2 * bb * bb) * dcy + a * bb * dcz);
      dcov[0] = Tm.c[0][0] * Tm.c[0][0] * da + Tm.c[0][0] * Tm.c[1][0] * db + Tm.c[1][0] * Tm.c[1][0] * dc;
      dcov[3] = Tm.c[0][1] * Tm.c[0][1] * da + Tm.c[0][1] * Tm.c[1][1] * db + Tm.c[1][1] * Tm.c[1][1] * dc;
      dcov[5] = Tm.c[0][2] * Tm.c[0][2] * da + Tm.c[0][2] * Tm.c[1][2] * db + Tm.c[1][2] * Tm.c[1][2] * dc;
      dcov[1] = 2 * Tm.c[0][0] * Tm.c[0][1] * da + (Tm.c[0][0] * Tm.c[1][1] + Tm.c[0][1] * Tm.c[1][0]) * db + 2 * Tm.c[1][0] * Tm.c[1][1] * dc;
      dcov[2] = 2 * Tm.c[0][0] * Tm.c[0][2] * da + (Tm.c[0][0] * Tm.c[1][2] + Tm.c[0][2] * Tm.c[1][0]) * db + 2 * Tm.c[1][0] * Tm.c[1][2] * dc;
      dcov[4] = 2 * Tm.c[0][2] * Tm.c[0][1] * da + (Tm.c[0][1] * Tm.c[1][2] + Tm.c[0][2] * Tm.c[1][1]) * db + 2 * Tm.c[1][1] * Tm.c[1][2] * dc;
    }
    for (int k = 0; k < 6; ++k) dL_dcov3D[6 * i + k] = dcov[k];
    double dT0[3], dT1[3];   /* RS backward.cu:257-268 */
    for (int k = 0; k < 3; ++k) {
      const double tv0 = Tm.c[0][0] * V.c[k][0] + Tm.c[0][1] * V.c[k][1] + Tm.c[0][2] * V.c[k][2];
      const double tv1 = Tm.c[1][0] * V.c[k][0] + Tm.c[1][1] * V.c[k][1] + Tm.c[1][2] * V.c[k][2];
      dT0[k] = 2 * tv0 * da + tv1 * db;
      dT1[k] = 2 * tv1 * dc + tv0 * db;
    }
    const double dJ00 = Wm.c[0][0] * dT0[0] + Wm.c[0][1] * dT0[1] + Wm.c[0][2] * dT0[2];
    const double dJ02 = Wm.c[2][0] * dT0[0] + Wm.c[2][1] * dT0[1] + Wm.c[2][2] * dT0[2];
    const double dJ11 = Wm.c[1][0] * dT1[0] + Wm.c[1][1] * dT1[1] + Wm.c[1][2] * dT1[2];
    const double dJ12 = Wm.c[2][0] * dT1[0] + Wm.c[2][1] * dT1[1] + Wm.c[2][2] * dT1[2];
    const d3 t = ax.t;
    const double tz = 1. / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    const d3 dLt = {xg * -fx * tz2 * dJ02, yg * -fy * tz2 * dJ12,
                    -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t.x) * tz3 * dJ02 + (2 * fy * t.y) * tz3 * dJ12};
    double tau[6] = {dLt.x, dLt.y, dLt.z, 0, 0, 0};   /* rho via I, theta via -skew(clamped t): :288-303 */
    const d3 th = cross(t, dLt);
    tau[3] += th.x; tau[4] += th.y; tau[5] += th.z;
    d3 dmean = xf_vec43T(dLt, view);   /* assigned, :311 */
    {   /* rotation part of W: :313-357 */
      const d3 dW1 = {J.c[0][0] * dT0[0], J.c[1][1] * dT1[0], J.c[0][2] * dT0[0] + J.c[1][2] * dT1[0]};
      const d3 dW2 = {J.c[0][0] * dT0[1], J.c[1][1] * dT1[1], J.c[0][2] * dT0[1] + J.c[1][2] * dT1[1]};
      const d3 dW3 = {J.c[0][0] * dT0[2], J.c[1][1] * dT1[2], J.c[0][2] * dT0[2] + J.c[1][2] * dT1[2]};
      const d3 c1 = {view[0], view[1], view[2]}, c2 = {view[4], view[5], view[6]}, c3 = {view[8], view[9], view[10]};
      const d3 x1 = cross(c1, dW1), x2 = cross(c2, dW2), x3 = cross(c3, dW3);
      tau[3] += x1.x + x2.x + x3.x; tau[4] += x1.y + x2.y + x3.y; tau[5] += x1.z + x2.z + x3.z;
    }
    /* preprocessCUDA bwd: RS backward.cu:487-571 */
    double mh[4];
    xf_point44(mean, proj, mh);
    const double m_w = 1.0 / (mh[3] + 0.0000001);
    const double mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    const double mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    const double d2x = dL_dmean2D[2 * i], d2y = dL_dmean2D[2 * i + 1];
    dmean.x += (proj[0] * m_w - proj[3] * mul1) * d2x + (proj[1] * m_w - proj[3] * mul2) * d2y;
    dmean.y += (proj[4] * m_w - proj[7] * mul1) * d2x + (proj[5] * m_w - proj[7] * mul2) * d2y;
    dmean.z += (proj[8] * m_w - proj[11] * mul1) * d2x + (proj[9] * m_w - proj[11] * mul2) * d2y;
    {
      const double al = 1.0 * m_w, be = -mh[0] * m_w * m_w, ga = -mh[1] * m_w * m_w;
      const double pa = proj_raw[0], pb = proj_raw[5], pe = proj_raw[11];
      const d3 pC = xf_point43(mean, view);   /* unclamped (q13) */
      const d3 d1 = {al * pa, 0., be * pe}, d2v = {0., al * pb, ga * pe};
      const d3 d1t = cross(pC, d1), d2t = cross(pC, d2v);
      tau[0] += d2x * d1.x + d2y * d2v.x; tau[1] += d2x * d1.y + d2y * d2v.y; tau[2] += d2x * d1.z + d2y * d2v.z;
      tau[3] += d2x * d1t.x + d2y * d2t.x; tau[4] += d2x * d1t.y + d2y * d2t.y; tau[5] += d2x * d1t.z + d2y * d2t.z;
      const double dd = dL_ddepth[i];
      dmean.x += dd * view[2]; dmean.y += dd * view[6]; dmean.z += dd * view[10];
      tau[2] += dd; tau[3] += dd * pC.y; tau[4] += dd * -pC.x;
    }
    dL_dmean3D[3 * i] = dmean.x; dL_dmean3D[3 * i + 1] = dmean.y; dL_dmean3D[3 * i + 2] = dmean.z;
    for (int k = 0; k < 6; ++k) dL_dtau[6 * i + k] = tau[k];
    {   /* computeCov3D bwd: RS backward.cu:362-452 (S keeps the un-zeroed s.z = mod * scale.z, q5) */
      const float* q = rots + 4 * i;
      const double r = q[0], x = q[1], y = q[2], z = q[3];
      const double sx = mod * scales[3 * i], sy = mod * scales[3 * i + 1], sz = mod * scales[3 * i + 2];
      m3 S = {{{sx, 0, 0}, {0, sy, 0}, {0, 0, sz}}};
      m3 M = m3mul(S, R);
      m3 dSig = {{{dcov[0], 0.5 * dcov[1], 0.5 * dcov[2]}, {0.5 * dcov[1], dcov[3], 0.5 * dcov[4]},
                  {0.5 * dcov[2], 0.5 * dcov[4], dcov[5]}}};
      m3 M2;
      for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) M2.c[j][k] = 2.0 * M.c[j][k];
      m3 dM = m3mul(M2, dSig);
      m3 Rt = m3t(R), dMt = m3t(dM);
      for (int k = 0; k < 3; ++k)
        dL_dscale[3 * i + k] = Rt.c[k][0] * dMt.c[k][0] + Rt.c[k][1] * dMt.c[k][1] + Rt.c[k][2] * dMt.c[k][2];
      if (surface) dL_dscale[3 * i + 2] = 0;   /* :419 */
      for (int k = 0; k < 3; ++k) { dMt.c[0][k] *= sx; dMt.c[1][k] *= sy; dMt.c[2][k] *= sz; }
      {   /* normal -> last row of dL_dRt through the view matrix, :427-436 */
        const d3 dn = {dL_dnormal[3 * i], dL_dnormal[3 * i + 1], dL_dnormal[3 * i + 2]};
        const d3 dw = xf_vec43T(dn, view);
        dMt.c[2][0] += dw.x; dMt.c[2][1] += dw.y; dMt.c[2][2] += dw.z;
      }
      dL_drot[4 * i + 0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
      dL_drot[4 * i + 1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
      dL_drot[4 * i + 2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
      dL_drot[4 * i + 3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
    }
  }
}
