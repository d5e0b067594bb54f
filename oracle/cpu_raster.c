/* TEST INFRASTRUCTURE — not part of the product.
 *
 * Plain-C (fp32, OpenMP over tiles / Gaussians) restatement of the reference 3DGS rasterizer with
 * camera-pose gradients, submodules/diff-gaussian-rasterization-w-pose (R3).  Each function cites
 * the reference lines it follows.  Uses: CPU-side oracle checks (tests/test_oracle_raster.py), a
 * GPU-box cross-check against the reference CUDA build, and bench.py's timed `cpu_baseline`
 * (the reference has no CPU rasterizer of its own).
 *
 * Parity: UNPINNED by reference golden vectors (none exist for this path); pinned instead against
 * the reference CUDA build run on the GPU box (tests/test_raster3_gpu.py::test_cpu_oracle_vs_reference).
 * gcc contracts a*b+c differently from nvcc, so integer outputs (radii, tile lists) may differ from
 * the CUDA reference on the rare Gaussian whose rounding straddles a tile edge; images and
 * gradients agree to fp32 tolerance.
 *
 * Only the colors_precomp + scale/rotation path (the one PINGS uses, GR:110,321) is restated.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float x, y, z; } f3;

/* R3 auxiliary.h:58-97 */
static f3 xf_point43(f3 p, const float* m) {
  f3 r = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
          m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
  return r;
}
static void xf_point44(f3 p, const float* m, float* o) {
  o[0] = m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12];
  o[1] = m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13];
  o[2] = m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14];
  o[3] = m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15];
}
static f3 xf_vec43T(f3 p, const float* m) {
  f3 r = {m[0] * p.x + m[1] * p.y + m[2] * p.z, m[4] * p.x + m[5] * p.y + m[6] * p.z,
          m[8] * p.x + m[9] * p.y + m[10] * p.z};
  return r;
}
static f3 cross(f3 a, f3 b) {
  f3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
  return r;
}
/* column-major 3x3 like glm::mat3: m[col][row] */
typedef struct { float c[3][3]; } m3;
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
static m3 quat_R(const float* q) { /* R3 forward.cu:129-142 (un-normalised quaternion) */
  float r = q[0], x = q[1], y = q[2], z = q[3];
  m3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
  return R;
}
/* R3 forward.cu:120-154 computeCov3D */
static void cov3d(const float* s, float mod, const float* q, float* c6) {
  m3 S = {{{mod * s[0], 0, 0}, {0, mod * s[1], 0}, {0, 0, mod * s[2]}}};
  m3 M = m3mul(S, quat_R(q));
  m3 Sg = m3mul(m3t(M), M);
  c6[0] = Sg.c[0][0]; c6[1] = Sg.c[0][1]; c6[2] = Sg.c[0][2];
  c6[3] = Sg.c[1][1]; c6[4] = Sg.c[1][2]; c6[5] = Sg.c[2][2];
}
typedef struct { f3 t; float txtz, tytz; m3 J, W, T, V; } cov2d_aux;
/* R3 forward.cu:76-115 computeCov2D */
static void cov2d(f3 mean, float fx, float fy, float tanx, float tany, const float* c6, const float* view,
                  float* abc, cov2d_aux* ax) {
  f3 t = xf_point43(mean, view);
  float limx = 1.3f * tanx, limy = 1.3f * tany;
  float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  m3 J = {{{fx / t.z, 0, -(fx * t.x) / (t.z * t.z)}, {0, fy / t.z, -(fy * t.y) / (t.z * t.z)}, {0, 0, 0}}};
  m3 W = {{{view[0], view[4], view[8]}, {view[1], view[5], view[9]}, {view[2], view[6], view[10]}}};
  m3 T = m3mul(W, J);
  m3 V = {{{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}}};
  m3 cov = m3mul(m3mul(m3t(T), m3t(V)), T);
  abc[0] = cov.c[0][0] + 0.3f;
  abc[1] = cov.c[0][1];
  abc[2] = cov.c[1][1] + 0.3f;
  if (ax) { ax->t = t; ax->txtz = txtz; ax->tytz = tytz; ax->J = J; ax->W = W; ax->T = T; ax->V = V; }
}
static float ndc2pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); } /* R3 auxiliary.h:41-44 */
static int clampi(int v, int hi) { return v < 0 ? 0 : (v > hi ? hi : v); }

typedef struct {
  int P, W, H, gx, gy;
  int64_t R;
  float *xy, *conic_o, *depth;   /* [P,2] [P,4] [P] */
  int* rect;                     /* [P,4] x0,y0,x1,y1 */
  uint32_t* point_list;          /* [R] */
  uint32_t* ranges;              /* [T,2] */
  float* final_T;                /* [W*H] */
  uint32_t* n_contrib;           /* [W*H] */
} cr_state;

typedef struct { uint64_t key; uint32_t val, seq; } kv;
static int kv_cmp(const void* a, const void* b) {
  const kv *x = (const kv*)a, *y = (const kv*)b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->seq < y->seq ? -1 : (x->seq > y->seq);
}

void cr_free(cr_state* s) {
  if (!s) return;
  free(s->xy); free(s->conic_o); free(s->depth); free(s->rect); free(s->point_list); free(s->ranges);
  free(s->final_T); free(s->n_contrib); free(s);
}
int64_t cr_num_rendered(const cr_state* s) { return s->R; }
const uint32_t* cr_point_list(const cr_state* s) { return s->point_list; }
const uint32_t* cr_ranges(const cr_state* s) { return s->ranges; }
const uint32_t* cr_n_contrib(const cr_state* s) { return s->n_contrib; }

/* Forward: R3 rasterizer_impl.cu:198-344 (pipeline), forward.cu:157-258 (preprocess), :263-392 (render),
 * rasterizer_impl.cu:70-138 (duplicateWithKeys, identifyTileRanges). Returns the state kept for backward. */
cr_state* cr_forward(int P, int W, int H, const float* bg, const float* means, const float* colors,
                     const float* opac, const float* scales, float mod, const float* rots, const float* view,
                     const float* proj, float tanx, float tany, float* out_color, float* out_depth,
                     float* out_opacity, int* radii, int* n_touched) {
  cr_state* s = (cr_state*)calloc(1, sizeof(cr_state));
  const int gx = (W + 15) / 16, gy = (H + 15) / 16, T = gx * gy;
  const float fx = W / (2.0f * tanx), fy = H / (2.0f * tany);
  s->P = P; s->W = W; s->H = H; s->gx = gx; s->gy = gy;
  s->xy = (float*)calloc((size_t)P * 2 + 1, 4);
  s->conic_o = (float*)calloc((size_t)P * 4 + 1, 4);
  s->depth = (float*)calloc((size_t)P + 1, 4);
  s->rect = (int*)calloc((size_t)P * 4 + 1, 4);
  uint32_t* tiles = (uint32_t*)calloc((size_t)P + 1, 4);
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    radii[i] = 0;
    n_touched[i] = 0;
    f3 p = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    f3 pv = xf_point43(p, view);
    if (pv.z <= 0.2f) continue; /* in_frustum, R3 auxiliary.h:139-164 */
    float ph[4];
    xf_point44(p, proj, ph);
    float pw = 1.0f / (ph[3] + 0.0000001f);
    float c6[6], abc[3];
    cov3d(scales + 3 * i, mod, rots + 4 * i, c6);
    cov2d(p, fx, fy, tanx, tany, c6, view, abc, 0);
    float det = abc[0] * abc[2] - abc[1] * abc[1];
    if (det == 0.0f) continue;
    float di = 1.f / det;
    float mid = 0.5f * (abc[0] + abc[2]);
    float l1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det)), l2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
    float rad = ceilf(3.f * sqrtf(fmaxf(l1, l2)));
    float px = ndc2pix(ph[0] * pw, W), py = ndc2pix(ph[1] * pw, H);
    int r = (int)rad; /* getRect, R3 auxiliary.h:46-56 */
    int x0 = clampi((int)((px - r) / 16), gx), y0 = clampi((int)((py - r) / 16), gy);
    int x1 = clampi((int)((px + r + 15) / 16), gx), y1 = clampi((int)((py + r + 15) / 16), gy);
    if ((x1 - x0) * (y1 - y0) == 0) continue;
    s->depth[i] = pv.z;
    radii[i] = r;
    s->xy[2 * i] = px; s->xy[2 * i + 1] = py;
    s->conic_o[4 * i] = abc[2] * di; s->conic_o[4 * i + 1] = -abc[1] * di; s->conic_o[4 * i + 2] = abc[0] * di;
    s->conic_o[4 * i + 3] = opac[i];
    s->rect[4 * i] = x0; s->rect[4 * i + 1] = y0; s->rect[4 * i + 2] = x1; s->rect[4 * i + 3] = y1;
    tiles[i] = (uint32_t)((x1 - x0) * (y1 - y0));
  }
  int64_t R = 0;
  for (int i = 0; i < P; ++i) R += tiles[i];
  s->R = R;
  kv* list = (kv*)malloc(sizeof(kv) * (size_t)(R > 0 ? R : 1));
  {
    int64_t off = 0;
    for (int i = 0; i < P; ++i) {
      if (radii[i] <= 0) continue;
      uint32_t db;
      memcpy(&db, &s->depth[i], 4);
      for (int y = s->rect[4 * i + 1]; y < s->rect[4 * i + 3]; ++y)
        for (int x = s->rect[4 * i]; x < s->rect[4 * i + 2]; ++x) {
          list[off].key = ((uint64_t)(y * gx + x) << 32) | db;
          list[off].val = (uint32_t)i;
          list[off].seq = (uint32_t)off;
          ++off;
        }
    }
  }
  qsort(list, (size_t)R, sizeof(kv), kv_cmp); /* stable via seq: cub::DeviceRadixSort is stable */
  s->point_list = (uint32_t*)malloc(4 * (size_t)(R > 0 ? R : 1));
  s->ranges = (uint32_t*)calloc((size_t)T * 2, 4);
  for (int64_t k = 0; k < R; ++k) {
    s->point_list[k] = list[k].val;
    uint32_t t = (uint32_t)(list[k].key >> 32);
    if (k == 0) s->ranges[2 * t] = 0;
    else {
      uint32_t pt = (uint32_t)(list[k - 1].key >> 32);
      if (pt != t) { s->ranges[2 * pt + 1] = (uint32_t)k; s->ranges[2 * t] = (uint32_t)k; }
    }
    if (k == R - 1) s->ranges[2 * t + 1] = (uint32_t)R;
  }
  free(list);
  free(tiles);
  s->final_T = (float*)malloc(4 * (size_t)W * H);
  s->n_contrib = (uint32_t*)malloc(4 * (size_t)W * H);
  const size_t HW = (size_t)W * H;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < T; ++tile) {
    const uint32_t b = s->ranges[2 * tile], e = s->ranges[2 * tile + 1];
    const int tx = tile % gx, ty = tile / gx;
    for (int ly = 0; ly < 16; ++ly)
      for (int lx = 0; lx < 16; ++lx) {
        const int x = tx * 16 + lx, y = ty * 16 + ly;
        if (x >= W || y >= H) continue;
        float Tt = 1.0f, C[3] = {0, 0, 0}, D = 0.f;
        uint32_t contributor = 0, last = 0;
        for (uint32_t k = b; k < e; ++k) {
          ++contributor;
          const uint32_t g = s->point_list[k];
          const float dx = s->xy[2 * g] - (float)x, dy = s->xy[2 * g + 1] - (float)y;
          const float* co = s->conic_o + 4 * g;
          const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          if (power > 0.0f) continue;
          const float alpha = fminf(0.99f, co[3] * expf(power));
          if (alpha < 1.0f / 255.0f) continue;
          const float test_T = Tt * (1 - alpha);
          if (test_T < 0.0001f) break;
          for (int ch = 0; ch < 3; ++ch) C[ch] += colors[3 * g + ch] * alpha * Tt;
          D += s->depth[g] * alpha * Tt;
          if (test_T > 0.5f) {
#pragma omp atomic
            n_touched[g] += 1;
          }
          Tt = test_T;
          last = contributor;
        }
        const size_t pix = (size_t)W * y + x;
        s->final_T[pix] = Tt;
        s->n_contrib[pix] = last;
        for (int ch = 0; ch < 3; ++ch) out_color[ch * HW + pix] = C[ch] + Tt * bg[ch];
        out_depth[pix] = D;
        out_opacity[pix] = 1 - Tt;
      }
  }
  return s;
}

/* Backward: R3 backward.cu:563-787 (renderCUDA), :150-346 (computeCov2DCUDA), :418-539 (preprocessCUDA),
 * :350-413 (computeCov3D).  All outputs must be zero-initialised by the caller ([P,*] arrays as in
 * R3 rasterize_points.cu:160-170).  dL_dconic is [P,4] (x, y, -, w), dL_dtau is [P,6]. */
void cr_backward(const cr_state* s, const float* bg, const float* means, const float* colors, const float* scales,
                 float mod, const float* rots, const float* view, const float* proj, const float* proj_raw,
                 float tanx, float tany, const int* radii, const float* dL_dpix, const float* dL_dpix_depth,
                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_ddepth,
                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dscale, float* dL_drot, float* dL_dtau) {
  const int P = s->P, W = s->W, H = s->H, gx = s->gx, T = s->gx * s->gy;
  const size_t HW = (size_t)W * H;
  const float fx = W / (2.0f * tanx), fy = H / (2.0f * tany);
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;
#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < T; ++tile) {
    const uint32_t b = s->ranges[2 * tile], e = s->ranges[2 * tile + 1];
    const int tx = tile % gx, ty = tile / gx;
    for (int ly = 0; ly < 16; ++ly)
      for (int lx = 0; lx < 16; ++lx) {
        const int x = tx * 16 + lx, y = ty * 16 + ly;
        if (x >= W || y >= H) continue;
        const size_t pix = (size_t)W * y + x;
        const float T_final = s->final_T[pix];
        float Tt = T_final;
        const uint32_t last_contributor = s->n_contrib[pix];
        float accum_rec[3] = {0, 0, 0}, accum_d = 0, last_alpha = 0, last_color[3] = {0, 0, 0}, last_depth = 0;
        const float dpix[3] = {dL_dpix[pix], dL_dpix[HW + pix], dL_dpix[2 * HW + pix]};
        const float dpd = dL_dpix_depth[pix];
        const float bg_dot = bg[0] * dpix[0] + bg[1] * dpix[1] + bg[2] * dpix[2];
        uint32_t contributor = e - b;
        for (uint32_t k = e; k-- > b;) {
          contributor--;
          if (contributor >= last_contributor) continue;
          const uint32_t g = s->point_list[k];
          const float dx = s->xy[2 * g] - (float)x, dy = s->xy[2 * g + 1] - (float)y;
          const float* co = s->conic_o + 4 * g;
          const float power = -0.5f * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
          if (power > 0.0f) continue;
          const float G = expf(power);
          const float alpha = fminf(0.99f, co[3] * G);
          if (alpha < 1.0f / 255.0f) continue;
          Tt = Tt / (1.f - alpha);
          const float dchannel_dcolor = alpha * Tt;
          float dL_dalpha = 0.0f;
          for (int ch = 0; ch < 3; ++ch) {
            const float c = colors[3 * g + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
            last_color[ch] = c;
            dL_dalpha += (c - accum_rec[ch]) * dpix[ch];
            const float v = dchannel_dcolor * dpix[ch];
#pragma omp atomic
            dL_dcolor[3 * g + ch] += v;
          }
          const float c_d = s->depth[g];
          accum_d = last_alpha * last_depth + (1.f - last_alpha) * accum_d;
          last_depth = c_d;
          dL_dalpha += (c_d - accum_d) * dpd;
          {
            const float v = dchannel_dcolor * dpd;
#pragma omp atomic
            dL_ddepth[g] += v;
          }
          dL_dalpha *= Tt;
          last_alpha = alpha;
          dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot;
          const float dL_dG = co[3] * dL_dalpha;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * co[0] - gdy * co[1];
          const float dG_ddely = -gdy * co[2] - gdx * co[1];
          const float v0 = dL_dG * dG_ddelx * ddelx_dx, v1 = dL_dG * dG_ddely * ddely_dy;
          const float v2 = -0.5f * gdx * dx * dL_dG, v3 = -0.5f * gdx * dy * dL_dG, v4 = -0.5f * gdy * dy * dL_dG;
          const float v5 = G * dL_dalpha;
#pragma omp atomic
          dL_dmean2D[3 * g] += v0;
#pragma omp atomic
          dL_dmean2D[3 * g + 1] += v1;
#pragma omp atomic
          dL_dconic[4 * g] += v2;
#pragma omp atomic
          dL_dconic[4 * g + 1] += v3;
#pragma omp atomic
          dL_dconic[4 * g + 3] += v4;
#pragma omp atomic
          dL_dopacity[g] += v5;
        }
      }
  }
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P; ++i) {
    if (!(radii[i] > 0)) continue;
    f3 mean = {means[3 * i], means[3 * i + 1], means[3 * i + 2]};
    float c6[6], abc[3];
    cov2d_aux ax;
    cov3d(scales + 3 * i, mod, rots + 4 * i, c6);
    cov2d(mean, fx, fy, tanx, tany, c6, view, abc, &ax);
    const float limx = 1.3f * tanx, limy = 1.3f * tany;
    const float xg = (ax.txtz < -limx || ax.txtz > limx) ? 0.f : 1.f, yg = (ax.tytz < -limy || ax.tytz > limy) ? 0.f : 1.f;
    const float a = abc[0], bb = abc[1], c = abc[2];
    const float dcx = dL_dconic[4 * i], dcy = dL_dconic[4 * i + 1], dcz = dL_dconic[4 * i + 3];
    const float denom = a * c - bb * bb;
    const float d2 = 1.0f / ((denom * denom) + 0.0000001f);
    float da = 0, db = 0, dc = 0, dcov[6] = {0, 0, 0, 0, 0, 0};
    const m3 Tm = ax.T, V = ax.V, Wm = ax.W, J = ax.J;
    if (d2 != 0) { /* R3 backward.cu:205-232 */
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
    float dT0[3], dT1[3]; /* R3 backward.cu:242-253 */
    for (int k = 0; k < 3; ++k) {
      const float tv0 = Tm.c[0][0] * V.c[k][0] + Tm.c[0][1] * V.c[k][1] + Tm.c[0][2] * V.c[k][2];
      const float tv1 = Tm.c[1][0] * V.c[k][0] + Tm.c[1][1] * V.c[k][1] + Tm.c[1][2] * V.c[k][2];
      dT0[k] = 2 * tv0 * da + tv1 * db;
      dT1[k] = 2 * tv1 * dc + tv0 * db;
    }
    const float dJ00 = Wm.c[0][0] * dT0[0] + Wm.c[0][1] * dT0[1] + Wm.c[0][2] * dT0[2];
    const float dJ02 = Wm.c[2][0] * dT0[0] + Wm.c[2][1] * dT0[1] + Wm.c[2][2] * dT0[2];
    const float dJ11 = Wm.c[1][0] * dT1[0] + Wm.c[1][1] * dT1[1] + Wm.c[1][2] * dT1[2];
    const float dJ12 = Wm.c[2][0] * dT1[0] + Wm.c[2][1] * dT1[1] + Wm.c[2][2] * dT1[2];
    const f3 t = ax.t;
    const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
    f3 dLt = {xg * -fx * tz2 * dJ02, yg * -fy * tz2 * dJ12,
              -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t.x) * tz3 * dJ02 + (2 * fy * t.y) * tz3 * dJ12};
    float tau[6] = {dLt.x, dLt.y, dLt.z, 0, 0, 0}; /* rho via identity, theta via -skew(t): R3 :273-293 */
    f3 th = cross(t, dLt);
    tau[3] += th.x; tau[4] += th.y; tau[5] += th.z;
    f3 dmean = xf_vec43T(dLt, view); /* assigned, R3 :297 */
    { /* rotation part of W: R3 :299-343 */
      f3 dW1 = {J.c[0][0] * dT0[0], J.c[1][1] * dT1[0], J.c[0][2] * dT0[0] + J.c[1][2] * dT1[0]};
      f3 dW2 = {J.c[0][0] * dT0[1], J.c[1][1] * dT1[1], J.c[0][2] * dT0[1] + J.c[1][2] * dT1[1]};
      f3 dW3 = {J.c[0][0] * dT0[2], J.c[1][1] * dT1[2], J.c[0][2] * dT0[2] + J.c[1][2] * dT1[2]};
      f3 c1 = {view[0], view[1], view[2]}, c2 = {view[4], view[5], view[6]}, c3 = {view[8], view[9], view[10]};
      f3 x1 = cross(c1, dW1), x2 = cross(c2, dW2), x3 = cross(c3, dW3);
      tau[3] += x1.x + x2.x + x3.x; tau[4] += x1.y + x2.y + x3.y; tau[5] += x1.z + x2.z + x3.z;
    }
    /* preprocessCUDA bwd: R3 :440-528 */
    float mh[4];
    xf_point44(mean, proj, mh);
    const float m_w = 1.0f / (mh[3] + 0.0000001f);
    const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
    const float d2x = dL_dmean2D[3 * i], d2y = dL_dmean2D[3 * i + 1];
    dmean.x += (proj[0] * m_w - proj[3] * mul1) * d2x + (proj[1] * m_w - proj[3] * mul2) * d2y;
    dmean.y += (proj[4] * m_w - proj[7] * mul1) * d2x + (proj[5] * m_w - proj[7] * mul2) * d2y;
    dmean.z += (proj[8] * m_w - proj[11] * mul1) * d2x + (proj[9] * m_w - proj[11] * mul2) * d2y;
    {
      const float al = 1.0f * m_w, be = -mh[0] * m_w * m_w, ga = -mh[1] * m_w * m_w;
      const float pa = proj_raw[0], pb = proj_raw[5], pe = proj_raw[11];
      const f3 pC = xf_point43(mean, view);
      f3 d1 = {al * pa, 0.f, be * pe}, d2v = {0.f, al * pb, ga * pe};
      f3 d1t = cross(pC, d1), d2t = cross(pC, d2v);
      tau[0] += d2x * d1.x + d2y * d2v.x; tau[1] += d2x * d1.y + d2y * d2v.y; tau[2] += d2x * d1.z + d2y * d2v.z;
      tau[3] += d2x * d1t.x + d2y * d2t.x; tau[4] += d2x * d1t.y + d2y * d2t.y; tau[5] += d2x * d1t.z + d2y * d2t.z;
      const float dd = dL_ddepth[i];
      dmean.x += dd * view[2]; dmean.y += dd * view[6]; dmean.z += dd * view[10];
      tau[2] += dd; tau[3] += dd * pC.y; tau[4] += dd * -pC.x;
    }
    dL_dmean3D[3 * i] = dmean.x; dL_dmean3D[3 * i + 1] = dmean.y; dL_dmean3D[3 * i + 2] = dmean.z;
    for (int k = 0; k < 6; ++k) dL_dtau[6 * i + k] = tau[k];
    { /* computeCov3D bwd: R3 :350-413 */
      const float* q = rots + 4 * i;
      const float r = q[0], x = q[1], y = q[2], z = q[3];
      const m3 R = quat_R(q);
      const float sx = mod * scales[3 * i], sy = mod * scales[3 * i + 1], sz = mod * scales[3 * i + 2];
      m3 S = {{{sx, 0, 0}, {0, sy, 0}, {0, 0, sz}}};
      m3 M = m3mul(S, R);
      m3 dSig = {{{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]}, {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                  {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}}};
      m3 M2;
      for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) M2.c[j][k] = 2.0f * M.c[j][k];
      m3 dM = m3mul(M2, dSig);
      m3 Rt = m3t(R), dMt = m3t(dM);
      for (int k = 0; k < 3; ++k)
        dL_dscale[3 * i + k] = Rt.c[k][0] * dMt.c[k][0] + Rt.c[k][1] * dMt.c[k][1] + Rt.c[k][2] * dMt.c[k][2];
      for (int k = 0; k < 3; ++k) { dMt.c[0][k] *= sx; dMt.c[1][k] *= sy; dMt.c[2][k] *= sz; }
      dL_drot[4 * i + 0] = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
      dL_drot[4 * i + 1] = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
      dL_drot[4 * i + 2] = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
      dL_drot[4 * i + 3] = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
    }
  }
}
