// C-ABI entry points of the rasterizer (see include/pings_b200.h).
#include <cstdlib>
#include "pg_rasters_bwd.cuh"

namespace pg {
struct HostSlot {
  unsigned long long* pinned = nullptr;
  cudaEvent_t ev = nullptr;
  int device = -1;
};
static thread_local HostSlot g_slot;

static int ensure_slot() {
  int dev = 0;
  PG_CUDA(cudaGetDevice(&dev));
  if (g_slot.pinned && g_slot.device == dev) return PG_OK;
  if (g_slot.pinned) {
    cudaFreeHost(g_slot.pinned);
    cudaEventDestroy(g_slot.ev);
    g_slot.pinned = nullptr;
  }
  PG_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&g_slot.pinned), 64, cudaHostAllocDefault));
  PG_CUDA(cudaEventCreateWithFlags(&g_slot.ev, cudaEventDisableTiming));
  g_slot.device = dev;
  return PG_OK;
}

static bool params_ok(const PgRasterParams* p) {
  if (!p) return false;
  if (p->variant != PG_VARIANT_3DGS && p->variant != PG_VARIANT_SURFEL) return false;
  if (p->P < 0 || p->image_width <= 0 || p->image_height <= 0) return false;
  if (!p->viewmatrix || !p->projmatrix || !p->bg) return false;
  if (p->variant == PG_VARIANT_SURFEL && (!p->patch_bbox || !p->config)) return false;
  return true;
}
}  // namespace pg

using namespace pg;

extern "C" {

size_t pg_raster_geom_bytes(int variant, int P) { return carve_geom(variant, P, nullptr, nullptr); }
size_t pg_raster_image_bytes(int variant, int width, int height) {
  return carve_image(variant, width, height, nullptr, nullptr);
}
size_t pg_raster_binning_bytes(int variant, int width, int height, int64_t num_rendered) {
  (void)variant;
  return carve_binning(num_rendered, width, height, nullptr, nullptr);
}

int pg_mark_visible(int variant, int P, const float* means3D, const float* viewmatrix,
                    const float* projmatrix, uint8_t* present, void* stream) {
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return PG_ERR_ARG;
  if (variant != PG_VARIANT_3DGS && variant != PG_VARIANT_SURFEL) return PG_ERR_ARG;
  if (variant == PG_VARIANT_SURFEL && P > 0 && !projmatrix) return PG_ERR_ARG;
  if (P == 0) return PG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (variant == PG_VARIANT_SURFEL)
    rs_mark_visible_kernel<<<div_up(P, 256), 256, 0, st>>>(P, means3D, viewmatrix, projmatrix, present);
  else
    r3_mark_visible_kernel<<<div_up(P, 256), 256, 0, st>>>(P, means3D, viewmatrix, present);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

// wait = false: everything is queued (count copy + event included) but the host does not wait for the count.
static int stage1_impl(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, size_t geom_bytes,
                       int32_t* radii, int64_t* num_rendered, void* stream, bool wait) {
  if (!params_ok(prm) || !g || !num_rendered) return PG_ERR_ARG;
  *num_rendered = 0;
  const int P = prm->P;
  if (P == 0) return PG_OK;
  if (!geom_ws || !radii || !g->means3D || !g->opacities) return PG_ERR_ARG;
  if (!g->colors_precomp && !g->shs) return PG_ERR_ARG;
  if (!g->cov3D_precomp && (!g->scales || !g->rotations)) return PG_ERR_ARG;
  if (!g->colors_precomp && !prm->campos) return PG_ERR_ARG;
  const bool surfel = prm->variant == PG_VARIANT_SURFEL;
  if (surfel && !g->rotations) return PG_ERR_ARG;  // RS forward.cu:263 reads rotations unconditionally
  const int W = prm->image_width, H = prm->image_height;
  const int gx = div_up(W, kTile), gy = div_up(H, kTile);
  if ((int64_t)gx * gy > 65535) return PG_ERR_UNSUPPORTED;
  if (P >= (1 << kBandIdBits)) return PG_ERR_UNSUPPORTED;   // band binning tags ids with 5 tile bits
  if (geom_bytes < carve_geom(prm->variant, P, nullptr, nullptr)) return PG_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  int rc = ensure_slot();
  if (rc != PG_OK) return rc;
  GeomWs ws;
  carve_geom(prm->variant, P, geom_ws, &ws);
  PG_CUDA(cudaMemsetAsync(ws.total, 0, geom_zero_bytes(P, band_path(gx, gy) && !sort_force_radix()), st));
  R3PreArgs a;
  a.P = P; a.D = prm->sh_degree; a.M = prm->sh_coeffs; a.W = W; a.H = H; a.gx = gx; a.gy = gy;
  a.means = g->means3D; a.scales = g->scales; a.rots = g->rotations; a.opac = g->opacities; a.shs = g->shs;
  a.cov3D_pre = g->cov3D_precomp; a.colors_pre = g->colors_precomp;
  a.view = prm->viewmatrix; a.proj = prm->projmatrix; a.campos = prm->campos;
  a.mod = prm->scale_modifier; a.tanx = prm->tanfovx; a.tany = prm->tanfovy;
  a.fx = W / (2.0f * prm->tanfovx);
  a.fy = H / (2.0f * prm->tanfovy);
  a.prefiltered = prm->prefiltered;
  {
    Phase ph("preprocess_fwd", st);
    if (surfel) {
      RSPreArgs b;
      b.P = P; b.D = a.D; b.M = a.M; b.W = W; b.H = H; b.gx = gx; b.gy = gy;
      b.means = a.means; b.scales = a.scales; b.rots = a.rots; b.opac = a.opac; b.shs = a.shs;
      b.cov3D_pre = a.cov3D_pre; b.colors_pre = a.colors_pre; b.view = a.view; b.proj = a.proj; b.campos = a.campos;
      b.patch_bbox = prm->patch_bbox; b.config = prm->config;
      b.mod = a.mod; b.tanx = a.tanx; b.tany = a.tany; b.fx = a.fx; b.fy = a.fy; b.prefiltered = a.prefiltered;
      rs_preprocess_kernel<<<div_up(P, 256), 256, 0, st>>>(b, radii, ws.rec, ws.ext, ws.zyv, ws.dkey, ws.tiles,
                                                          ws.wts, ws.rect, ws.clamped, ws.total);
    } else {
      r3_preprocess_kernel<<<div_up(P, 256), 256, 0, st>>>(a, radii, ws.rec, ws.dkey, ws.tiles, ws.wts, ws.rect,
                                                          ws.clamped, ws.total);
    }
  }
  PG_LAUNCH_CHECK(prm->debug, st);
  PG_CUDA(cudaMemcpyAsync(g_slot.pinned, ws.total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  PG_CUDA(cudaEventRecord(g_slot.ev, st));
  // depth ordering runs while the host waits for the count
  {
    Phase ph("depth_sort", st);
    PG_CUDA(depth_order_and_offsets(ws, P, bin_fast_path(gx * gy), st, /*need_offsets=*/!band_path(gx, gy)));
  }
  PG_LAUNCH_CHECK(prm->debug, st);
  if (!wait) return PG_OK;
  PG_CUDA(cudaEventSynchronize(g_slot.ev));
  const unsigned long long R = *g_slot.pinned;
  if (R > 0xfffffff0ull) return PG_ERR_UNSUPPORTED;
  *num_rendered = (int64_t)R;
  return PG_OK;
}

int pg_raster_forward_stage1(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, size_t geom_bytes,
                             int32_t* radii, int64_t* num_rendered, void* stream) {
  return stage1_impl(prm, g, geom_ws, geom_bytes, radii, num_rendered, stream, true);
}

// num_rendered: what the binning workspace was carved for; cap: ~0 when num_rendered is the true count, else the
// capacity the device-side count is checked against (optimistic launch).
static int stage2_impl(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, void* binning_ws,
                       size_t binning_bytes, void* image_ws, int64_t num_rendered, unsigned long long cap,
                       const PgRasterFwdOut* out, void* stream) {
  if (!params_ok(prm) || !g || !out || num_rendered < 0) return PG_ERR_ARG;
  const bool surfel = prm->variant == PG_VARIANT_SURFEL;
  if (!image_ws || !out->out_color || !out->out_depth || !out->out_opacity) return PG_ERR_ARG;
  if (surfel && !out->out_normal) return PG_ERR_ARG;
  const int P = prm->P, W = prm->image_width, H = prm->image_height;
  const int gx = div_up(W, kTile), gy = div_up(H, kTile);
  if ((int64_t)gx * gy > 65535) return PG_ERR_UNSUPPORTED;
  if (P > 0 && (!geom_ws || (surfel ? !out->contributions : !out->n_touched))) return PG_ERR_ARG;
  if (num_rendered > 0 && (!binning_ws || binning_bytes < carve_binning(num_rendered, W, H, nullptr, nullptr)))
    return PG_ERR_WORKSPACE;
  cudaStream_t st = (cudaStream_t)stream;
  GeomWs gw;
  BinWs bw;
  ImageWs iw;
  carve_geom(prm->variant, P, geom_ws, &gw);
  carve_binning(num_rendered, W, H, binning_ws, &bw);
  carve_image(prm->variant, W, H, image_ws, &iw);
  const uint32_t* point_list = nullptr;
  if (band_path(gx, gy) && num_rendered > 0 && P > 0) {
    PG_CUDA(bin_instances(gw, bw, iw, P, W, H, num_rendered, &point_list, st, cap));   // times each kernel itself
  } else {
    Phase ph("binning", st);
    PG_CUDA(bin_instances(gw, bw, iw, P, W, H, num_rendered, &point_list, st));
  }
  PG_LAUNCH_CHECK(prm->debug, st);
  {
    Phase ph("render_fwd", st);
    // launched as the programmatic dependent of the last binning kernel (a plain launch when there was none)
    if (surfel)
      PG_CUDA(launch_pdl(rs_render_fwd_kernel, (unsigned)(gx * gy), (unsigned)kRThreads, 0, st, (const uint2*)iw.ranges,
                         point_list, W, H, gx, (const float4*)gw.rec, (const float4*)gw.ext, (const float*)gw.zyv,
                         (const float*)prm->bg, (const float*)prm->config, out->out_color, out->out_normal,
                         out->out_depth, out->out_opacity, iw.final_T, iw.final_D, iw.n_contrib, out->contributions));
    else
      PG_CUDA(launch_pdl(r3_render_fwd_kernel, (unsigned)(gx * gy), (unsigned)kRThreads, 0, st, (const uint2*)iw.ranges,
                         point_list, W, H, gx, (const float4*)gw.rec, (const float*)prm->bg, out->out_color,
                         out->out_depth, out->out_opacity, iw.final_T, iw.n_contrib, out->n_touched));
  }
  PG_LAUNCH_CHECK(prm->debug, st);
  return PG_OK;
}

int pg_raster_forward_stage2(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, void* binning_ws,
                             size_t binning_bytes, void* image_ws, int64_t num_rendered,
                             const PgRasterFwdOut* out, void* stream) {
  return stage2_impl(prm, g, geom_ws, binning_ws, binning_bytes, image_ws, num_rendered, ~0ull, out, stream);
}

int pg_raster_forward_optimistic(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, size_t geom_bytes,
                                 int32_t* radii, void* binning_ws, size_t binning_bytes, int64_t capacity,
                                 void* image_ws, const PgRasterFwdOut* out, int64_t* num_rendered, void* stream) {
  if (!params_ok(prm) || !num_rendered || capacity <= 0 || capacity > 0xfffffff0ll) return PG_ERR_ARG;
  if (prm->P == 0) return PG_ERR_UNSUPPORTED;
  if (!band_path(div_up(prm->image_width, kTile), div_up(prm->image_height, kTile))) return PG_ERR_UNSUPPORTED;
  int rc = stage1_impl(prm, g, geom_ws, geom_bytes, radii, num_rendered, stream, false);
  if (rc != PG_OK) return rc;
  rc = stage2_impl(prm, g, geom_ws, binning_ws, binning_bytes, image_ws, capacity, (unsigned long long)capacity, out, stream);
  if (rc != PG_OK) return rc;
  // the ONE host wait of the forward, after everything is queued
  PG_CUDA(cudaEventSynchronize(g_slot.ev));
  const unsigned long long R = *g_slot.pinned;
  if (R > 0xfffffff0ull) return PG_ERR_UNSUPPORTED;
  *num_rendered = (int64_t)R;
  return R > (unsigned long long)capacity ? PG_ERR_WORKSPACE : PG_OK;
}

int pg_raster_backward(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, void* binning_ws,
                       void* image_ws, int64_t num_rendered, const PgRasterBwdIn* in, const PgRasterBwdOut* out,
                       void* stream) {
  if (!params_ok(prm) || !g || !in || !out || num_rendered < 0) return PG_ERR_ARG;
  const bool surfel = prm->variant == PG_VARIANT_SURFEL;
  if (surfel && (!in->dL_dout_normal || !in->dL_dout_opac)) return PG_ERR_ARG;
  const int P = prm->P, W = prm->image_width, H = prm->image_height;
  if (P == 0) {
    if (out->dL_dtau) PG_CUDA(cudaMemsetAsync(out->dL_dtau, 0, 6 * sizeof(float), (cudaStream_t)stream));
    return PG_OK;
  }
  if (!geom_ws || !image_ws || !in->dL_dout_color || !in->dL_dout_depth || !in->radii || !prm->projmatrix_raw)
    return PG_ERR_ARG;
  if (!out->dL_dmeans2D || !out->dL_dcolors || !out->dL_dopacity || !out->dL_dmeans3D || !out->dL_dcov3D ||
      !out->dL_dscales || !out->dL_drotations || !out->dL_dtau)
    return PG_ERR_ARG;
  if (g->shs && !out->dL_dsh) return PG_ERR_ARG;
  const int gx = div_up(W, kTile), gy = div_up(H, kTile);
  cudaStream_t st = (cudaStream_t)stream;
  GeomWs gw;
  BinWs bw;
  ImageWs iw;
  carve_geom(prm->variant, P, geom_ws, &gw);
  carve_binning(num_rendered, W, H, binning_ws, &bw);
  carve_image(prm->variant, W, H, image_ws, &iw);
  PG_CUDA(cudaMemsetAsync(gw.gacc, 0, sizeof(float4) * gw.gacc_count, st));
  PG_CUDA(cudaMemsetAsync(gw.tau_acc, 0, sizeof(double) * 8, st));
  if (num_rendered > 0) {
    const uint32_t* point_list = sorted_point_list(bw, W, H);
    Phase ph("render_bwd", st);
    if (surfel) {
      // 2 warps x quads by default (fewest per-surfel reductions; 882 vs 971 us at C3 once the accum_rec state
      // became two scalars per pixel); PG_RS_BWD_WARPS=4 selects 4 warps x pixel pairs (dev A/B knob)
      static const int nw = [] { const char* e = getenv("PG_RS_BWD_WARPS"); return (e && atoi(e) == 4) ? 4 : 2; }();
      if (nw == 4)
        rs_render_bwd_kernel<4><<<gx * gy, 128, 0, st>>>(iw.ranges, point_list, W, H, gx, gw.rec, gw.ext, gw.zyv,
                                                         prm->bg, prm->config, iw.final_T, iw.final_D, iw.n_contrib,
                                                         in->dL_dout_color, in->dL_dout_normal, in->dL_dout_depth,
                                                         in->dL_dout_opac, gw.gacc);
      else
        rs_render_bwd_kernel<2><<<gx * gy, 64, 0, st>>>(iw.ranges, point_list, W, H, gx, gw.rec, gw.ext, gw.zyv,
                                                        prm->bg, prm->config, iw.final_T, iw.final_D, iw.n_contrib,
                                                        in->dL_dout_color, in->dL_dout_normal, in->dL_dout_depth,
                                                        in->dL_dout_opac, gw.gacc);
    } else
      r3_render_bwd_kernel<<<gx * gy, kRThreads, 0, st>>>(iw.ranges, point_list, W, H, gx, gw.rec, prm->bg,
                                                          iw.final_T, iw.n_contrib, in->dL_dout_color,
                                                          in->dL_dout_depth, gw.gacc);
    PG_LAUNCH_CHECK(prm->debug, st);
  }
  R3BwdPreArgs a;
  a.P = P; a.D = prm->sh_degree; a.M = prm->sh_coeffs; a.W = W; a.H = H;
  a.means = g->means3D; a.scales = g->scales; a.rots = g->rotations; a.shs = g->shs; a.cov3D_pre = g->cov3D_precomp;
  a.view = prm->viewmatrix; a.proj = prm->projmatrix; a.proj_raw = prm->projmatrix_raw; a.campos = prm->campos;
  a.radii = in->radii; a.clamped = gw.clamped; a.gacc = gw.gacc;
  a.mod = prm->scale_modifier; a.tanx = prm->tanfovx; a.tany = prm->tanfovy;
  a.fx = W / (2.0f * prm->tanfovx);
  a.fy = H / (2.0f * prm->tanfovy);
  a.dL_dmeans2D = out->dL_dmeans2D; a.dL_dcolors = out->dL_dcolors; a.dL_dopacity = out->dL_dopacity;
  a.dL_dmeans3D = out->dL_dmeans3D; a.dL_dcov3D = out->dL_dcov3D; a.dL_dsh = g->shs ? out->dL_dsh : nullptr;
  a.dL_dscales = out->dL_dscales; a.dL_drots = out->dL_drotations;
  a.dL_dtau_full = out->dL_dtau_full; a.tau_acc = gw.tau_acc; a.dL_dtau = out->dL_dtau;
  a.ext = gw.ext; a.zyv = gw.zyv; a.config = prm->config;
  {
    Phase ph("preprocess_bwd", st);
    if (surfel) r3_preprocess_bwd_kernel<true><<<div_up(P, 256), 256, 0, st>>>(a);
    else r3_preprocess_bwd_kernel<false><<<div_up(P, 256), 256, 0, st>>>(a);
  }
  PG_LAUNCH_CHECK(prm->debug, st);
  return PG_OK;
}

int pg_raster_debug_views(int variant, int P, int width, int height, int64_t num_rendered, void* geom_ws,
                          void* binning_ws, void* image_ws, const uint32_t** point_list, const uint32_t** ranges,
                          const float** final_T, const uint32_t** n_contrib, const float** records,
                          const uint32_t** depth_sorted_idx) {
  GeomWs gw;
  BinWs bw;
  ImageWs iw;
  carve_geom(variant, P, geom_ws, &gw);
  carve_binning(num_rendered, width, height, binning_ws, &bw);
  carve_image(variant, width, height, image_ws, &iw);
  if (point_list) *point_list = sorted_point_list(bw, width, height);
  if (ranges) *ranges = reinterpret_cast<const uint32_t*>(iw.ranges);
  if (final_T) *final_T = iw.final_T;
  if (n_contrib) *n_contrib = iw.n_contrib;
  if (records) *records = reinterpret_cast<const float*>(gw.rec);
  if (depth_sorted_idx) *depth_sorted_idx = gw.sorted_idx;
  return PG_OK;
}

int pg_raster_debug_gacc(int variant, int P, void* geom_ws, const float** gacc, int* stride_float4) {
  if (!geom_ws || !gacc || P < 0) return PG_ERR_ARG;
  GeomWs gw;
  carve_geom(variant, P, geom_ws, &gw);
  *gacc = reinterpret_cast<const float*>(gw.gacc);
  if (stride_float4) *stride_float4 = variant == PG_VARIANT_SURFEL ? 4 : 3;
  return PG_OK;
}

int64_t pg_raster_band_records(int width, int height, int64_t num_rendered, void* binning_ws, void* stream) {
  const int gx = div_up(width, kTile), gy = div_up(height, kTile);
  if (!binning_ws || num_rendered <= 0 || !band_path(gx, gy)) return -1;
  BinWs bw;
  carve_binning(num_rendered, width, height, binning_ws, &bw);
  uint32_t q = 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (cudaMemcpyAsync(&q, bw.band.lstart + band_lists(gx, gy), sizeof(q), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
      cudaStreamSynchronize(st) != cudaSuccess)
    return -1;
  return (int64_t)q;
}

}  // extern "C"
