// TEST INFRASTRUCTURE — not part of the product.
//
// Thin C-ABI shim around the UNMODIFIED reference Gaussian-surfel rasterizer
// (submodules/diff-gaussian-surfel-rasterization-w-pose/cuda_rasterizer/*), compiled from the
// sources where they lie under /root/reference by oracle/build_ref.py into
// oracle/_ref/libref_rs.so.  Only tests/, __graft_entry__.smoke() and bench.py's comparator
// leg may load it.  This file contains no reference code: it calls
// CudaRasterizer::Rasterizer::{forward,backward,markVisible} through the reference's own headers
// and exposes the internal buffers by re-deriving the chunk layout with its fromChunk() helpers.
#include <cuda_runtime.h>
#include <cstdint>
#include <functional>
#include <vector>
#include "cuda_rasterizer/rasterizer.h"
#include "cuda_rasterizer/rasterizer_impl.h"

namespace {
struct Buf {
  char* p = nullptr;
  size_t cap = 0;
  char* get(size_t n) {
    if (n > cap) {
      if (p) cudaFree(p);
      cudaMalloc(&p, n);
      cudaMemset(p, 0, n);
      cap = n;
    }
    return p;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};
struct RefState {
  Buf geom, bin, img;
  int P = 0, W = 0, H = 0, R = 0;
};
}  // namespace

extern "C" {

void* ref_rs_create() { return new RefState(); }
void ref_rs_destroy(void* h) {
  auto* s = static_cast<RefState*>(h);
  s->geom.release();
  s->bin.release();
  s->img.release();
  delete s;
}

int ref_rs_forward(void* h, int P, int D, int M, const float* bg, int W, int H, const float* means3D,
                   const float* shs, const float* colors, const float* opac, const float* scales, float mod,
                   const float* rots, const float* cov3D_pre, const float* view, const float* proj,
                   const float* prcppoint, const float* patchbbox, const float* campos, float tanx, float tany,
                   int prefiltered, float* config, float* out_color, float* out_normal, float* out_depth,
                   float* out_opac, int* radii, float* contributions) {
  auto* s = static_cast<RefState*>(h);
  s->P = P; s->W = W; s->H = H;
  auto gf = [s](size_t n) { return s->geom.get(n); };
  auto bf = [s](size_t n) { return s->bin.get(n); };
  auto imf = [s](size_t n) { return s->img.get(n); };
  s->R = CudaRasterizer::Rasterizer::forward(gf, bf, imf, P, D, M, bg, W, H, means3D, shs, colors, opac, scales,
                                              mod, rots, cov3D_pre, view, proj, prcppoint, patchbbox, campos, tanx,
                                              tany, prefiltered != 0, config, out_color, out_normal, out_depth,
                                              out_opac, radii, contributions, false);
  cudaError_t e = cudaDeviceSynchronize();
  return e == cudaSuccess ? s->R : -1;
}

int ref_rs_views(void* h, const uint32_t** point_list, const uint64_t** keys, const uint32_t** ranges,
                 const uint32_t** n_contrib, const float** final_T, const float** means2D, const float** depths,
                 const float** conic_opacity, const uint32_t** tiles_touched, const float** normal,
                 const float** Jinv) {
  auto* s = static_cast<RefState*>(h);
  char* g = s->geom.p;
  char* b = s->bin.p;
  char* i = s->img.p;
  auto gs = CudaRasterizer::GeometryState::fromChunk(g, s->P);
  auto is = CudaRasterizer::ImageState::fromChunk(i, (size_t)s->W * s->H);
  if (point_list || keys) {
    if (s->R > 0) {
      auto bs = CudaRasterizer::BinningState::fromChunk(b, s->R);
      if (point_list) *point_list = bs.point_list;
      if (keys) *keys = bs.point_list_keys;
    } else {
      if (point_list) *point_list = nullptr;
      if (keys) *keys = nullptr;
    }
  }
  if (ranges) *ranges = reinterpret_cast<const uint32_t*>(is.ranges);
  if (n_contrib) *n_contrib = is.n_contrib;
  if (final_T) *final_T = is.accum_alpha;
  if (means2D) *means2D = reinterpret_cast<const float*>(gs.means2D);
  if (depths) *depths = gs.depths;
  if (conic_opacity) *conic_opacity = reinterpret_cast<const float*>(gs.conic_opacity);
  if (tiles_touched) *tiles_touched = gs.tiles_touched;
  if (normal) *normal = gs.normal;
  if (Jinv) *Jinv = gs.Jinv;
  return 0;
}

int ref_rs_backward(void* h, int D, int M, const float* bg, const float* means3D, const float* shs,
                    const float* colors, const float* scales, float mod, const float* rots,
                    const float* cov3D_pre, const float* view, const float* proj, const float* proj_raw,
                    const float* campos, const float* prcppoint, const float* patchbbox, float tanx, float tany,
                    const int* radii, const float* dL_dpix, const float* dL_dpix_normal,
                    const float* dL_dpix_depth, const float* dL_dpix_opac, float* dL_dmean2D, float* dL_dconic,
                    float* dL_dopacity, float* dL_dcolor, float* dL_dnormal, float* dL_ddepth, float* dL_dmean3D,
                    float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dtau,
                    float* config) {
  auto* s = static_cast<RefState*>(h);
  CudaRasterizer::Rasterizer::backward(s->P, D, M, s->R, bg, s->W, s->H, means3D, shs, colors, scales, mod, rots,
                                       cov3D_pre, view, proj, proj_raw, campos, prcppoint, patchbbox, tanx, tany,
                                       radii, s->geom.p, s->bin.p, s->img.p, dL_dpix, dL_dpix_normal, dL_dpix_depth,
                                       dL_dpix_opac, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_dnormal,
                                       dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dtau, false,
                                       config);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

int ref_rs_copy(void* dst, const void* src, size_t bytes) {
  return cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToDevice) == cudaSuccess ? 0 : -1;
}

int ref_rs_mark_visible(int P, float* means3D, float* view, float* proj, bool* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : -1;
}

}  // extern "C"
