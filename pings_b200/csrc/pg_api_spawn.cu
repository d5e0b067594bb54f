// C-ABI entry points of the fused spawn_gaussians epilogue (see include/pings_b200.h).
#include "pg_spawn.cuh"

using namespace pg;

extern "C" {

int pg_spawn_prepare(int64_t n, const int64_t* sidx, const float* position, const float* orientation,
                     const float* cam_origin, const float* geo_feature, int Fg, const float* color_feature, int Fc,
                     int dist_concat, int view_concat, int xy_only, float* geo_in, float* color_in, void* stream) {
  if (n < 0 || Fg <= 0 || !geo_feature || !geo_in) return PG_ERR_ARG;
  if (color_in && (!color_feature || Fc <= 0)) return PG_ERR_ARG;
  if ((dist_concat || (view_concat && color_in)) && cam_origin && (!position || !orientation)) return PG_ERR_ARG;
  if (n == 0) return PG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  spawn_prepare_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(n, sidx, position, orientation, cam_origin, geo_feature, Fg,
                                                                color_feature, Fc, dist_concat, view_concat, xy_only, geo_in,
                                                                color_in);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_spawn_prepare_backward(int64_t n, const int64_t* sidx, const float* d_in, int in_w, int F, float* d_feature,
                              void* stream) {
  if (n < 0 || !d_in || !d_feature || F <= 0 || in_w < F) return PG_ERR_ARG;
  if (n == 0) return PG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  spawn_prepare_bwd_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(n, sidx, d_in, in_w, F, d_feature);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

size_t pg_spawn_state_words(int64_t NK) { return (size_t)((NK + 255) / 256) + 4; }
size_t pg_spawn_count_index(int64_t NK) { return (size_t)((NK + 255) / 256) + 1; }

static bool spawn_args_ok(const PgSpawnArgs* a) {
  return a && a->n >= 0 && a->K > 0 && a->position && a->orientation && a->xyz_raw && a->rot_raw && a->scale_raw &&
         a->alpha_raw && a->color_raw && a->n * a->K < (1ll << 30);
}

int pg_spawn_epilogue_forward(const PgSpawnArgs* a, const PgSpawnOut* o, void* stream) {
  if (!spawn_args_ok(a) || !o || !o->xyz || !o->scale || !o->rot || !o->alpha || !o->color || !o->alpha_all || !o->cpos ||
      !o->state)
    return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t NK = a->n * a->K;
  const uint32_t nblocks = (uint32_t)((NK + 255) / 256);
  PG_CUDA(cudaMemsetAsync(o->state, 0, pg_spawn_state_words(NK) * sizeof(uint32_t), st));
  if (NK == 0) return PG_OK;
  spawn_epilogue_kernel<<<nblocks, 256, 0, st>>>(*a, *o, nblocks);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_spawn_epilogue_backward(const PgSpawnArgs* a, const PgSpawnGrads* g, void* stream) {
  if (!spawn_args_ok(a) || !g || !g->cpos || !g->d_xyz_raw || !g->d_rot_raw || !g->d_scale_raw || !g->d_alpha_raw ||
      !g->d_color_raw)
    return PG_ERR_ARG;
  const int64_t NK = a->n * a->K;
  if (NK == 0) return PG_OK;
  cudaStream_t st = (cudaStream_t)stream;
  spawn_epilogue_bwd_kernel<<<(unsigned)((NK + 255) / 256), 256, 0, st>>>(*a, *g);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

}  // extern "C"
