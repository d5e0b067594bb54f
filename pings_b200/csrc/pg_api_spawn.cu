// C-ABI entry points of the fused spawn_gaussians epilogue (see include/pings_b200.h).
#include <mutex>
#include "pg_spawn.cuh"

using namespace pg;

// ---- scratch of pg_spawn_epilogue_forward_static (no allocation anywhere on that path) -----------------------------
// The look-back state words of spawn_epilogue_kernel for up to PG_SPAWN_STATIC_MAX_NK candidate Gaussians: 65,540 words
// = 256 KiB of device memory, held by every GPU that loads the library until the process exits.  In HOST code the
// name below is only a stand-in for the symbol, NOT a device address: the address comes from cudaGetSymbolAddress,
// looked up once per device with that device current.
constexpr size_t kSpawnStaticWords = (size_t)(PG_SPAWN_STATIC_MAX_NK / 256) + 4;
static __device__ __align__(256) uint32_t g_spawn_static_state[kSpawnStaticWords];

static int spawn_static_state(uint32_t** out) {
  constexpr int kMaxDev = 64;
  static uint32_t* ptrs[kMaxDev] = {};
  static std::mutex mu;
  int dev = -1;
  PG_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDev) {
    snprintf(g_last_error, sizeof(g_last_error), "pg_spawn_epilogue_forward_static: device ordinal %d out of range", dev);
    return PG_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    void* p = nullptr;
    PG_CUDA(cudaGetSymbolAddress(&p, g_spawn_static_state));
    ptrs[dev] = static_cast<uint32_t*>(p);
  }
  *out = ptrs[dev];
  return PG_OK;
}

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

int pg_spawn_epilogue_forward_static(const PgSpawnArgs* a, const PgSpawnOut* o, uint32_t* count_nan_out, void* stream) {
  if (!spawn_args_ok(a) || !o || !o->xyz || !o->scale || !o->rot || !o->alpha || !o->color || !o->alpha_all || !o->cpos ||
      !count_nan_out)
    return PG_ERR_ARG;
  const int64_t NK = a->n * a->K;
  if (NK > (int64_t)PG_SPAWN_STATIC_MAX_NK) {
    snprintf(g_last_error, sizeof(g_last_error),
             "pg_spawn_epilogue_forward_static: n*K = %lld candidate Gaussians exceed the static scratch limit "
             "PG_SPAWN_STATIC_MAX_NK = %lld; use pg_spawn_epilogue_forward with caller-provided state",
             (long long)NK, (long long)PG_SPAWN_STATIC_MAX_NK);
    return PG_ERR_UNSUPPORTED;
  }
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* state = nullptr;
  const int rc = spawn_static_state(&state);
  if (rc != PG_OK) return rc;
  // the array keeps the previous call's words: the look-back status words, the ticket, the count and the flag of
  // THIS call's extent must start at zero
  PG_CUDA(cudaMemsetAsync(state, 0, pg_spawn_state_words(NK) * sizeof(uint32_t), st));
  if (NK > 0) {
    PgSpawnOut oo = *o;
    oo.state = state;
    const uint32_t nblocks = (uint32_t)((NK + 255) / 256);
    spawn_epilogue_kernel<<<nblocks, 256, 0, st>>>(*a, oo, nblocks);
    PG_LAUNCH_CHECK(0, st);
  }
  PG_CUDA(cudaMemcpyAsync(count_nan_out, state + pg_spawn_count_index(NK), 2 * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
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
