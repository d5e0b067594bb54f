// C-ABI entry points of the neural-point path (see include/pings_b200.h).
#include <cstdlib>
#include "pg_sdf.cuh"
#include "pg_mlp_tc.cuh"

using namespace pg;

namespace {

bool map_ok(const PgNeuralMap* m, bool tf) {
  if (!m || m->buffer_size <= 0 || !m->buffer_pt_index || !m->neural_points || m->n_points <= 0) return false;
  if (!m->neighbor_dx || m->K <= 0 || m->K > 256 || !(m->resolution > 0.f)) return false;
  if (tf && (!m->point_ts_create || !m->travel_dist)) return false;
  return true;
}

NpMapDev to_dev(const PgNeuralMap* m, int time_filtering) {
  NpMapDev d;
  d.buffer_size = m->buffer_size;
  {
    const unsigned long long bs = (unsigned long long)m->buffer_size;
    // floor(2^64 / bs); for a power of two (2^64 - 1) / bs is one short, which the correction step absorbs
    d.bs_magic = bs ? ~0ull / bs : 0ull;
  }
  d.table = reinterpret_cast<const long long*>(m->buffer_pt_index);
  d.table32 = m->buffer_pt_index32;
  d.pts = m->neural_points;
  d.n_points = m->n_points;
  d.dx = reinterpret_cast<const long long*>(m->neighbor_dx);
  d.K = m->K;
  d.p0 = m->primes[0]; d.p1 = m->primes[1]; d.p2 = m->primes[2];
  d.res = m->resolution;
  d.max_valid = m->max_valid_dist2;
  d.ts_create = m->point_ts_create;
  d.travel = m->travel_dist;
  d.cur_ts = m->cur_ts;
  d.diff_local = m->diff_travel_dist_local;
  d.time_filtering = time_filtering;
  return d;
}

bool query_ok(const PgQueryArgs* a) {
  if (!a || a->nn_k <= 0 || a->nn_k > kNpMaxK) return false;
  if (a->F_geo < 0 || a->F_color < 0 || (a->F_geo == 0 && a->F_color == 0)) return false;
  if (!a->points || (a->F_geo && !a->geo_features) || (a->F_color && !a->color_features)) return false;
  if (a->use_only_measured && !a->free_gs_mask) return false;
  if (a->use_only_valid && !a->valid_gs_mask) return false;
  if (a->query_locally && !a->global2local) return false;
  if (a->after_pgo && !a->orientations) return false;
  return true;
}

NpQueryDev to_dev(const PgQueryArgs* a) {
  NpQueryDev d;
  d.nn_k = a->nn_k; d.query_locally = a->query_locally; d.use_only_measured = a->use_only_measured;
  d.use_only_valid = a->use_only_valid; d.weighted_first = a->weighted_first; d.after_pgo = a->after_pgo;
  d.F_geo = a->F_geo; d.F_color = a->F_color;
  d.free_mask = a->free_gs_mask; d.valid_mask = a->valid_gs_mask;
  d.g2l = reinterpret_cast<const long long*>(a->global2local);
  d.points = a->points; d.quats = a->orientations; d.cert = a->certainties;
  d.geo = a->geo_features; d.color = a->color_features;
  return d;
}

bool mlp_ok(const PgMlp* m) {
  if (!m || m->in_dim <= 0 || m->hidden_dim <= 0 || m->out_dim <= 0) return false;
  if (m->hidden_level < 1 || m->hidden_level + 1 > PG_MLP_MAX_LAYERS) return false;
  for (int l = 0; l <= m->hidden_level; ++l)
    if (!m->weights[l]) return false;
  return true;
}

MlpDev to_dev(const PgMlp* m) {
  MlpDev d;
  d.in_dim = m->in_dim; d.hidden = m->hidden_dim; d.levels = m->hidden_level; d.out_dim = m->out_dim;
  d.leaky = m->leaky_relu;
  for (int l = 0; l < PG_MLP_MAX_LAYERS; ++l) {
    d.W[l] = l <= m->hidden_level ? m->weights[l] : nullptr;
    d.B[l] = l <= m->hidden_level ? m->biases[l] : nullptr;
  }
  return d;
}

constexpr size_t kMaxDynSmem = 227 * 1024;
constexpr int64_t kTcAutoMinRows = 4096;

template <typename F>
int with_cpl(int K, F&& f) {
  if (K <= 32) return f(std::integral_constant<int, 1>{});
  if (K <= 96) return f(std::integral_constant<int, 3>{});
  if (K <= 128) return f(std::integral_constant<int, 4>{});
  return f(std::integral_constant<int, 8>{});
}

}  // namespace

__global__ void __launch_bounds__(256) np_slot_mirror_kernel(long long n, const long long* __restrict__ t64,
                                                             int* __restrict__ t32) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) t32[i] = (int)t64[i];
}

extern "C" {

int pg_np_build_slot_mirror(int64_t buffer_size, const int64_t* table64, int32_t* table32, void* stream) {
  if (buffer_size <= 0 || !table64 || !table32) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  np_slot_mirror_kernel<<<kSmCount * 16, 256, 0, st>>>(buffer_size, reinterpret_cast<const long long*>(table64), table32);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_np_radius_search(const PgNeuralMap* map, int64_t N, const float* q, int time_filtering, float* out_dist2,
                        int64_t* out_idx, void* stream) {
  if (N < 0 || !map_ok(map, time_filtering != 0)) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!q || !out_dist2 || !out_idx) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)N * map->K;
  const long long blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return PG_ERR_UNSUPPORTED;
  Phase ph("np_radius_search", st);
  np_radius_search_kernel<<<(unsigned)blocks, 256, 0, st>>>(to_dev(map, time_filtering), N, q, out_dist2,
                                                          reinterpret_cast<long long*>(out_idx));
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_np_query_feature(const PgNeuralMap* map, const PgQueryArgs* args, int64_t N, const float* q,
                        const PgQueryOut* out, void* stream) {
  if (N < 0 || !query_ok(args) || !map_ok(map, args->time_filtering != 0) || !out) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!q || !out->weight || !out->nn_counts) return PG_ERR_ARG;
  if (out->certainty && !args->certainties) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const long long blocks = (N + kNpWarps - 1) / kNpWarps;
  if (blocks > 0x7fffffffLL) return PG_ERR_UNSUPPORTED;
  const NpMapDev m = to_dev(map, args->time_filtering);
  const NpQueryDev a = to_dev(args);
  float* geo = args->F_geo ? out->geo_vec : nullptr;
  float* col = args->F_color ? out->color_vec : nullptr;
  Phase ph("np_query_feature", st);
  return with_cpl(map->K, [&](auto cpl) {
    np_query_kernel<decltype(cpl)::value><<<(unsigned)blocks, kNpWarps * 32, 0, st>>>(
        m, a, N, q, geo, col, out->weight, reinterpret_cast<long long*>(out->nn_counts), out->certainty, out->knn_idx,
        out->knn_dist2);
    PG_LAUNCH_CHECK(0, st);
    return PG_OK;
  });
}

int pg_np_accumulate_stability(int64_t N, int32_t nn_k, const int32_t* knn_idx, const float* weight,
                               const int32_t* query_ts, float* certainties, int32_t* ts_update, void* stream) {
  if (N < 0 || nn_k <= 0) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!knn_idx || !weight || !certainties) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = (long long)N * nn_k;
  Phase ph("np_accumulate", st);
  np_accumulate_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(total, nn_k, knn_idx, weight, query_ts,
                                                                      certainties, ts_update);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_np_query_feature_backward(const PgQueryArgs* args, int64_t N, const float* q, const int32_t* knn_idx,
                                 const float* knn_dist2, const float* weight, const int64_t* nn_counts,
                                 const float* d_geo_vec, const float* d_color_vec, const float* d_weight,
                                 float* d_geo_features, float* d_color_features, float* d_query, void* stream) {
  if (N < 0 || !query_ok(args)) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!q || !knn_idx || !knn_dist2 || !weight || !nn_counts) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const long long blocks = (N + kNpWarps - 1) / kNpWarps;
  Phase ph("np_query_feature_bwd", st);
  np_query_bwd_kernel<<<(unsigned)blocks, kNpWarps * 32, 0, st>>>(
      to_dev(args), N, q, knn_idx, knn_dist2, weight, reinterpret_cast<const long long*>(nn_counts),
      args->F_geo ? d_geo_vec : nullptr, args->F_color ? d_color_vec : nullptr, d_weight, d_geo_features,
      d_color_features, d_query);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_mlp_forward_impl(const PgMlp* mlp, int64_t N, const float* x, float* y, float* hidden_save, int impl,
                        void* stream) {
  if (N < 0 || !mlp_ok(mlp) || impl < PG_MLP_AUTO || impl > PG_MLP_TCGEN05) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!x || !y) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const MlpDev m = to_dev(mlp);
  const bool tc_ok = tc_supported(m);
  if (impl == PG_MLP_TCGEN05 && !tc_ok) return PG_ERR_UNSUPPORTED;
  // large batches (mlp_batch, spawn_gaussians) go to the tensor cores; small ones are latency bound
  if (impl == PG_MLP_TCGEN05 || (impl == PG_MLP_AUTO && tc_ok && N >= kTcAutoMinRows)) {
    Phase ph("mlp_forward_tc", st);
    PG_CUDA(mlp_tc_forward(m, N, x, y, hidden_save, st));
    PG_LAUNCH_CHECK(0, st);
    return PG_OK;
  }
  const size_t smem = mlp_smem_floats(m) * sizeof(float);
  if (smem > kMaxDynSmem) return PG_ERR_UNSUPPORTED;
  PG_CUDA(cudaFuncSetAttribute(mlp_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  Phase ph("mlp_forward", st);
  mlp_forward_kernel<<<(unsigned)((N + kMlpRows - 1) / kMlpRows), kMlpThreads, smem, st>>>(m, N, x, y, hidden_save);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_mlp_forward(const PgMlp* mlp, int64_t N, const float* x, float* y, float* hidden_save, void* stream) {
  return pg_mlp_forward_impl(mlp, N, x, y, hidden_save, PG_MLP_AUTO, stream);
}

int pg_mlp_backward(const PgMlp* mlp, int64_t N, const float* x, const float* hidden_save, const float* dY,
                    float* dX, float* const* dW, float* const* dB, void* stream) {
  if (N < 0 || !mlp_ok(mlp) || !dW) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!x || !hidden_save || !dY) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const MlpDev m = to_dev(mlp);
  MlpGradDev g;
  for (int l = 0; l < PG_MLP_MAX_LAYERS; ++l) {
    g.dW[l] = l <= m.levels ? dW[l] : nullptr;
    g.dB[l] = (dB && l <= m.levels) ? dB[l] : nullptr;
    if (l <= m.levels && !g.dW[l]) return PG_ERR_ARG;
  }
  const size_t smem = mlp_bwd_smem_floats(m) * sizeof(float);
  if (smem > kMaxDynSmem) return PG_ERR_UNSUPPORTED;
  PG_CUDA(cudaFuncSetAttribute(mlp_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  Phase ph("mlp_backward", st);
  mlp_backward_kernel<<<(unsigned)((N + kMlpRows - 1) / kMlpRows), kMlpThreads, smem, st>>>(m, g, N, x, hidden_save, dY,
                                                                                         dX);
  PG_LAUNCH_CHECK(0, st);
  return PG_OK;
}

int pg_sdf_query(const PgNeuralMap* map, const PgQueryArgs* args, const PgMlp* mlp, float sdf_scale, int64_t N,
                 const float* q, float* sdf, float* sdf_std, int64_t* nn_counts, void* stream) {
  if (N < 0 || !query_ok(args) || !map_ok(map, args->time_filtering != 0) || !mlp_ok(mlp)) return PG_ERR_ARG;
  if (args->F_geo <= 0 || mlp->in_dim != args->F_geo + 3 || mlp->out_dim != 1) return PG_ERR_ARG;
  if (N == 0) return PG_OK;
  if (!q || !sdf) return PG_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  const NpMapDev m = to_dev(map, args->time_filtering);
  const NpQueryDev a = to_dev(args);
  const MlpDev md = to_dev(mlp);
  const size_t smem = sdf_smem_floats(md) * sizeof(float);
  if (smem > kMaxDynSmem) return PG_ERR_UNSUPPORTED;
  // Measured on B200 (C1, 115,200 queries): warp-per-query 0.227 ms, batched 0.256 ms - the batched kernel's 16
  // resident warps per SM (128 registers) hide the probe latency worse than its cheaper decoder saves, so the
  // warp-per-query kernel stays the default; PG_SDF_IMPL=batch selects the batched one (dev A/B knob).
  static const int use_batch = [] { const char* e = getenv("PG_SDF_IMPL"); return (e && e[0] == 'b') ? 1 : 0; }();
  const size_t smem_b = sdf_batch_smem_floats(md, args->nn_k) * sizeof(float);
  if (use_batch && sdf_batch_supported(md) && N >= 4096 && smem_b <= kMaxDynSmem) {
    const long long nbatch = (N + 31) / 32;
    long long blocks = (nbatch + kSdfWarps - 1) / kSdfWarps;
    const long long cap = (long long)kSmCount * 4;
    if (blocks > cap) blocks = cap;
    Phase ph("sdf_query", st);
    return with_cpl(map->K, [&](auto cpl) {
      auto launch = [&](auto kern) {
        PG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
        kern<<<(unsigned)blocks, kSdfWarps * 32, smem_b, st>>>(m, a, md, sdf_scale, N, q, sdf,
                                                               args->weighted_first ? nullptr : sdf_std,
                                                               reinterpret_cast<long long*>(nn_counts));
        PG_LAUNCH_CHECK(0, st);
        return PG_OK;
      };
      if (md.hidden == 64) return launch(sdf_query_batch_kernel<decltype(cpl)::value, 64>);
      return launch(sdf_query_batch_kernel<decltype(cpl)::value, 32>);
    });
  }
  // PG_SDF_IMPL=warp keeps the warp-per-query kernel for shapes the half-warp kernel also covers (A/B, parity tests)
  static const int force_warp = [] { const char* e = getenv("PG_SDF_IMPL"); return (e && e[0] == 'w') ? 1 : 0; }();
  if (!force_warp && sdf_half_supported(md, a, map->K) && map->n_points < (1ll << 31)) {
    // persistent grid = what is resident (a partial second wave of persistent CTAs would double the time)
    static int per_sm = 0;
    if (!per_sm) {
      PG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sdf_query_half_kernel, kSdfWarps * 32, 0));
      if (per_sm < 1) per_sm = 1;
    }
    long long blocks = ((N + 1) / 2 + kSdfWarps - 1) / kSdfWarps;
    const long long cap = (long long)kSmCount * per_sm;
    if (blocks > cap) blocks = cap;
    Phase ph("sdf_query", st);
    sdf_query_half_kernel<<<(unsigned)blocks, kSdfWarps * 32, 0, st>>>(m, a, md, sdf_scale, N, q, sdf,
                                                                       reinterpret_cast<long long*>(nn_counts));
    PG_LAUNCH_CHECK(0, st);
    return PG_OK;
  }
  long long blocks = (N + kSdfWarps - 1) / kSdfWarps;
  const long long cap = (long long)kSmCount * 5;  // persistent CTAs (weights staged once per CTA): 5 resident per SM (48 registers)
  if (blocks > cap) blocks = cap;
  Phase ph("sdf_query", st);
  return with_cpl(map->K, [&](auto cpl) {
    auto kern = sdf_query_kernel<decltype(cpl)::value>;
    PG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)blocks, kSdfWarps * 32, smem, st>>>(m, a, md, sdf_scale, N, q, sdf,
                                                         args->weighted_first ? nullptr : sdf_std,
                                                         reinterpret_cast<long long*>(nn_counts));
    PG_LAUNCH_CHECK(0, st);
    return PG_OK;
  });
}

}  // extern "C"
