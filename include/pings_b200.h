/*
 * pings_b200 C-ABI — the drop-in boundary of the B200-native PINGS hot path.
 *
 * Every entry point takes plain device pointers + sizes + a cudaStream_t (as void*),
 * returns an int status (PG_OK = 0) and never throws.  No torch types cross this line;
 * the Python host (pings_b200/*.py) allocates every buffer as a torch tensor and passes
 * data_ptr()s through ctypes.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   R3 = submodules/diff-gaussian-rasterization-w-pose
 *   RS = submodules/diff-gaussian-surfel-rasterization-w-pose
 *   FS = submodules/fused-ssim
 *   NP = model/neural_gaussians.py, DEC = model/decoder.py
 */
#ifndef PINGS_B200_H_
#define PINGS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PG_OK 0
#define PG_ERR_ARG 1        /* bad argument (null pointer, negative size, bad combination) */
#define PG_ERR_CUDA 2       /* a CUDA runtime call failed; see pg_last_error() */
#define PG_ERR_UNSUPPORTED 3 /* size outside what the kernels support (e.g. > 65535 tiles) */
#define PG_ERR_WORKSPACE 4  /* caller-provided workspace too small */

#define PG_VARIANT_3DGS 0
#define PG_VARIANT_SURFEL 1

/* ABI version; bumped when any struct below changes. */
int pg_abi_version(void);
/* Last CUDA error string recorded by this library on the calling thread ("" if none). */
const char* pg_last_error(void);

/* Opt-in per-phase device timing (no reference counterpart; the reference only has host
 * wall-clock timers, utils/tools.py:551-558).  When enabled on the calling thread every group of
 * launches is bracketed by a CUDA-event pair on the launching stream; pg_timing_drain() waits for
 * them and returns (static) phase names + milliseconds in launch order. */
int pg_timing_enable(int on);
int pg_timing_drain(int max_phases, const char** names, float* ms);

/* ------------------------------------------------------------------------------------------
 * Rasterizer (3DGS: R3/rasterize_points.h:18-39,41-65,67-70; surfel: RS/rasterize_points.cu:36-60,134-162)
 * ---------------------------------------------------------------------------------------- */

/* Camera + image settings: the fields of GaussianRasterizationSettings
 * (R3/diff_gaussian_rasterization/__init__.py:173-186, RS/.../__init__.py:180-196).
 * All pointers are DEVICE pointers to fp32.  The 4x4 matrices are the 16 floats of the
 * torch tensor as stored (i.e. the transposed matrix in row-major = column-major math
 * matrix), exactly what the reference kernels index as m[0],m[4],m[8],m[12]. */
typedef struct PgRasterParams {
  int32_t variant;          /* PG_VARIANT_* */
  int32_t P;                /* number of Gaussians */
  int32_t sh_degree;        /* D */
  int32_t sh_coeffs;        /* M = sh.size(1), 0 when colours are precomputed */
  int32_t image_width, image_height;
  float tanfovx, tanfovy;
  float scale_modifier;
  int32_t prefiltered;
  int32_t debug;            /* 1: synchronise + check after every launch (CHECK_CUDA, R3 auxiliary.h:166-173) */
  const float* bg;              /* [3] */
  const float* viewmatrix;      /* [16] */
  const float* projmatrix;      /* [16] */
  const float* projmatrix_raw;  /* [16]  (backward only) */
  const float* campos;          /* [3]   (SH path only) */
  /* surfel-only (ignored for 3DGS) */
  const float* patch_bbox;      /* [4] */
  const float* prcppoint;       /* [2] */
  const float* config;          /* [5] surface, normalize_depth, perpix_depth, default, front_only */
} PgRasterParams;

/* Per-Gaussian inputs; a NULL pointer means "absent" exactly like the reference's empty
 * tensors (R3 forward.cu:207,243). */
typedef struct PgGaussians {
  const float* means3D;        /* [P,3] */
  const float* shs;            /* [P,M,3] or NULL */
  const float* colors_precomp; /* [P,3] or NULL */
  const float* opacities;      /* [P,1] */
  const float* scales;         /* [P,3] or NULL */
  const float* rotations;      /* [P,4] or NULL */
  const float* cov3D_precomp;  /* [P,6] or NULL */
} PgGaussians;

/* Workspace sizes (bytes).  geom: per-Gaussian state kept for backward; image: per-pixel
 * state kept for backward + tile ranges; binning: sorted instance lists (R = num_rendered).
 * Replaces the resize callbacks of R3/rasterize_points.cu:27-33 — no callbacks cross the ABI. */
size_t pg_raster_geom_bytes(int variant, int P);
size_t pg_raster_image_bytes(int variant, int width, int height);
size_t pg_raster_binning_bytes(int variant, int width, int height, int64_t num_rendered);

typedef struct PgRasterFwdOut {
  float* out_color;     /* [3,H,W]  */
  float* out_depth;     /* [1,H,W]  */
  float* out_opacity;   /* [1,H,W]  (3DGS: 1-T; surfel: opac) */
  int32_t* radii;       /* [P]      */
  int32_t* n_touched;   /* [P] int, zero-initialised by the caller (3DGS) */
  float* out_normal;    /* [3,H,W]  surfel only */
  float* contributions; /* [P] f32, zero-initialised by the caller (surfel) */
} PgRasterFwdOut;

/* Stage 1: preprocess (K1/K2) + depth ordering + offsets.  Enqueues on `stream`, then waits
 * only for the 8-byte instance count and returns it in *num_rendered (the one host sync of
 * the forward, cf. R3 rasterizer_impl.cu:283-284). `radii` is written here. */
int pg_raster_forward_stage1(const PgRasterParams* prm, const PgGaussians* g,
                             void* geom_ws, size_t geom_bytes,
                             int32_t* radii, int64_t* num_rendered, void* stream);
/* Stage 2: duplicate + tile binning + ranges + alpha-blend (K5-K8). */
int pg_raster_forward_stage2(const PgRasterParams* prm, const PgGaussians* g,
                             void* geom_ws, void* binning_ws, size_t binning_bytes,
                             void* image_ws, int64_t num_rendered,
                             const PgRasterFwdOut* out, void* stream);
/* Stage 1 + stage 2 queued back to back, the host waits for the instance count only AFTER the last launch (the
 * reference blocks between its preprocess and its sort, R3 rasterizer_impl.cu:283-284, because it sizes the binning
 * buffer from the count).  The caller passes a binning workspace carved for `capacity` instances
 * (pg_raster_binning_bytes(variant, W, H, capacity); e.g. the previous frame's count plus a margin); the kernels
 * check the device-side count against it and leave without writing when it does not fit.
 * Returns PG_OK with *num_rendered = R when R <= capacity (outputs complete; the backward / debug views take
 * num_rendered = R with this same workspace), PG_ERR_WORKSPACE with *num_rendered = R when R > capacity (the caller
 * allocates for R and calls pg_raster_forward_stage2 - stage 1's results in geom_ws stay valid), PG_ERR_UNSUPPORTED
 * for tile grids off the band-binning path (use the two-stage calls). */
int pg_raster_forward_optimistic(const PgRasterParams* prm, const PgGaussians* g, void* geom_ws, size_t geom_bytes,
                                 int32_t* radii, void* binning_ws, size_t binning_bytes, int64_t capacity,
                                 void* image_ws, const PgRasterFwdOut* out, int64_t* num_rendered, void* stream);

typedef struct PgRasterBwdIn {
  const float* dL_dout_color;   /* [3,H,W] */
  const float* dL_dout_depth;   /* [1,H,W] */
  const float* dL_dout_normal;  /* [3,H,W] surfel only */
  const float* dL_dout_opac;    /* [1,H,W] surfel only */
  const int32_t* radii;         /* [P] from forward */
} PgRasterBwdIn;

typedef struct PgRasterBwdOut {   /* all caller-allocated, need NOT be zeroed */
  float* dL_dmeans2D;   /* [P,3] */
  float* dL_dcolors;    /* [P,3] */
  float* dL_dopacity;   /* [P,1] */
  float* dL_dmeans3D;   /* [P,3] */
  float* dL_dcov3D;     /* [P,6] */
  float* dL_dsh;        /* [P,M,3] or NULL */
  float* dL_dscales;    /* [P,3] */
  float* dL_drotations; /* [P,4] */
  float* dL_dtau;       /* [6]  = sum over Gaussians of the reference's [P,6] (rho, theta)   */
  float* dL_dtau_full;  /* optional [P,6] per-Gaussian pose gradient (reference layout) or NULL */
} PgRasterBwdOut;

/* Backward (K9-K12): R3/rasterize_points.cu:124-211, RS/rasterize_points.cu:133-233. */
int pg_raster_backward(const PgRasterParams* prm, const PgGaussians* g,
                       void* geom_ws, void* binning_ws, void* image_ws, int64_t num_rendered,
                       const PgRasterBwdIn* in, const PgRasterBwdOut* out, void* stream);

/* markVisible (R3/rasterize_points.cu:213-232; RS/rasterize_points.cu:235-254). */
int pg_mark_visible(int variant, int P, const float* means3D, const float* viewmatrix,
                    const float* projmatrix, uint8_t* present, void* stream);

/* Debug/parity accessors into the opaque workspaces (device pointers, valid after stage 2). */
int pg_raster_debug_views(int variant, int P, int width, int height, int64_t num_rendered,
                          void* geom_ws, void* binning_ws, void* image_ws,
                          const uint32_t** point_list, const uint32_t** ranges /* uint2[T] */,
                          const float** final_T, const uint32_t** n_contrib,
                          const float** records /* float4[3P] */, const uint32_t** depth_sorted_idx);

/* Debug/parity accessor: the per-Gaussian gradient accumulation records the blend backward fills (valid after
 * pg_raster_backward).  *gacc points at P * stride float4: 3DGS stride 3 = (dmean2D.xy, dconic.x, dconic.y)
 * (dconic.w, dopacity, ddepth, 0) (dcolor.rgb, 0); surfel stride 4 adds (dnormal.xyz (x10), 0) and keeps the
 * sum of dL_dpixD in the 4th lane of record 1 (the per-pixel-depth term of RS backward.cu:894-897 is applied
 * in the per-surfel kernel).  These are the reference's dL_dconic / dL_ddepths / dL_dnormal scratch tensors
 * (RS rasterize_points.cu:176-188), which it never returns either. */
int pg_raster_debug_gacc(int variant, int P, void* geom_ws, const float** gacc, int* stride_float4);

/* Measurement accessor: number of (Gaussian, tile row, 32-tile band) records the band binning of the
 * last stage 2 on this workspace produced (synchronises `stream`); -1 when the tile grid used another
 * binning path.  Used by bench.py for the algorithmic byte count; no reference counterpart. */
int64_t pg_raster_band_records(int width, int height, int64_t num_rendered, void* binning_ws, void* stream);


/* ------------------------------------------------------------------------------------------
 * Neural-point map: voxel-hash kNN, top-k + inverse-distance interpolation, decoder MLP
 * (NP = model/neural_gaussians.py, DEC = model/decoder.py, utils/mapper.py:2273-2289).
 * The reference implements these as chains of eager torch ops; there is no native symbol to
 * replace, so the entry points mirror the Python methods one-to-one.
 * ---------------------------------------------------------------------------------------- */

/* The global map + search neighbourhood read by NeuralPoints.radius_neighborhood_search (NP:1061-1115). */
typedef struct PgNeuralMap {
  int64_t buffer_size;              /* config.buffer_size (NP:52) */
  const int64_t* buffer_pt_index;   /* [buffer_size] int64 slot table, -1 = empty (NP:86-88) */
  const int32_t* buffer_pt_index32; /* optional int32 mirror of the slot table (pg_np_build_slot_mirror): when non-NULL
                                       the kernels probe it instead (4-B slots, half the footprint: 400 MB at the
                                       default buffer_size 1e8); the caller rebuilds it whenever the table changes */
  const float* neural_points;       /* [n_points,3] */
  int64_t n_points;
  const int64_t* neighbor_dx;       /* [K,3] cell offsets (NP:1026-1059) */
  int32_t K;
  int64_t primes[3];                /* NP:80-82 */
  float resolution;                 /* voxel_size_m */
  float max_valid_dist2;            /* 3((n+1) res)^2 rounded to fp32 */
  const int32_t* point_ts_create;   /* [n_points]; only read when time filtering is on (NP:1088-1094) */
  const float* travel_dist;         /* [n_ts] */
  int32_t cur_ts;
  float diff_travel_dist_local;
} PgNeuralMap;

/* int32 copy of the int64 slot table (indices fit: n_points < 2^31). */
int pg_np_build_slot_mirror(int64_t buffer_size, const int64_t* table64, int32_t* table32, void* stream);

/* radius_neighborhood_search: dist2 [N,K] f32 and idx [N,K] i64 exactly as NP:1061-1115 returns them. */
int pg_np_radius_search(const PgNeuralMap* map, int64_t N, const float* query_xyz, int time_filtering,
                        float* out_dist2, int64_t* out_idx, void* stream);

/* Arguments of NeuralPoints.query_feature (NP:506-725). `points/orientations/certainties/*_features`
 * are the arrays indexed by the FINAL index: the local map when query_locally != 0, else the global one. */
typedef struct PgQueryArgs {
  int32_t nn_k;
  int32_t time_filtering;
  int32_t query_locally;
  int32_t use_only_measured;        /* mask free_gs_mask (NP:544-546) */
  int32_t use_only_valid;           /* mask ~valid_gs_mask (NP:548-550) */
  int32_t weighted_first;           /* config.weighted_first (NP:701-710) */
  int32_t after_pgo;                /* rotate neighbour vectors by the point orientation (NP:627-630) */
  int32_t F_geo, F_color;           /* feature widths; 0 = not queried */
  const uint8_t* free_gs_mask;      /* [n_points] */
  const uint8_t* valid_gs_mask;     /* [n_points] */
  const int64_t* global2local;      /* [n_points+1] (NP:455-466) */
  const float* points;              /* [n,3] */
  const float* orientations;        /* [n,4] (w,x,y,z) */
  const float* certainties;         /* [n] */
  const float* geo_features;        /* [n+1,F_geo] */
  const float* color_features;      /* [n+1,F_color] */
} PgQueryArgs;

typedef struct PgQueryOut {
  float* geo_vec;      /* [N,F_geo+3] if weighted_first else [N,nn_k,F_geo+3]; NULL if F_geo == 0 */
  float* color_vec;    /* same with F_color */
  float* weight;       /* [N,nn_k] */
  int64_t* nn_counts;  /* [N] counted before the top-k cut (NP:557-558) */
  float* certainty;    /* [N] */
  int32_t* knn_idx;    /* [N,nn_k] final index of each kept neighbour, -1 = invalid */
  float* knn_dist2;    /* [N,nn_k] sorted ascending, 9e3 for invalid slots (NP:562) */
} PgQueryOut;

int pg_np_query_feature(const PgNeuralMap* map, const PgQueryArgs* args, int64_t N, const float* query_xyz,
                        const PgQueryOut* out, void* stream);

/* Side effects of query_feature(accumulate_stability=True) (NP:664-689): certainties scatter_add and
 * ts_update amax.  Run after pg_np_query_feature on the same stream. query_ts may be NULL. */
int pg_np_accumulate_stability(int64_t N, int32_t nn_k, const int32_t* knn_idx, const float* weight,
                               const int32_t* query_ts, float* certainties, int32_t* ts_update, void* stream);

/* Backward of query_feature w.r.t. the feature tables and the query coordinates (autograd of NP:506-725:
 * through the gathered features, the neighbour vectors and the IDW weights).  d_*_features are
 * ACCUMULATED into (caller zero-fills); d_query [N,3] is overwritten; any of them may be NULL. */
int pg_np_query_feature_backward(const PgQueryArgs* args, int64_t N, const float* query_xyz,
                                 const int32_t* knn_idx, const float* knn_dist2, const float* weight,
                                 const int64_t* nn_counts, const float* d_geo_vec, const float* d_color_vec,
                                 const float* d_weight, float* d_geo_features, float* d_color_features,
                                 float* d_query, void* stream);

/* Decoder MLP (DEC:14-104): hidden_level x (Linear + ReLU | LeakyReLU(0.01)) + Linear.
 * weights[l] is nn.Linear.weight ([out,in] row-major), biases[l] may be NULL (mlp_bias_on = False). */
#define PG_MLP_MAX_LAYERS 6
typedef struct PgMlp {
  int32_t in_dim, hidden_dim, hidden_level, out_dim;
  int32_t leaky_relu;
  const float* weights[PG_MLP_MAX_LAYERS];
  const float* biases[PG_MLP_MAX_LAYERS];
} PgMlp;

/* y[N,out_dim] = mlp(x[N,in_dim]).  hidden_save (optional) receives the post-activation hidden
 * states [hidden_level][N,hidden_dim] needed by the backward. */
int pg_mlp_forward(const PgMlp* mlp, int64_t N, const float* x, float* y, float* hidden_save, void* stream);
/* Same with an explicit kernel choice.  PG_MLP_TCGEN05: tcgen05/TMEM kernel with 3xTF32 split
 * accumulation (fp32-level accuracy; in_dim, hidden_dim <= 128, out_dim <= 64), which PG_MLP_AUTO
 * picks for batches of >= 4096 rows (Decoder.mlp_batch, DEC:84-98); PG_MLP_SIMT: fp32 FMA kernel. */
#define PG_MLP_AUTO 0
#define PG_MLP_SIMT 1
#define PG_MLP_TCGEN05 2
int pg_mlp_forward_impl(const PgMlp* mlp, int64_t N, const float* x, float* y, float* hidden_save, int impl,
                        void* stream);
/* dX (optional) overwritten; dW[l], dB[l] ACCUMULATED into (caller zero-fills); dB[l] may be NULL. */
int pg_mlp_backward(const PgMlp* mlp, int64_t N, const float* x, const float* hidden_save, const float* dY,
                    float* dX, float* const* dW, float* const* dB, void* stream);

/* Fused inference SDF query = Mapper.sdf under no_grad (utils/mapper.py:2273-2289):
 * kNN -> top-k -> IDW -> decoder -> sdf * sdf_scale (-> weighted mean over k if !weighted_first).
 * sdf_std (optional, !weighted_first only) as utils/mapper.py:2280-2284. */
int pg_sdf_query(const PgNeuralMap* map, const PgQueryArgs* args, const PgMlp* mlp, float sdf_scale, int64_t N,
                 const float* query_xyz, float* sdf, float* sdf_std, int64_t* nn_counts, void* stream);

/* ------------------------------------------------------------------------------------------
 * spawn_gaussians epilogue (gaussian_splatting/gaussian_renderer/__init__.py:551-764 with the quaternion helpers of
 * utils/tools.py:743-751,811-844): what the reference runs as ~60 eager torch kernels between the five decoder
 * heads and the rasterizer.  Raw head outputs are [n, out*K] row-major (Gaussian r = i*K + k reads `out` floats at
 * out*r, the reference's .view(NK, -1)).
 * ---------------------------------------------------------------------------------------- */
typedef struct PgSpawnArgs {
  int64_t n;                 /* spawning neural points */
  int32_t K;                 /* Gaussians per point (Decoder.out_k) */
  const int64_t* sidx;       /* [n] row of each spawning point in the point arrays, or NULL (identity) */
  const float* position;     /* [N,3] */
  const float* orientation;  /* [N,4] (w,x,y,z) */
  const float* color;        /* [N,3] or NULL */
  const float* cam_origin;   /* [3] or NULL */
  const float *xyz_raw, *rot_raw, *scale_raw, *alpha_raw, *color_raw;   /* [n,3K] [n,4K] [n,3K] [n,K] [n,3K] */
  float disp_range;          /* displacement_range_ratio * resolution (GR:605) */
  float unit_scale;          /* unit_scale_ratio * resolution (GR:661) */
  float max_scale;           /* max_scale_ratio * resolution (GR:654) */
  float z_far;               /* > 0: dist_adaptive_scale (GR:656-659), else off */
  float scale_filter;        /* >= 0: scale_filter_on with this threshold (GR:739-751), else off */
  int32_t surfel;            /* third scale = 1e-7 (GR:667-670) */
  int32_t residual;          /* clamp(base + 0.1 tanh(.), 0, 1) (GR:706-711) else sigmoid(.) */
  int32_t alpha_filter;      /* keep alpha > 0 only (GR:727-737) */
  int32_t xy_only;           /* view_direction_xy_only (GR:592) */
} PgSpawnArgs;

typedef struct PgSpawnOut {
  float *xyz, *scale, *rot, *alpha, *color;   /* compacted, capacity n*K rows: [.,3] [.,3] [.,4] [.,1] [.,3] */
  float* alpha_all;                           /* dense [n*K] (GR:717) */
  int32_t* cpos;                              /* dense [n*K]: compact row of the Gaussian, -1 when dropped */
  uint8_t* free_out;                          /* compacted gaussian_free_mask or NULL */
  const uint8_t* free_mask;                   /* [N] per point or NULL */
  uint32_t* state;                            /* [pg_spawn_state_words(n*K)] scratch; after the call state[blocks+1] =
                                                 number of kept Gaussians, state[blocks+2] = 1 if a rotation is NaN */
} PgSpawnOut;

typedef struct PgSpawnGrads {
  const float *d_xyz, *d_scale, *d_rot, *d_alpha, *d_color;   /* compacted cotangents, any may be NULL */
  const float* d_alpha_all;                                   /* dense [n*K] or NULL */
  const int32_t* cpos;                                        /* from the forward */
  float *d_xyz_raw, *d_rot_raw, *d_scale_raw, *d_alpha_raw, *d_color_raw;   /* dense, fully written */
} PgSpawnGrads;

/* Gather the feature rows of the spawning points into the decoder inputs: geo_in [n, Fg (+1 view distance if
 * dist_concat)], color_in [n, Fc (+3 view direction in the point frame if view_concat)] (GR:563-569,587-597,677-698);
 * color_in may be NULL. */
int pg_spawn_prepare(int64_t n, const int64_t* sidx, const float* position, const float* orientation,
                     const float* cam_origin, const float* geo_feature, int Fg, const float* color_feature, int Fc,
                     int dist_concat, int view_concat, int xy_only, float* geo_in, float* color_in, void* stream);
/* d_feature[sidx[i], :F] += d_in[i, :F] (adjoint of the gather; d_in rows are in_w wide). */
int pg_spawn_prepare_backward(int64_t n, const int64_t* sidx, const float* d_in, int in_w, int F, float* d_feature,
                              void* stream);
size_t pg_spawn_state_words(int64_t NK);
/* Word index of the kept-Gaussian count inside `state` (the NaN flag follows it). */
size_t pg_spawn_count_index(int64_t NK);
int pg_spawn_epilogue_forward(const PgSpawnArgs* args, const PgSpawnOut* out, void* stream);
int pg_spawn_epilogue_backward(const PgSpawnArgs* args, const PgSpawnGrads* grads, void* stream);

/* ------------------------------------------------------------------------------------------
 * Tracker registration step (utils/tracker.py:353-607 registration_step, :610-759 implicit_reg /
 * implicit_color_reg).  pg_registration_weights: per source point the validity test (nn mask, gradient-norm window,
 * sdf_std bound), the residual sdf_pred [/ |grad|] - label and the product of the Geman-McClure distance / gradient
 * weights, the normal-consistency weight 0.5 + |n . grad/|grad|| and an optional extra weight; invalid points get
 * residual = weight = 0 (no compaction); sums[4] (double) = {valid count, sum w, sum |residual|, -}.
 * pg_normal_equations: J = [p x g, g]; out28 (double) = the 21 upper-triangle entries of sum w J^T J (row-major),
 * the 6 entries of sum w J r, and sum w r^2.  weight may be NULL (all ones).
 * ---------------------------------------------------------------------------------------- */
int pg_registration_weights(int64_t N, const float* sdf_pred, const float* sdf_grad, const float* sdf_labels,
                            const float* normals, const uint8_t* mask, const float* sdf_std, const float* w_extra,
                            float min_grad_norm, float max_grad_norm, float max_sdf_std, float gm_dist, float gm_grad,
                            int div_grad_norm, float* residual, float* weight, uint8_t* valid, double* sums, void* stream);
int pg_normal_equations(int64_t N, const float* points, const float* grad, const float* residual, const float* weight,
                        double* out28, void* stream);

/* ------------------------------------------------------------------------------------------
 * Gradient exchange of the data-parallel step (the reference is single-GPU, SURVEY.md 0.3, so
 * there is no reference symbol): in-place all-reduce (sum, then * scale) of a flat fp32 buffer that lives in
 * SYMMETRIC memory mapped for NVSwitch multicast on every rank, as ONE kernel: cross-GPU start barrier over the
 * signal pads, multimem.ld_reduce of this rank's 1/world slice (the switch adds the replicas), multimem.st of the
 * sum to all replicas, end barrier.  multicast_ptr: this rank's multicast address of element 0 (16-byte aligned);
 * signal_pad_ptrs_dev: device array [world] of pointers to the ranks' zero-initialised signal pads of
 * signal_pad_bytes each (torch.distributed._symmetric_memory provides all three); numel % 4 == 0.  Every rank must
 * call with the same numel / world / max_ctas (0 = 16 at world <= 2, else 8), each on its own stream; calls on one buffer must not overlap.
 * ---------------------------------------------------------------------------------------- */
int pg_multimem_allreduce_f32(void* multicast_ptr, const void* signal_pad_ptrs_dev, size_t signal_pad_bytes, int64_t numel,
                              int rank, int world, float scale, int max_ctas, void* stream);

/* Allocation-free variant of pg_spawn_epilogue_forward.  It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  The look-back scratch (`out->state`, ignored here and may be NULL) lives
 * in a file-scope `static __device__` array inside the library: 256 KiB per GPU that loads the library, held until
 * the process exits, enough for n*K <= PG_SPAWN_STATIC_MAX_NK candidate Gaussians; larger calls return
 * PG_ERR_UNSUPPORTED with the limit in pg_last_error().  The call does no cudaMalloc / cudaFree and calls no
 * allocating library routine: one cudaMemsetAsync of the scratch words it uses, the kernel, and one 8-byte
 * device-to-device copy on `stream`.  count_nan_out: device pointer to 2 words = {number of kept Gaussians, 1 if a
 * rotation is NaN} (what the allocating entry point leaves at state[pg_spawn_count_index(n*K)]).  All other inputs and
 * outputs are the caller's, exactly as for pg_spawn_epilogue_forward, and the results are identical.
 * Every call on a device shares the one array and the array keeps the last call's contents (the call clears what it
 * needs): calls that OVERLAP on the same GPU (different streams, concurrent host threads) race - serialise them. */
#define PG_SPAWN_STATIC_MAX_NK (1ll << 24)
int pg_spawn_epilogue_forward_static(const PgSpawnArgs* args, const PgSpawnOut* out, uint32_t* count_nan_out,
                                     void* stream);

/* ------------------------------------------------------------------------------------------
 * Post-raster loss assembly (utils/mapper.py:1239,1252-1267): loss = mean|rgb - gt_rgb| +
 * lambda_depth * mean over valid pixels of |gt_depth - depth| (or |1/gt_depth - 1/depth| when
 * inverse_depth), valid = depth_min < gt_depth < depth_max [and alpha > min_alpha].  depth/gt_depth
 * /alpha may be NULL (no depth term / no alpha mask).  The forward also writes the backward seeds
 * (seed_rgb [3,H,W], seed_depth [H,W]) and the reduction scalars sums[3] (double); loss[3] =
 * {total, rgb l1, depth l1}.  The backward scales the seeds by *upstream (device scalar).
 * ---------------------------------------------------------------------------------------- */
int pg_photo_depth_loss_forward(int H, int W, const float* rgb, const float* gt_rgb, const float* depth,
                                const float* gt_depth, const float* alpha, float lambda_depth, float depth_min,
                                float depth_max, float min_alpha, int inverse_depth, float* seed_rgb, float* seed_depth,
                                double* sums, float* loss, void* stream);
int pg_photo_depth_loss_backward(int H, int W, float lambda_depth, const float* upstream, const double* sums,
                                 const float* seed_rgb, const float* seed_depth, float* d_rgb, float* d_depth,
                                 void* stream);

/* depth2normal (gaussian_splatting/utils/point_utils.py:83-148): camera-frame normal map [3,H,W] from a depth map
 * [H,W] and the visibility mask (alpha > min_alpha); cx, cy, fx, fy are in pixels of the (down-scaled) image
 * (prcppoint * size / img_scale, f / img_scale).  alpha (optional) multiplies the output, as GR:381-382 does with
 * the detached rendered alpha.  The backward overwrites d_depth [H,W]. */
int pg_depth2normal_forward(int H, int W, float cx, float cy, float fx, float fy, const float* depth, const uint8_t* mask,
                            const float* alpha, float* normal, void* stream);
int pg_depth2normal_backward(int H, int W, float cx, float cy, float fx, float fy, const float* depth, const uint8_t* mask,
                             const float* alpha, const float* d_normal, float* d_depth, void* stream);
/* Affine exposure correction (GR:449-458): out[3,H,W] = mat[3,3] * rgb + offset[3] per pixel.  The backward
 * overwrites d_img (may be NULL) and d_mat_off[12] = (d_mat row-major, d_offset). */
int pg_exposure_affine_forward(int H, int W, const float* img, const float* mat, const float* offset, float* out,
                               void* stream);
int pg_exposure_affine_backward(int H, int W, const float* img, const float* mat, const float* d_out, float* d_img,
                                float* d_mat_off, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused SSIM (FS/ssim.h:7-26: fusedssim, fusedssim_backward).  img*: [B,C,H,W] contiguous fp32.
 * dm_* may be NULL in the forward (train == false, FS/ssim.cu:458-460).  In the backward dL_dmap
 * may be NULL, in which case the map gradient is the constant dL_dmap_scalar (the backward of the
 * `.mean()` in FS/fused_ssim/__init__.py:42) — one plane less to read.
 * ---------------------------------------------------------------------------------------- */
int pg_ssim_forward(int B, int C, int H, int W, float C1, float C2, const float* img1, const float* img2,
                    float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, void* stream);
int pg_ssim_backward(int B, int C, int H, int W, float C1, float C2, const float* img1, const float* img2,
                     const float* dL_dmap, float dL_dmap_scalar, const float* dm_dmu1, const float* dm_dsigma1_sq,
                     const float* dm_dsigma12, float* dL_dimg1, void* stream);

/* fused_ssim(img1, img2, padding="same") as the caller uses it (FS/fused_ssim/__init__.py:34-42: `map.mean()`): the
 * mean of the SSIM map is reduced inside the forward (sum_scratch: one double of device scratch), the map itself is
 * never written, and the backward takes the upstream gradient of the mean as a device scalar.  dm_* may be NULL
 * (no backward wanted). */
int pg_ssim_mean_forward(int B, int C, int H, int W, float C1, float C2, const float* img1, const float* img2,
                         float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, double* sum_scratch, float* mean,
                         void* stream);
int pg_ssim_mean_backward(int B, int C, int H, int W, const float* img1, const float* img2, const float* upstream,
                          const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1,
                          void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PINGS_B200_H_ */
