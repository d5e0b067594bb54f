"""ctypes binding of the C-ABI shared library (include/pings_b200.h).

There is no CPU fallback: if libpings_b200.so is missing the import of any op fails loudly.
Build it with `python -c "import __graft_entry__ as g; g.build()"` or `pings_b200/build.py`.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PINGS_B200_LIB") or os.path.join(_HERE, "libpings_b200.so")   # env override: dev A/B of two builds

PG_OK = 0
PG_VARIANT_3DGS = 0
PG_VARIANT_SURFEL = 1
_ERR = {1: "bad argument", 2: "CUDA error", 3: "unsupported size", 4: "workspace too small"}

fp = C.c_void_p  # all device pointers travel as void*


class PgRasterParams(C.Structure):
    _fields_ = [
        ("variant", C.c_int32), ("P", C.c_int32), ("sh_degree", C.c_int32), ("sh_coeffs", C.c_int32),
        ("image_width", C.c_int32), ("image_height", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("bg", fp), ("viewmatrix", fp), ("projmatrix", fp), ("projmatrix_raw", fp), ("campos", fp),
        ("patch_bbox", fp), ("prcppoint", fp), ("config", fp),
    ]


class PgGaussians(C.Structure):
    _fields_ = [("means3D", fp), ("shs", fp), ("colors_precomp", fp), ("opacities", fp), ("scales", fp),
                ("rotations", fp), ("cov3D_precomp", fp)]


class PgRasterFwdOut(C.Structure):
    _fields_ = [("out_color", fp), ("out_depth", fp), ("out_opacity", fp), ("radii", fp), ("n_touched", fp),
                ("out_normal", fp), ("contributions", fp)]


class PgRasterBwdIn(C.Structure):
    _fields_ = [("dL_dout_color", fp), ("dL_dout_depth", fp), ("dL_dout_normal", fp), ("dL_dout_opac", fp),
                ("radii", fp)]


class PgRasterBwdOut(C.Structure):
    _fields_ = [("dL_dmeans2D", fp), ("dL_dcolors", fp), ("dL_dopacity", fp), ("dL_dmeans3D", fp),
                ("dL_dcov3D", fp), ("dL_dsh", fp), ("dL_dscales", fp), ("dL_drotations", fp), ("dL_dtau", fp),
                ("dL_dtau_full", fp)]


PG_MLP_MAX_LAYERS = 6


class PgNeuralMap(C.Structure):
    _fields_ = [("buffer_size", C.c_int64), ("buffer_pt_index", fp), ("buffer_pt_index32", fp), ("neural_points", fp),
                ("n_points", C.c_int64),
                ("neighbor_dx", fp), ("K", C.c_int32), ("primes", C.c_int64 * 3), ("resolution", C.c_float),
                ("max_valid_dist2", C.c_float), ("point_ts_create", fp), ("travel_dist", fp), ("cur_ts", C.c_int32),
                ("diff_travel_dist_local", C.c_float)]


class PgQueryArgs(C.Structure):
    _fields_ = [("nn_k", C.c_int32), ("time_filtering", C.c_int32), ("query_locally", C.c_int32),
                ("use_only_measured", C.c_int32), ("use_only_valid", C.c_int32), ("weighted_first", C.c_int32),
                ("after_pgo", C.c_int32), ("F_geo", C.c_int32), ("F_color", C.c_int32),
                ("free_gs_mask", fp), ("valid_gs_mask", fp), ("global2local", fp), ("points", fp),
                ("orientations", fp), ("certainties", fp), ("geo_features", fp), ("color_features", fp)]


class PgQueryOut(C.Structure):
    _fields_ = [("geo_vec", fp), ("color_vec", fp), ("weight", fp), ("nn_counts", fp), ("certainty", fp),
                ("knn_idx", fp), ("knn_dist2", fp)]


class PgMlp(C.Structure):
    _fields_ = [("in_dim", C.c_int32), ("hidden_dim", C.c_int32), ("hidden_level", C.c_int32), ("out_dim", C.c_int32),
                ("leaky_relu", C.c_int32), ("weights", fp * PG_MLP_MAX_LAYERS), ("biases", fp * PG_MLP_MAX_LAYERS)]


class PgSpawnArgs(C.Structure):
    _fields_ = [("n", C.c_int64), ("K", C.c_int32), ("sidx", fp), ("position", fp), ("orientation", fp), ("color", fp),
                ("cam_origin", fp), ("xyz_raw", fp), ("rot_raw", fp), ("scale_raw", fp), ("alpha_raw", fp),
                ("color_raw", fp), ("disp_range", C.c_float), ("unit_scale", C.c_float), ("max_scale", C.c_float),
                ("z_far", C.c_float), ("scale_filter", C.c_float), ("surfel", C.c_int32), ("residual", C.c_int32),
                ("alpha_filter", C.c_int32), ("xy_only", C.c_int32)]


class PgSpawnOut(C.Structure):
    _fields_ = [("xyz", fp), ("scale", fp), ("rot", fp), ("alpha", fp), ("color", fp), ("alpha_all", fp), ("cpos", fp),
                ("free_out", fp), ("free_mask", fp), ("state", fp)]


class PgSpawnGrads(C.Structure):
    _fields_ = [("d_xyz", fp), ("d_scale", fp), ("d_rot", fp), ("d_alpha", fp), ("d_color", fp), ("d_alpha_all", fp),
                ("cpos", fp), ("d_xyz_raw", fp), ("d_rot_raw", fp), ("d_scale_raw", fp), ("d_alpha_raw", fp),
                ("d_color_raw", fp)]


class PingsB200LibraryMissing(ImportError):
    pass


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PingsB200LibraryMissing(
            f"{LIB_PATH} not found: the CUDA extension is not built and pings_b200 has no CPU fallback. "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` at the repo root.")
    L = C.CDLL(LIB_PATH)
    L.pg_abi_version.restype = C.c_int
    L.pg_last_error.restype = C.c_char_p
    L.pg_timing_enable.restype = C.c_int
    L.pg_timing_enable.argtypes = [C.c_int]
    L.pg_timing_drain.restype = C.c_int
    L.pg_timing_drain.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_float)]
    L.pg_raster_geom_bytes.restype = C.c_size_t
    L.pg_raster_geom_bytes.argtypes = [C.c_int, C.c_int]
    L.pg_raster_image_bytes.restype = C.c_size_t
    L.pg_raster_image_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    L.pg_raster_binning_bytes.restype = C.c_size_t
    L.pg_raster_binning_bytes.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64]
    L.pg_mark_visible.restype = C.c_int
    L.pg_mark_visible.argtypes = [C.c_int, C.c_int, fp, fp, fp, fp, fp]
    L.pg_raster_forward_stage1.restype = C.c_int
    L.pg_raster_forward_stage1.argtypes = [C.POINTER(PgRasterParams), C.POINTER(PgGaussians), fp, C.c_size_t, fp,
                                           C.POINTER(C.c_int64), fp]
    L.pg_raster_forward_stage2.restype = C.c_int
    L.pg_raster_forward_stage2.argtypes = [C.POINTER(PgRasterParams), C.POINTER(PgGaussians), fp, fp, C.c_size_t,
                                           fp, C.c_int64, C.POINTER(PgRasterFwdOut), fp]
    L.pg_raster_forward_optimistic.restype = C.c_int
    L.pg_raster_forward_optimistic.argtypes = [C.POINTER(PgRasterParams), C.POINTER(PgGaussians), fp, C.c_size_t, fp, fp,
                                               C.c_size_t, C.c_int64, fp, C.POINTER(PgRasterFwdOut),
                                               C.POINTER(C.c_int64), fp]
    L.pg_raster_backward.restype = C.c_int
    L.pg_raster_backward.argtypes = [C.POINTER(PgRasterParams), C.POINTER(PgGaussians), fp, fp, fp, C.c_int64,
                                     C.POINTER(PgRasterBwdIn), C.POINTER(PgRasterBwdOut), fp]
    L.pg_raster_debug_views.restype = C.c_int
    L.pg_raster_debug_views.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int64, fp, fp, fp] + [C.POINTER(fp)] * 6
    L.pg_raster_debug_gacc.restype = C.c_int
    L.pg_raster_debug_gacc.argtypes = [C.c_int, C.c_int, fp, C.POINTER(fp), C.POINTER(C.c_int)]
    L.pg_raster_band_records.restype = C.c_int64
    L.pg_raster_band_records.argtypes = [C.c_int, C.c_int, C.c_int64, fp, fp]
    L.pg_photo_depth_loss_forward.restype = C.c_int
    L.pg_photo_depth_loss_forward.argtypes = [C.c_int, C.c_int, fp, fp, fp, fp, fp, C.c_float, C.c_float, C.c_float,
                                              C.c_float, C.c_int, fp, fp, fp, fp, fp]
    L.pg_photo_depth_loss_backward.restype = C.c_int
    L.pg_photo_depth_loss_backward.argtypes = [C.c_int, C.c_int, C.c_float, fp, fp, fp, fp, fp, fp, fp]
    L.pg_np_build_slot_mirror.restype = C.c_int
    L.pg_np_build_slot_mirror.argtypes = [C.c_int64, fp, fp, fp]
    L.pg_np_radius_search.restype = C.c_int
    L.pg_np_radius_search.argtypes = [C.POINTER(PgNeuralMap), C.c_int64, fp, C.c_int, fp, fp, fp]
    L.pg_np_query_feature.restype = C.c_int
    L.pg_np_query_feature.argtypes = [C.POINTER(PgNeuralMap), C.POINTER(PgQueryArgs), C.c_int64, fp,
                                      C.POINTER(PgQueryOut), fp]
    L.pg_np_accumulate_stability.restype = C.c_int
    L.pg_np_accumulate_stability.argtypes = [C.c_int64, C.c_int32, fp, fp, fp, fp, fp, fp]
    L.pg_np_query_feature_backward.restype = C.c_int
    L.pg_np_query_feature_backward.argtypes = [C.POINTER(PgQueryArgs), C.c_int64] + [fp] * 12
    L.pg_mlp_forward.restype = C.c_int
    L.pg_mlp_forward.argtypes = [C.POINTER(PgMlp), C.c_int64, fp, fp, fp, fp]
    L.pg_mlp_forward_impl.restype = C.c_int
    L.pg_mlp_forward_impl.argtypes = [C.POINTER(PgMlp), C.c_int64, fp, fp, fp, C.c_int, fp]
    L.pg_mlp_backward.restype = C.c_int
    L.pg_mlp_backward.argtypes = [C.POINTER(PgMlp), C.c_int64, fp, fp, fp, fp, C.POINTER(fp), C.POINTER(fp), fp]
    L.pg_sdf_query.restype = C.c_int
    L.pg_sdf_query.argtypes = [C.POINTER(PgNeuralMap), C.POINTER(PgQueryArgs), C.POINTER(PgMlp), C.c_float, C.c_int64,
                               fp, fp, fp, fp, fp]
    L.pg_ssim_forward.restype = C.c_int
    L.pg_ssim_forward.argtypes = [C.c_int] * 4 + [C.c_float] * 2 + [fp] * 7
    L.pg_ssim_backward.restype = C.c_int
    L.pg_ssim_backward.argtypes = [C.c_int] * 4 + [C.c_float] * 2 + [fp, fp, fp, C.c_float, fp, fp, fp, fp, fp]
    L.pg_ssim_mean_forward.restype = C.c_int
    L.pg_ssim_mean_forward.argtypes = [C.c_int] * 4 + [C.c_float] * 2 + [fp] * 8
    L.pg_ssim_mean_backward.restype = C.c_int
    L.pg_ssim_mean_backward.argtypes = [C.c_int] * 4 + [fp] * 8
    L.pg_depth2normal_forward.restype = C.c_int
    L.pg_depth2normal_forward.argtypes = [C.c_int, C.c_int] + [C.c_float] * 4 + [fp] * 5
    L.pg_depth2normal_backward.restype = C.c_int
    L.pg_depth2normal_backward.argtypes = [C.c_int, C.c_int] + [C.c_float] * 4 + [fp] * 6
    L.pg_exposure_affine_forward.restype = C.c_int
    L.pg_exposure_affine_forward.argtypes = [C.c_int, C.c_int] + [fp] * 5
    L.pg_exposure_affine_backward.restype = C.c_int
    L.pg_exposure_affine_backward.argtypes = [C.c_int, C.c_int] + [fp] * 6
    L.pg_spawn_prepare.restype = C.c_int
    L.pg_spawn_prepare.argtypes = [C.c_int64, fp, fp, fp, fp, fp, C.c_int, fp, C.c_int, C.c_int, C.c_int, C.c_int, fp, fp, fp]
    L.pg_spawn_prepare_backward.restype = C.c_int
    L.pg_spawn_prepare_backward.argtypes = [C.c_int64, fp, fp, C.c_int, C.c_int, fp, fp]
    L.pg_spawn_state_words.restype = C.c_size_t
    L.pg_spawn_state_words.argtypes = [C.c_int64]
    L.pg_spawn_count_index.restype = C.c_size_t
    L.pg_spawn_count_index.argtypes = [C.c_int64]
    L.pg_spawn_epilogue_forward.restype = C.c_int
    L.pg_spawn_epilogue_forward.argtypes = [C.POINTER(PgSpawnArgs), C.POINTER(PgSpawnOut), fp]
    L.pg_registration_weights.restype = C.c_int
    L.pg_registration_weights.argtypes = [C.c_int64, fp, fp, fp, fp, fp, fp, fp, C.c_float, C.c_float, C.c_float, C.c_float,
                                          C.c_float, C.c_int, fp, fp, fp, fp, fp]
    L.pg_normal_equations.restype = C.c_int
    L.pg_normal_equations.argtypes = [C.c_int64, fp, fp, fp, fp, fp, fp]
    L.pg_multimem_allreduce_f32.restype = C.c_int
    L.pg_multimem_allreduce_f32.argtypes = [fp, fp, C.c_size_t, C.c_int64, C.c_int, C.c_int, C.c_float, C.c_int, fp]
    L.pg_spawn_epilogue_forward_static.restype = C.c_int
    L.pg_spawn_epilogue_forward_static.argtypes = [C.POINTER(PgSpawnArgs), C.POINTER(PgSpawnOut), fp, fp]
    L.pg_spawn_epilogue_backward.restype = C.c_int
    L.pg_spawn_epilogue_backward.argtypes = [C.POINTER(PgSpawnArgs), C.POINTER(PgSpawnGrads), fp]
    _declare_optional(L)
    _lib = L
    return L


def _declare_optional(L: C.CDLL) -> None:
    """Entry points added by later sources (ssim, neural points); declared when present."""
    for name, (res, args) in OPTIONAL_SIGNATURES.items():
        if hasattr(L, name):
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args


OPTIONAL_SIGNATURES: dict = {}


def timing_enable(on: bool) -> None:
    lib().pg_timing_enable(int(on))


def timing_drain(max_phases: int = 4096) -> list[tuple[str, float]]:
    """[(phase name, ms)] recorded on this thread since the last drain (synchronises on the events)."""
    names = (C.c_char_p * max_phases)()
    ms = (C.c_float * max_phases)()
    n = lib().pg_timing_drain(max_phases, names, ms)
    if n < 0:
        raise RuntimeError("pings_b200: timing drain failed")
    return [(names[i].decode(), float(ms[i])) for i in range(n)]


def check(rc: int, what: str) -> None:
    if rc != PG_OK:
        msg = lib().pg_last_error().decode() if rc == 2 else ""
        raise RuntimeError(f"pings_b200: {what} failed: {_ERR.get(rc, rc)} {msg}")


class _NoGuard:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


_NO_GUARD = _NoGuard()


def device_guard(dev):
    """`torch.cuda.device(dev)` only when another device is current (the context manager costs ~10 us per call)."""
    idx = dev.index
    if idx is None or idx == torch.cuda.current_device():
        return _NO_GUARD
    return torch.cuda.device(dev)


def raw_stream(dev) -> int:
    """cudaStream_t of torch's current stream on `dev` as an int."""
    idx = dev.index
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device() if idx is None else idx)


def ptr(t) -> int | None:
    """data_ptr of a tensor, or NULL for None / empty tensors (the reference's 'absent' inputs)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
