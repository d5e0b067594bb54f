"""Host shim with the positional signatures of the reference's pybind modules
(`diff_gaussian_rasterization._C`, R3/ext.cpp:15-18 + R3/rasterize_points.cu:35-232).

It owns no arithmetic: it allocates outputs/workspaces as torch tensors (so autograd keeps
them alive exactly like the reference's geomBuffer/binningBuffer/imgBuffer) and calls the
C-ABI library on torch's current stream under a device guard.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib as L


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype is torch.float32 and t.is_contiguous():   # the common case: nothing to do, no new tensor object
        return t
    if t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


_stream = L.raw_stream
_guard = L.device_guard

# Instance-count history per (variant, W, H): (instances per Gaussian - a slowly decaying maximum over the recent calls,
# so that a batch of views that alternate does not overflow on every other view - and a validity flag).  The forward
# carves its binning workspace for the predicted count (+ margin) and launches both stages before it waits for the
# true one (pg_raster_forward_optimistic); a wrong guess costs one re-run of stage 2, never a wrong result.
_R_HISTORY: dict = {}
_OPTIMISTIC = os.environ.get("PG_RASTER_OPTIMISTIC", "1") != "0"
_LAST_CAPACITY = [0]     # instance capacity the most recent forward's binning workspace was carved for (diagnostics)


def last_binning_capacity() -> int:
    """What `pg_raster_band_records` / any carve of the most recent forward's binning workspace must be given as
    num_rendered (point_list itself sits at offset 0 whatever the capacity)."""
    return _LAST_CAPACITY[0]


def _forward_stages(V, lib, prm, g, geom, radii, img, out, P, W, H, dev, what):
    """-> (num_rendered, binning workspace).  Both stages of the forward."""
    st = _stream(dev)
    n = C.c_int64(0)
    hist = _R_HISTORY.get((V, W, H)) if _OPTIMISTIC else None
    if hist is not None and hist[0] > 0:
        cap = int(hist[0] * P * 1.25) + 65536
        binning = torch.empty(lib.pg_raster_binning_bytes(V, W, H, cap), dtype=torch.uint8, device=dev)
        rc = lib.pg_raster_forward_optimistic(C.byref(prm), C.byref(g), geom.data_ptr(), geom.numel(), radii.data_ptr(),
                                              binning.data_ptr(), binning.numel(), cap, img.data_ptr(), C.byref(out),
                                              C.byref(n), st)
        rendered = int(n.value)
        if rc == L.PG_OK:
            _R_HISTORY[(V, W, H)] = (max(rendered / P, 0.97 * hist[0]), 1)
            _LAST_CAPACITY[0] = cap
            return rendered, binning
        if rc != 4 and rc != 3:     # PG_ERR_WORKSPACE: guess too small (stage 2 again below); 3: not on the band path
            L.check(rc, what + " (optimistic)")
        if rc == 3:
            _R_HISTORY[(V, W, H)] = (0, 0)
            L.check(lib.pg_raster_forward_stage1(C.byref(prm), C.byref(g), geom.data_ptr(), geom.numel(),
                                                 radii.data_ptr(), C.byref(n), st), what + " (stage 1)")
            rendered = int(n.value)
    else:
        L.check(lib.pg_raster_forward_stage1(C.byref(prm), C.byref(g), geom.data_ptr(), geom.numel(),
                                             radii.data_ptr(), C.byref(n), st), what + " (stage 1)")
        rendered = int(n.value)
        if hist is None:
            _R_HISTORY[(V, W, H)] = (max(rendered / P, 1e-9), 1)
    binning = torch.empty(lib.pg_raster_binning_bytes(V, W, H, rendered), dtype=torch.uint8, device=dev)
    L.check(lib.pg_raster_forward_stage2(C.byref(prm), C.byref(g), geom.data_ptr(), binning.data_ptr(),
                                         binning.numel(), img.data_ptr(), rendered, C.byref(out), st), what + " (stage 2)")
    _LAST_CAPACITY[0] = rendered
    h_ = _R_HISTORY.get((V, W, H), (0, 0))
    if h_[0] > 0:
        _R_HISTORY[(V, W, H)] = (max(rendered / P, 0.97 * h_[0]), 1)
    return rendered, binning


def _params(variant, P, W, H, tanx, tany, mod, prefiltered, debug, degree, M, bg, view, proj, proj_raw, campos,
            patch_bbox=None, prcppoint=None, config=None) -> L.PgRasterParams:
    p = L.PgRasterParams()
    p.variant = variant
    p.P = P
    p.sh_degree = int(degree)
    p.sh_coeffs = int(M)
    p.image_width = int(W)
    p.image_height = int(H)
    p.tanfovx = float(tanx)
    p.tanfovy = float(tany)
    p.scale_modifier = float(mod)
    p.prefiltered = int(bool(prefiltered))
    p.debug = int(bool(debug))
    p.bg = L.ptr(bg)
    p.viewmatrix = L.ptr(view)
    p.projmatrix = L.ptr(proj)
    p.projmatrix_raw = L.ptr(proj_raw)
    p.campos = L.ptr(campos)
    p.patch_bbox = L.ptr(patch_bbox)
    p.prcppoint = L.ptr(prcppoint)
    p.config = L.ptr(config)
    return p


def _gauss(means3D, sh, colors, opacity, scales, rotations, cov3D) -> L.PgGaussians:
    g = L.PgGaussians()
    g.means3D = L.ptr(means3D)
    g.shs = L.ptr(sh)
    g.colors_precomp = L.ptr(colors)
    g.opacities = L.ptr(opacity)
    g.scales = L.ptr(scales)
    g.rotations = L.ptr(rotations)
    g.cov3D_precomp = L.ptr(cov3D)
    return g


def mark_visible(means3D, viewmatrix, projmatrix, variant=L.PG_VARIANT_3DGS):
    """R3/rasterize_points.cu:213-232."""
    lib = L.lib()
    P = means3D.size(0)
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    if P:
        means3D, viewmatrix, projmatrix = _f32c(means3D), _f32c(viewmatrix), _f32c(projmatrix)
        with _guard(means3D.device):
            L.check(lib.pg_mark_visible(variant, P, L.ptr(means3D), L.ptr(viewmatrix), L.ptr(projmatrix),
                                        present.data_ptr(), _stream(means3D.device)), "mark_visible")
    return present


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy, image_height, image_width, sh,
                        degree, campos, prefiltered, debug):
    """R3/rasterize_points.cu:35-122 (same positional arguments, same 9-tuple)."""
    lib = L.lib()
    if means3D.ndim != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    fopt = dict(dtype=torch.float32, device=dev)
    # With P > 0 the kernels write every pixel and every radii entry, so only the accumulated outputs are
    # zero-filled (each fill is one more launch per step); P == 0 returns zeros like the reference.
    alloc = torch.empty if P != 0 else torch.zeros
    out_color = alloc(3, H, W, **fopt)
    out_depth = alloc(1, H, W, **fopt)
    out_opacity = alloc(1, H, W, **fopt)
    radii = alloc(P, dtype=torch.int32, device=dev)
    n_touched = torch.zeros(P, dtype=torch.int32, device=dev)
    bopt = dict(dtype=torch.uint8, device=dev)
    rendered = 0
    if P == 0:
        geom, binning, img = torch.empty(0, **bopt), torch.empty(0, **bopt), torch.empty(0, **bopt)
    else:
        M = sh.size(1) if sh.numel() != 0 else 0
        background, means3D, colors, opacity = _f32c(background), _f32c(means3D), _f32c(colors), _f32c(opacity)
        scales, rotations, cov3D_precomp, sh = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp), _f32c(sh)
        viewmatrix, projmatrix, campos = _f32c(viewmatrix), _f32c(projmatrix), _f32c(campos)
        prm = _params(L.PG_VARIANT_3DGS, P, W, H, tan_fovx, tan_fovy, scale_modifier, prefiltered, debug, degree, M,
                      background, viewmatrix, projmatrix, projmatrix_raw, campos)
        g = _gauss(means3D, sh, colors, opacity, scales, rotations, cov3D_precomp)
        with _guard(dev):
            geom = torch.empty(lib.pg_raster_geom_bytes(L.PG_VARIANT_3DGS, P), **bopt)
            img = torch.empty(lib.pg_raster_image_bytes(L.PG_VARIANT_3DGS, W, H), **bopt)
            out = L.PgRasterFwdOut()
            out.out_color, out.out_depth, out.out_opacity = out_color.data_ptr(), out_depth.data_ptr(), out_opacity.data_ptr()
            out.radii, out.n_touched = radii.data_ptr(), n_touched.data_ptr()
            rendered, binning = _forward_stages(L.PG_VARIANT_3DGS, lib, prm, g, geom, radii, img, out, P, W, H, dev,
                                                "raster forward")
    return rendered, out_color, radii, geom, binning, img, out_depth, out_opacity, n_touched


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_depths, sh, degree, campos, geomBuffer, R, binningBuffer,
                                 imageBuffer, debug, full_tau=False, param_grads_out=None):
    """R3/rasterize_points.cu:124-211.  Returns the same 9-tuple; the last element is the pose
    gradient already reduced over Gaussians as a [1,6] tensor (the reference returns [P,6] and
    sums it in Python, R3 diff_gaussian_rasterization/__init__.py:152); pass full_tau=True for
    the per-Gaussian [P,6] layout.

    param_grads_out (optional): a flat fp32 tensor of 14*P elements; the Gaussian-parameter
    gradients are then written straight into it as [means3D 3P | colors 3P | opacity P | scales 3P |
    rotations 4P] and returned as views, so a data-parallel caller can all-reduce one buffer without
    a gather copy (pings_b200.parallel.FlatGradAllReduce layout)."""
    lib = L.lib()
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    fopt = dict(dtype=torch.float32, device=dev)
    if param_grads_out is not None:
        flat = param_grads_out
        if flat.dtype != torch.float32 or flat.device != dev or flat.numel() != 14 * P or not flat.is_contiguous():
            raise RuntimeError("param_grads_out must be a contiguous fp32 tensor of 14*P elements on the input device")
        flat = flat.view(-1)
        dL_dmeans3D = flat[0:3 * P].view(P, 3)
        dL_dcolors = flat[3 * P:6 * P].view(P, 3)
        dL_dopacity = flat[6 * P:7 * P].view(P, 1)
        dL_dscales = flat[7 * P:10 * P].view(P, 3)
        dL_drotations = flat[10 * P:14 * P].view(P, 4)
    else:
        dL_dmeans3D = torch.empty(P, 3, **fopt)
        dL_dcolors = torch.empty(P, 3, **fopt)
        dL_dopacity = torch.empty(P, 1, **fopt)
        dL_dscales = torch.empty(P, 3, **fopt)
        dL_drotations = torch.empty(P, 4, **fopt)
    dL_dmeans2D = torch.empty(P, 3, **fopt)
    dL_dcov3D = torch.empty(P, 6, **fopt)
    dL_dsh = torch.empty(P, M, 3, **fopt)
    # written in full by the last block of the per-Gaussian backward; the library is not called when P == 0
    dL_dtau = torch.empty(1, 6, **fopt) if P != 0 else torch.zeros(1, 6, **fopt)
    tau_full = torch.empty(P, 6, **fopt) if full_tau else None
    if P != 0:
        background, means3D, colors = _f32c(background), _f32c(means3D), _f32c(colors)
        scales, rotations, cov3D_precomp, sh = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp), _f32c(sh)
        viewmatrix, projmatrix, projmatrix_raw, campos = _f32c(viewmatrix), _f32c(projmatrix), _f32c(projmatrix_raw), _f32c(campos)
        dL_dout_color, dL_dout_depths = _f32c(dL_dout_color), _f32c(dL_dout_depths)
        prm = _params(L.PG_VARIANT_3DGS, P, W, H, tan_fovx, tan_fovy, scale_modifier, False, debug, degree, M,
                      background, viewmatrix, projmatrix, projmatrix_raw, campos)
        g = _gauss(means3D, sh, colors, None, scales, rotations, cov3D_precomp)
        bi = L.PgRasterBwdIn()
        bi.dL_dout_color, bi.dL_dout_depth, bi.radii = dL_dout_color.data_ptr(), dL_dout_depths.data_ptr(), radii.data_ptr()
        bo = L.PgRasterBwdOut()
        bo.dL_dmeans2D, bo.dL_dcolors, bo.dL_dopacity = dL_dmeans2D.data_ptr(), dL_dcolors.data_ptr(), dL_dopacity.data_ptr()
        bo.dL_dmeans3D, bo.dL_dcov3D, bo.dL_dsh = dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), L.ptr(dL_dsh)
        bo.dL_dscales, bo.dL_drotations = dL_dscales.data_ptr(), dL_drotations.data_ptr()
        bo.dL_dtau = dL_dtau.data_ptr()
        bo.dL_dtau_full = L.ptr(tau_full)
        with _guard(dev):
            L.check(lib.pg_raster_backward(C.byref(prm), C.byref(g), geomBuffer.data_ptr(), L.ptr(binningBuffer),
                                           imageBuffer.data_ptr(), int(R), C.byref(bi), C.byref(bo), _stream(dev)),
                    "raster backward")
        if not scales.numel():
            dL_dscales.zero_()
            dL_drotations.zero_()
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
            tau_full if full_tau else dL_dtau)


def rasterize_gaussians_surfel(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                               cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, prcppoint, patchbbox, tan_fovx,
                               tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug, config):
    """RS/rasterize_points.cu:35-131 (same positional arguments, same 10-tuple)."""
    lib = L.lib()
    if means3D.ndim != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    fopt = dict(dtype=torch.float32, device=dev)
    alloc = torch.empty if P != 0 else torch.zeros   # see rasterize_gaussians
    out_color = alloc(3, H, W, **fopt)
    out_normal = alloc(3, H, W, **fopt)
    out_depth = alloc(1, H, W, **fopt)
    out_opac = alloc(1, H, W, **fopt)
    radii = alloc(P, dtype=torch.int32, device=dev)
    contributions = torch.zeros(P, **fopt)
    bopt = dict(dtype=torch.uint8, device=dev)
    rendered = 0
    if P == 0:
        geom, binning, img = torch.empty(0, **bopt), torch.empty(0, **bopt), torch.empty(0, **bopt)
    else:
        V = L.PG_VARIANT_SURFEL
        M = sh.size(1) if sh.numel() != 0 else 0
        background, means3D, colors, opacity = _f32c(background), _f32c(means3D), _f32c(colors), _f32c(opacity)
        scales, rotations, cov3D_precomp, sh = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp), _f32c(sh)
        viewmatrix, projmatrix, campos = _f32c(viewmatrix), _f32c(projmatrix), _f32c(campos)
        prcppoint, patchbbox, config = (_f32c(t.to(dev)) for t in (prcppoint, patchbbox, config))
        prm = _params(V, P, W, H, tan_fovx, tan_fovy, scale_modifier, prefiltered, debug, degree, M, background,
                      viewmatrix, projmatrix, projmatrix_raw, campos, patchbbox, prcppoint, config)
        g = _gauss(means3D, sh, colors, opacity, scales, rotations, cov3D_precomp)
        with _guard(dev):
            geom = torch.empty(lib.pg_raster_geom_bytes(V, P), **bopt)
            img = torch.empty(lib.pg_raster_image_bytes(V, W, H), **bopt)
            out = L.PgRasterFwdOut()
            out.out_color, out.out_depth, out.out_opacity = out_color.data_ptr(), out_depth.data_ptr(), out_opac.data_ptr()
            out.radii, out.out_normal, out.contributions = radii.data_ptr(), out_normal.data_ptr(), contributions.data_ptr()
            rendered, binning = _forward_stages(V, lib, prm, g, geom, radii, img, out, P, W, H, dev, "surfel forward")
    return rendered, out_color, out_normal, out_depth, out_opac, radii, contributions, geom, binning, img


def rasterize_gaussians_surfel_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                        cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, prcppoint, patchbbox,
                                        tan_fovx, tan_fovy, dL_dout_color, dL_dout_normal, dL_dout_depth, dL_dout_opac,
                                        sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug, config,
                                        full_tau=False, param_grads_out=None):
    """RS/rasterize_points.cu:133-233.  Same 9-tuple; the pose gradient is already reduced to [1,6]
    unless full_tau=True; param_grads_out as in rasterize_gaussians_backward."""
    lib = L.lib()
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    fopt = dict(dtype=torch.float32, device=dev)
    if param_grads_out is not None:
        flat = param_grads_out
        if flat.dtype != torch.float32 or flat.device != dev or flat.numel() != 14 * P or not flat.is_contiguous():
            raise RuntimeError("param_grads_out must be a contiguous fp32 tensor of 14*P elements on the input device")
        flat = flat.view(-1)
        dL_dmeans3D = flat[0:3 * P].view(P, 3)
        dL_dcolors = flat[3 * P:6 * P].view(P, 3)
        dL_dopacity = flat[6 * P:7 * P].view(P, 1)
        dL_dscales = flat[7 * P:10 * P].view(P, 3)
        dL_drotations = flat[10 * P:14 * P].view(P, 4)
    else:
        dL_dmeans3D = torch.empty(P, 3, **fopt)
        dL_dcolors = torch.empty(P, 3, **fopt)
        dL_dopacity = torch.empty(P, 1, **fopt)
        dL_dscales = torch.empty(P, 3, **fopt)
        dL_drotations = torch.empty(P, 4, **fopt)
    dL_dmeans2D = torch.empty(P, 3, **fopt)
    dL_dcov3D = torch.empty(P, 6, **fopt)
    dL_dsh = torch.empty(P, M, 3, **fopt)
    # written in full by the last block of the per-Gaussian backward; the library is not called when P == 0
    dL_dtau = torch.empty(1, 6, **fopt) if P != 0 else torch.zeros(1, 6, **fopt)
    tau_full = torch.empty(P, 6, **fopt) if full_tau else None
    if P != 0:
        V = L.PG_VARIANT_SURFEL
        background, means3D, colors = _f32c(background), _f32c(means3D), _f32c(colors)
        scales, rotations, cov3D_precomp, sh = _f32c(scales), _f32c(rotations), _f32c(cov3D_precomp), _f32c(sh)
        viewmatrix, projmatrix, projmatrix_raw, campos = _f32c(viewmatrix), _f32c(projmatrix), _f32c(projmatrix_raw), _f32c(campos)
        prcppoint, patchbbox, config = (_f32c(t.to(dev)) for t in (prcppoint, patchbbox, config))
        dL_dout_color, dL_dout_normal = _f32c(dL_dout_color), _f32c(dL_dout_normal)
        dL_dout_depth, dL_dout_opac = _f32c(dL_dout_depth), _f32c(dL_dout_opac)
        prm = _params(V, P, W, H, tan_fovx, tan_fovy, scale_modifier, False, debug, degree, M, background, viewmatrix,
                      projmatrix, projmatrix_raw, campos, patchbbox, prcppoint, config)
        g = _gauss(means3D, sh, colors, None, scales, rotations, cov3D_precomp)
        bi = L.PgRasterBwdIn()
        bi.dL_dout_color, bi.dL_dout_depth, bi.radii = dL_dout_color.data_ptr(), dL_dout_depth.data_ptr(), radii.data_ptr()
        bi.dL_dout_normal, bi.dL_dout_opac = dL_dout_normal.data_ptr(), dL_dout_opac.data_ptr()
        bo = L.PgRasterBwdOut()
        bo.dL_dmeans2D, bo.dL_dcolors, bo.dL_dopacity = dL_dmeans2D.data_ptr(), dL_dcolors.data_ptr(), dL_dopacity.data_ptr()
        bo.dL_dmeans3D, bo.dL_dcov3D, bo.dL_dsh = dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), L.ptr(dL_dsh)
        bo.dL_dscales, bo.dL_drotations = dL_dscales.data_ptr(), dL_drotations.data_ptr()
        bo.dL_dtau = dL_dtau.data_ptr()
        bo.dL_dtau_full = L.ptr(tau_full)
        with _guard(dev):
            L.check(lib.pg_raster_backward(C.byref(prm), C.byref(g), geomBuffer.data_ptr(), L.ptr(binningBuffer),
                                           imageBuffer.data_ptr(), int(R), C.byref(bi), C.byref(bo), _stream(dev)),
                    "surfel backward")
        if not scales.numel():
            dL_dscales.zero_()
            dL_drotations.zero_()
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
            tau_full if full_tau else dL_dtau)


def debug_views(variant, P, W, H, R, geom, binning, img):
    """Typed torch views of the opaque workspaces, for the parity tests."""
    lib = L.lib()
    outs = [C.c_void_p() for _ in range(6)]
    L.check(lib.pg_raster_debug_views(variant, P, W, H, R, geom.data_ptr(), L.ptr(binning), img.data_ptr(),
                                      *[C.byref(o) for o in outs]), "debug views")

    def view(buf, p, n, dtype):
        if n == 0 or not p.value:
            return torch.empty(0, dtype=dtype, device=buf.device)
        off = p.value - buf.data_ptr()
        nbytes = n * torch.empty(0, dtype=dtype).element_size()
        return buf[off:off + nbytes].view(dtype)

    T = ((W + 15) // 16) * ((H + 15) // 16)
    return dict(
        point_list=view(binning, outs[0], R, torch.int32),
        ranges=view(img, outs[1], 2 * T, torch.int32).view(T, 2),
        final_T=view(img, outs[2], W * H, torch.float32),
        n_contrib=view(img, outs[3], W * H, torch.int32),
        records=view(geom, outs[4], 12 * P, torch.float32).view(P, 3, 4),
        depth_sorted_idx=view(geom, outs[5], P, torch.int32),
    )


def debug_blend_sums(variant, P, geom):
    """The blend backward's per-Gaussian sums (valid after the backward ran on `geom`), as typed views:
    means2D [P,2] (without the surfel per-pixel-depth term), conic [P,3] = (x, y, w), opacity [P], depths [P],
    colors [P,3], and for surfels normal [P,3] (x10) and dpixD_sum [P]."""
    lib = L.lib()
    p, stride = C.c_void_p(), C.c_int(0)
    L.check(lib.pg_raster_debug_gacc(variant, P, geom.data_ptr(), C.byref(p), C.byref(stride)), "debug gacc")
    off = p.value - geom.data_ptr()
    rec = geom[off:off + 16 * stride.value * P].view(torch.float32).view(P, stride.value, 4)
    out = dict(means2D=rec[:, 0, :2], conic=torch.stack([rec[:, 0, 2], rec[:, 0, 3], rec[:, 1, 0]], 1),
               opacity=rec[:, 1, 1], depths=rec[:, 1, 2], colors=rec[:, 2, :3])
    if stride.value == 4:
        out["normal"] = rec[:, 3, :3]
        out["dpixD_sum"] = rec[:, 1, 3]
    return out
