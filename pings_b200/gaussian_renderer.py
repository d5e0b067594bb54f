"""Mirror of the reference render front-end, gaussian_splatting/gaussian_renderer/__init__.py (GR):
`render()` (GR:27-466) and `spawn_gaussians()` (GR:469-778) with the same signatures, return
dictionaries and early-out rules, running on the pings_b200 ops: `markVisible` + the rasterizers behind
the C-ABI, and the decoder-MLP kernels for the five Gaussian-spawning heads.

Differences from the reference, none of which changes results:
* the three rasterizer packages are imported from pings_b200 (the reference imports the CUDA
  extensions by their top-level names, GR:86-98); `gs_type == "2d_gs"` is not supported (that
  rasterizer is not vendored by the reference either);
* the `get_time()` host synchronisations around every stage (GR:103,203) are dropped.
The elementwise epilogue of `spawn_gaussians` runs as fused CUDA kernels
(pings_b200/csrc/pg_spawn.cuh).
"""
from __future__ import annotations

import math
import ctypes as C
from typing import Dict, Optional

import torch

from . import _lib as L
from .decoder import mlp_multi
from .imgops import depth2normal, exposure_affine


# ---- spawn_gaussians (GR:469-778) ------------------------------------------------------------------
# The elementwise epilogue between the five decoder heads and the rasterizer (activations, quaternion rotate /
# multiply of utils/tools.py:743-751,811-844, view-direction transform, colour residual, alpha > 0 compaction) runs
# as three CUDA kernels behind the C-ABI (pings_b200/csrc/pg_spawn.cuh): one gather before the heads, one fused
# epilogue after them (single-pass ordered compaction, device-side count), and their adjoints.
def _stream(dev) -> int:
    return L.raw_stream(dev)


class _SpawnPrepare(torch.autograd.Function):
    """(geo_feature [N+1,Fg], color_feature [N+1,Fc]) -> decoder inputs of the n spawning points:
    geo_in [n,Fg], color_in [n,Fc(+3 view direction in the point frame)]  (GR:563-569,587-597,690-698)."""

    @staticmethod
    def forward(ctx, geo_feature, color_feature, sidx, n, position, orientation, cam_origin, view_concat, xy_only):
        lib = L.lib()
        dev = geo_feature.device
        gf, cf = geo_feature.detach().contiguous(), color_feature.detach().contiguous()
        Fg, Fc = gf.shape[1], cf.shape[1]
        use_view = bool(view_concat and cam_origin is not None)
        geo_in = torch.empty(n, Fg, dtype=torch.float32, device=dev)
        color_in = torch.empty(n, Fc + (3 if use_view else 0), dtype=torch.float32, device=dev)
        with L.device_guard(dev):
            L.check(lib.pg_spawn_prepare(n, L.ptr(sidx), L.ptr(position), L.ptr(orientation), L.ptr(cam_origin),
                                         gf.data_ptr(), Fg, cf.data_ptr(), Fc, 0, int(use_view), int(bool(xy_only)),
                                         geo_in.data_ptr(), color_in.data_ptr(), _stream(dev)), "spawn prepare")
        ctx.sidx, ctx.n = sidx, n
        ctx.shapes = (geo_feature.shape, color_feature.shape)
        return geo_in, color_in

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, d_geo_in, d_color_in):
        lib = L.lib()
        gshape, cshape = ctx.shapes
        outs = [None, None]
        for k, (d, shape) in enumerate(((d_geo_in, gshape), (d_color_in, cshape))):
            if d is None or not ctx.needs_input_grad[k]:
                continue
            d = d.contiguous()
            g = torch.zeros(shape, dtype=torch.float32, device=d.device)
            with L.device_guard(d.device):
                L.check(lib.pg_spawn_prepare_backward(ctx.n, L.ptr(ctx.sidx), d.data_ptr(), d.shape[1], shape[1],
                                                      g.data_ptr(), _stream(d.device)), "spawn prepare backward")
            outs[k] = g
        return outs[0], outs[1], None, None, None, None, None, None, None


class _SpawnEpilogue(torch.autograd.Function):
    """Raw outputs of the five heads -> (xyz, scale, rot, alpha, color) of the kept Gaussians + dense alpha_all."""

    @staticmethod
    def forward(ctx, xyz_raw, rot_raw, scale_raw, alpha_raw, color_raw, c):
        lib = L.lib()
        dev = xyz_raw.device
        raws = [t.detach().to(torch.float32).contiguous() for t in (xyz_raw, rot_raw, scale_raw, alpha_raw, color_raw)]
        n, K = c["n"], c["K"]
        NK = n * K
        a = L.PgSpawnArgs()
        a.n, a.K = n, K
        a.sidx, a.position, a.orientation = L.ptr(c["sidx"]), c["position"].data_ptr(), c["orientation"].data_ptr()
        a.color, a.cam_origin = L.ptr(c["color"]), L.ptr(c["cam_origin"])
        a.xyz_raw, a.rot_raw, a.scale_raw, a.alpha_raw, a.color_raw = [t.data_ptr() for t in raws]
        a.disp_range, a.unit_scale, a.max_scale = c["disp_range"], c["unit_scale"], c["max_scale"]
        a.z_far, a.scale_filter = c["z_far"], c["scale_filter"]
        a.surfel, a.residual, a.alpha_filter, a.xy_only = c["surfel"], c["residual"], c["alpha_filter"], c["xy_only"]
        f32 = dict(dtype=torch.float32, device=dev)
        xyz, scale, rot = torch.empty(NK, 3, **f32), torch.empty(NK, 3, **f32), torch.empty(NK, 4, **f32)
        alpha, color, alpha_all = torch.empty(NK, 1, **f32), torch.empty(NK, 3, **f32), torch.empty(NK, 1, **f32)
        cpos = torch.empty(NK, dtype=torch.int32, device=dev)
        free_out = torch.empty(NK, dtype=torch.uint8, device=dev) if c["free_mask"] is not None else None
        state = torch.empty(lib.pg_spawn_state_words(NK), dtype=torch.int32, device=dev)
        o = L.PgSpawnOut()
        o.xyz, o.scale, o.rot, o.alpha, o.color = (t.data_ptr() for t in (xyz, scale, rot, alpha, color))
        o.alpha_all, o.cpos, o.free_out, o.free_mask = alpha_all.data_ptr(), cpos.data_ptr(), L.ptr(free_out), L.ptr(c["free_mask"])
        o.state = state.data_ptr()
        with L.device_guard(dev):
            L.check(lib.pg_spawn_epilogue_forward(C.byref(a), C.byref(o), _stream(dev)), "spawn epilogue")
        ci = lib.pg_spawn_count_index(NK)
        host = state[ci:ci + 2].cpu()    # the one host read of spawn_gaussians: kept count + NaN flag (GR:305-306 assert)
        M, nan_flag = int(host[0]), int(host[1])
        assert nan_flag == 0, "NaN in rotation"
        ctx.args, ctx.keep = a, (raws, c, cpos)
        ctx.NK = NK
        outs = (xyz[:M], scale[:M], rot[:M], alpha[:M], color[:M], alpha_all)
        fm = free_out[:M].view(torch.bool) if free_out is not None else None
        if fm is not None:
            ctx.mark_non_differentiable(cpos, fm)
        else:
            ctx.mark_non_differentiable(cpos)
        return (*outs, cpos, fm)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, d_xyz, d_scale, d_rot, d_alpha, d_color, d_alpha_all, _cpos, _fm):
        lib = L.lib()
        raws, c, cpos = ctx.keep
        dev = cpos.device
        K, n, NK = c["K"], c["n"], ctx.NK
        cont = lambda t: None if t is None else t.to(torch.float32).contiguous()
        d_xyz, d_scale, d_rot, d_alpha, d_color, d_alpha_all = map(cont, (d_xyz, d_scale, d_rot, d_alpha, d_color, d_alpha_all))
        f32 = dict(dtype=torch.float32, device=dev)
        g_xyz, g_rot, g_scale = torch.empty(n, 3 * K, **f32), torch.empty(n, 4 * K, **f32), torch.empty(n, 3 * K, **f32)
        g_alpha, g_color = torch.empty(n, K, **f32), torch.empty(n, 3 * K, **f32)
        g = L.PgSpawnGrads()
        g.d_xyz, g.d_scale, g.d_rot, g.d_alpha, g.d_color = (L.ptr(t) for t in (d_xyz, d_scale, d_rot, d_alpha, d_color))
        g.d_alpha_all, g.cpos = L.ptr(d_alpha_all), cpos.data_ptr()
        g.d_xyz_raw, g.d_rot_raw, g.d_scale_raw = g_xyz.data_ptr(), g_rot.data_ptr(), g_scale.data_ptr()
        g.d_alpha_raw, g.d_color_raw = g_alpha.data_ptr(), g_color.data_ptr()
        with L.device_guard(dev):
            L.check(lib.pg_spawn_epilogue_backward(C.byref(ctx.args), C.byref(g), _stream(dev)), "spawn epilogue backward")
        return g_xyz, g_rot, g_scale, g_alpha, g_color, None


def spawn_gaussians(neural_points_data: Dict, decoders: Dict, visible_mask: Optional[torch.Tensor] = None,
                    cam_origin: Optional[torch.Tensor] = None, dist_concat_on: bool = False,
                    view_concat_on: bool = False, alpha_filter_on: bool = True, scale_filter_on: bool = False,
                    z_far: float = 100.0, dist_adaptive_scale: bool = False, learn_color_residual: bool = True,
                    view_direction_xy_only: bool = True, gs_type: str = "gaussian_surfel",
                    displacement_range_ratio: float = 1.0, max_scale_ratio: float = 1.0,
                    unit_scale_ratio: float = 0.2, scale_filter_ratio: float = 0.2, record_shifted: bool = False):
    position = neural_points_data["position"]
    orientation = neural_points_data["orientation"]
    color = neural_points_data.get("color")
    geo_features = neural_points_data["geo_feature"]
    color_features = neural_points_data["color_feature"]
    resolution = float(neural_points_data["resolution"])
    free_mask = neural_points_data.get("free_mask")
    valid_mask = neural_points_data.get("valid_mask")
    if gs_type == "2d_gs":
        raise NotImplementedError("2d_gs: diff_surfel_rasterization is not vendored by the reference (SURVEY.md 0.5)")
    if not position.is_cuda:
        raise RuntimeError("pings_b200.spawn_gaussians: expected CUDA tensors (there is no CPU fallback)")

    spawn_mask = None
    if visible_mask is not None and valid_mask is not None:
        spawn_mask = visible_mask & valid_mask
    elif visible_mask is not None:
        spawn_mask = visible_mask
    elif valid_mask is not None:
        spawn_mask = valid_mask
    # the spawning points are addressed through an index list (the kernels gather; GR:556-569 materialises every
    # masked tensor): one nonzero, i.e. one host read for the row count
    sidx = torch.nonzero(spawn_mask).view(-1) if spawn_mask is not None else None
    n_pts = int(sidx.shape[0]) if sidx is not None else int(position.shape[0])
    if n_pts < 10:
        return None

    xyz_mlp, scale_mlp, rot_mlp = decoders["gauss_xyz"], decoders["gauss_scale"], decoders["gauss_rot"]
    alpha_mlp, color_mlp = decoders["gauss_alpha"], decoders["gauss_color"]
    K = xyz_mlp.out_k
    f32c = lambda t: None if t is None else t.detach().to(torch.float32).contiguous()
    pos_c, quat_c, col_c = f32c(position), f32c(orientation), f32c(color)
    cam_c = f32c(cam_origin)
    geo_in, color_in = _SpawnPrepare.apply(geo_features, color_features, sidx, n_pts, pos_c, quat_c, cam_c,
                                           bool(view_concat_on), bool(view_direction_xy_only))
    # the heads that read the gathered geo features run as ONE autograd node (decoder.mlp_multi: same kernels, a
    # quarter of the Function / backward-node overhead); GR:609-676 calls them one by one
    shared_alpha = not (dist_concat_on and cam_c is not None)
    heads = [xyz_mlp, rot_mlp, scale_mlp] + ([alpha_mlp] if shared_alpha else [])
    outs = mlp_multi(heads, geo_in)
    xyz_raw, rot_raw, scale_raw = outs[0], outs[1], outs[2]
    if shared_alpha:
        alpha_raw = outs[3]
    else:                                        # GR:677-678: the view distance feeds the alpha head only
        p_sel = pos_c if sidx is None else pos_c.index_select(0, sidx)
        vd = p_sel - cam_c
        if view_direction_xy_only:
            vd = torch.cat((vd[:, :2], torch.zeros_like(vd[:, 2:])), 1)
        alpha_raw = alpha_mlp.mlp_batch(torch.cat((geo_in, vd.norm(dim=1, keepdim=True)), dim=1))
    color_raw = color_mlp.mlp_batch(color_in)

    shifted_position = None
    if record_shifted:   # GR:614-626 (visualisation aid; note the reference's [n, 3, K] view of the [n, 3K] output)
        disp_all = (displacement_range_ratio * resolution * torch.tanh(xyz_raw.detach())).view(n_pts, 3, K)
        max_disp, max_idx = torch.max(torch.norm(disp_all, dim=1), dim=1)
        disp_max = torch.gather(disp_all, dim=2, index=max_idx.unsqueeze(1).expand(-1, 3).unsqueeze(2)).squeeze(2)
        flag = max_disp > 2.0 * resolution
        p_sel = pos_c if sidx is None else pos_c.index_select(0, sidx)
        shifted_position = p_sel[flag] + disp_max[flag]

    consts = dict(n=n_pts, K=int(K), sidx=sidx, position=pos_c, orientation=quat_c,
                  color=col_c if learn_color_residual else None, cam_origin=cam_c,
                  disp_range=float(displacement_range_ratio * resolution), unit_scale=float(unit_scale_ratio * resolution),
                  max_scale=float(max_scale_ratio * resolution),
                  z_far=float(z_far) if (dist_adaptive_scale and cam_c is not None) else -1.0,
                  scale_filter=float(scale_filter_ratio * resolution) if scale_filter_on else -1.0,
                  surfel=int(gs_type == "gaussian_surfel"), residual=int(bool(learn_color_residual and color is not None)),
                  alpha_filter=int(bool(alpha_filter_on)), xy_only=int(bool(view_direction_xy_only)),
                  free_mask=None if free_mask is None else free_mask.detach().contiguous().view(torch.uint8))
    fn_out = _SpawnEpilogue.apply(xyz_raw, rot_raw, scale_raw, alpha_raw, color_raw, consts)
    g_xyz, g_scale, g_rot, g_alpha, g_color, alpha_all, cpos, gaussian_free_mask = fn_out
    return {
        "gaussian_xyz": g_xyz, "gaussian_scale": g_scale, "gaussian_rot": g_rot,
        "gaussian_alpha": g_alpha, "gaussian_color": g_color, "alpha_all": alpha_all,
        "gaussian_free_mask": gaussian_free_mask, "local_view_gaussian_count": g_xyz.shape[0],
        "shifted_position": shifted_position,
    }


_SURFEL_CFG: Dict = {}


# ---- render (GR:27-466) --------------------------------------------------------------------------
def render(viewpoint_camera, cam_pose: Optional[torch.Tensor], neural_points_data: Optional[Dict],
           decoders: Optional[Dict], gaussians: Optional[Dict], bg_color: torch.Tensor,
           scaling_modifier: float = 1.0, down_rate: int = 0, min_visible_neural_point_ratio: float = 0.0,
           verbose: bool = False, replay_mode: bool = False, dist_concat_on: bool = False,
           view_concat_on: bool = False, correct_exposure: bool = True, correct_exposure_affine: bool = True,
           learn_color_residual: bool = False, front_only_on: bool = True, d2n_on: bool = False,
           gs_type: str = "gaussian_surfel", use_median_depth: bool = False, min_alpha: float = 1e-3,
           displacement_range_ratio: float = 1.0, max_scale_ratio: float = 1.0, unit_scale_ratio: float = 0.2):
    """Same contract as the reference: returns the result dict, or None when there is nothing to render
    (no visible neural points GR:221-224, too few in replay mode GR:228-232, inputs None GR:264-265,
    <= 10 Gaussians GR:291-292)."""
    if gs_type == "gaussian_surfel":
        from .diff_gaussian_surfel_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    elif gs_type == "2d_gs":
        raise NotImplementedError("2d_gs: diff_surfel_rasterization is not vendored by the reference (SURVEY.md 0.5)")
    else:
        if gs_type != "3d_gs":
            print("wrong gs type selected, use the default one 3d gs")
            gs_type = "3d_gs"
        from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    dtype = torch.float32
    device = viewpoint_camera.device
    img_scale = 2 ** down_rate
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    z_far = viewpoint_camera.zfar

    if cam_pose is not None:  # GR:117-131
        cam_pose = cam_pose.to(dtype=dtype, device=device)
        T_cw = torch.linalg.inv(cam_pose)
        wvt = T_cw.T
        viewpoint_camera.world_view_transform = wvt
        viewpoint_camera.full_proj_transform = wvt @ viewpoint_camera.projection_matrix
        viewpoint_camera.camera_center = torch.linalg.inv(wvt)[3, :3]
        viewpoint_camera.R = T_cw[:3, :3]
        viewpoint_camera.T = T_cw[:3, 3]

    W = int(viewpoint_camera.image_width / img_scale)   # full-resolution tanfov is kept (quirk q20)
    H = int(viewpoint_camera.image_height / img_scale)
    common = dict(image_height=H, image_width=W, tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color,
                  scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
                  projmatrix=viewpoint_camera.full_proj_transform, projmatrix_raw=viewpoint_camera.projection_matrix,
                  sh_degree=0, campos=viewpoint_camera.camera_center, prefiltered=False, debug=False)
    if gs_type == "gaussian_surfel":
        key = (str(device), bool(front_only_on))   # GR:137-142; cached: torch.tensor(list, device=cuda) is a blocking H2D copy
        cfg = _SURFEL_CFG.get(key)
        if cfg is None:
            cfg = _SURFEL_CFG[key] = torch.tensor([True, True, True, True, front_only_on], dtype=dtype, device=device)
        raster_settings = GaussianRasterizationSettings(patch_bbox=viewpoint_camera.full_patch(down_rate),
                                                        prcppoint=viewpoint_camera.prcppoint, config=cfg, **common)
    else:
        raster_settings = GaussianRasterizationSettings(**common)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    if neural_points_data is None or decoders is None:
        return None
    visible = rasterizer.markVisible(neural_points_data["position"])
    n_total = visible.shape[0]
    n_visible = int(torch.sum(visible).item())
    if n_visible == 0:
        if verbose:
            print("[Render] No visible neural points, skip this frame {}".format(viewpoint_camera.uid))
        return None
    ratio = 1.0 * n_visible / n_total
    if ratio < min_visible_neural_point_ratio and replay_mode:
        if verbose:
            print("[Render] Too small ratio of visible neural points, skip this frame {}".format(viewpoint_camera.uid))
        return None
    spawn = spawn_gaussians(neural_points_data, decoders, visible, viewpoint_camera.camera_center, dist_concat_on,
                            view_concat_on, z_far=z_far, learn_color_residual=learn_color_residual, gs_type=gs_type,
                            displacement_range_ratio=displacement_range_ratio, max_scale_ratio=max_scale_ratio,
                            unit_scale_ratio=unit_scale_ratio)
    if spawn is None:
        e = lambda c: torch.empty((0, c), dtype=dtype, device=device)
        g_xyz, g_scale, g_rot, g_alpha, g_color = e(3), e(3), e(4), e(1), e(3)
        results = {}
    else:
        g_xyz, g_scale, g_rot = spawn["gaussian_xyz"], spawn["gaussian_scale"], spawn["gaussian_rot"]
        g_alpha, g_color = spawn["gaussian_alpha"], spawn["gaussian_color"]
        spawn["visible_neural_point_ratio"] = ratio
        results = spawn

    if gaussians is not None and gaussians["gaussian_xyz"].shape[0] > 10:   # surrounding (frozen) Gaussians
        means3D = torch.cat((g_xyz, gaussians["gaussian_xyz"]), 0)
        opacity = torch.cat((g_alpha, gaussians["gaussian_alpha"]), 0)
        scales = torch.cat((g_scale, gaussians["gaussian_scale"]), 0)
        rotations = torch.cat((g_rot, gaussians["gaussian_rot"]), 0)
        colors = torch.cat((g_color, gaussians["gaussian_color"]), 0)
    else:
        means3D, opacity, scales, rotations, colors = g_xyz, g_alpha, g_scale, g_rot, g_color
    if means3D.shape[0] <= 10:
        return None

    screenspace_points = torch.zeros_like(means3D, requires_grad=True, dtype=dtype, device=device)
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    # GR:305-306 asserts that no rotation is NaN: the spawned ones were checked inside the fused epilogue (its flag
    # travels with the kept count), the frozen surrounding ones are checked here only when present
    if gaussians is not None and gaussians["gaussian_xyz"].shape[0] > 10:
        assert not bool(torch.isnan(gaussians["gaussian_rot"]).any()), "NaN in rotation"

    if gs_type == "gaussian_surfel":
        image, normal, depth, alpha, radii, contributions = rasterizer(
            means3D=means3D, means2D=screenspace_points, colors_precomp=colors, opacities=opacity, scales=scales,
            rotations=rotations, theta=viewpoint_camera.cam_rot_delta, rho=viewpoint_camera.cam_trans_delta)
        alpha_d = alpha.detach()
        mask_vis = alpha_d > min_alpha
        d2n = depth2normal(depth, mask_vis, viewpoint_camera, img_scale=img_scale, alpha=alpha_d) if d2n_on else None
        results.update({"rend_normal": normal, "surf_depth": depth, "rend_alpha": alpha, "surf_normal": d2n,
                        "rend_dist": None, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
                        "radii": radii, "contributions": contributions})
    else:
        image, radii, depth, alpha, n_touched = rasterizer(
            means3D=means3D, means2D=screenspace_points, colors_precomp=colors, opacities=opacity, scales=scales,
            rotations=rotations, theta=viewpoint_camera.cam_rot_delta, rho=viewpoint_camera.cam_trans_delta)
        alpha_d = alpha.detach()
        mask_vis = alpha_d > min_alpha
        # GR:426-437: depth / alpha where visible, 0 elsewhere (the reference edits the tensor in place)
        depth = torch.where(mask_vis, depth / torch.where(mask_vis, alpha_d, torch.ones_like(alpha_d)), depth)
        d2n = depth2normal(depth, mask_vis, viewpoint_camera, img_scale=img_scale, alpha=alpha_d) if d2n_on else None
        depth = torch.where(mask_vis, depth, torch.zeros_like(depth))
        results.update({"rend_normal": None, "surf_depth": depth, "rend_alpha": alpha, "surf_normal": d2n,
                        "rend_dist": None, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
                        "radii": radii})

    if correct_exposure:   # GR:449-462
        if correct_exposure_affine:
            image = exposure_affine(image, viewpoint_camera.exposure_mat, viewpoint_camera.exposure_offset)
        else:
            image = torch.exp(viewpoint_camera.exposure_a) * image + viewpoint_camera.exposure_b
    results.update({"render": image})
    return results
