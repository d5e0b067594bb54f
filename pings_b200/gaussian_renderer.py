"""Mirror of the reference render front-end, gaussian_splatting/gaussian_renderer/__init__.py (GR):
`render()` (GR:27-466) and `spawn_gaussians()` (GR:469-778) with the same signatures, return
dictionaries and early-out rules, running on the pings_b200 ops: `markVisible` + the rasterizers behind
the C-ABI, and the decoder-MLP kernels for the five Gaussian-spawning heads.

Differences from the reference, none of which changes results:
* the three rasterizer packages are imported from pings_b200 (the reference imports the CUDA
  extensions by their top-level names, GR:86-98); `gs_type == "2d_gs"` is not supported (that
  rasterizer is not vendored by the reference either);
* the `get_time()` host synchronisations around every stage (GR:103,203) are dropped.
The elementwise epilogue of `spawn_gaussians` is still eager torch (its fusion is the first "next"
row of SURVEY.md §8f).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch


# ---- quaternion helpers (utils/tools.py:743-751, 811-844) --------------------------------------
def apply_quaternion_rotation(quat: torch.Tensor, points: torch.Tensor) -> torch.Tensor:
    """Passive rotation p' = q p q^-1 with q = (w, x, y, z)."""
    quat_w = quat[..., 0].unsqueeze(-1)
    quat_xyz = -quat[..., 1:]
    t = 2 * torch.linalg.cross(quat_xyz, points)
    return points + quat_w * t + torch.linalg.cross(quat_xyz, t)


def quat_multiply(q1: torch.Tensor, q2: torch.Tensor) -> torch.Tensor:
    w1, x1, y1, z1 = torch.unbind(q1, dim=1)
    w2, x2, y2, z2 = torch.unbind(q2, dim=1)
    w = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2
    x = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2
    y = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2
    z = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2
    return torch.stack((w, x, y, z), dim=1)


def quat_inverse(quat: torch.Tensor) -> torch.Tensor:
    w, x, y, z = torch.unbind(quat, dim=1)
    return torch.stack((w, -x, -y, -z), dim=1)


def depth2normal(depth, mask, camera, img_scale: int = 1):
    """gaussian_splatting/utils/point_utils.py:83-148: normal map (camera frame, [3,H,W]) from a depth map."""
    camD = depth.permute([1, 2, 0])
    mask = mask.permute([1, 2, 0])
    H, W, _ = camD.shape
    device = camD.device
    h, w = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float32),
                          torch.arange(W, device=device, dtype=torch.float32), indexing="ij")
    p = torch.stack([w, h], dim=-1)
    p[..., 0:1] -= camera.prcppoint[0] * camera.image_width / img_scale
    p[..., 1:2] -= camera.prcppoint[1] * camera.image_height / img_scale
    p = p * camD
    K = torch.tensor([camera.fx / img_scale, 0, 0, camera.fy / img_scale]).reshape(2, 2)
    Kinv = torch.inverse(K).to(torch.float32).to(device)
    p = p @ Kinv.t()
    camPos = torch.cat([p, camD], -1)
    p = torch.nn.functional.pad(camPos[None], [0, 0, 1, 1, 1, 1], mode="replicate")
    mask = torch.nn.functional.pad(mask[None].to(torch.float32), [0, 0, 1, 1, 1, 1], mode="replicate").to(torch.bool)
    p_c = p[:, 1:-1, 1:-1, :] * mask[:, 1:-1, 1:-1, :]
    p_u = (p[:, :-2, 1:-1, :] - p_c) * mask[:, :-2, 1:-1, :]
    p_l = (p[:, 1:-1, :-2, :] - p_c) * mask[:, 1:-1, :-2, :]
    p_b = (p[:, 2:, 1:-1, :] - p_c) * mask[:, 2:, 1:-1, :]
    p_r = (p[:, 1:-1, 2:, :] - p_c) * mask[:, 1:-1, 2:, :]
    n = (torch.linalg.cross(p_u, p_l) + torch.linalg.cross(p_r, p_u) + torch.linalg.cross(p_b, p_r)
         + torch.linalg.cross(p_l, p_b))[0]
    mask = mask[0, 1:-1, 1:-1, :]
    n = torch.nn.functional.normalize(n, dim=-1)
    return (n * mask).permute([2, 0, 1])


# ---- spawn_gaussians (GR:469-778) ------------------------------------------------------------------
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
    resolution = neural_points_data["resolution"]
    free_mask = neural_points_data.get("free_mask")
    valid_mask = neural_points_data.get("valid_mask")
    stability = neural_points_data.get("stability")
    gaussian_free_mask = None

    spawn_mask = None
    if visible_mask is not None and valid_mask is not None:
        spawn_mask = visible_mask & valid_mask
    elif visible_mask is not None:
        spawn_mask = visible_mask
    elif valid_mask is not None:
        spawn_mask = valid_mask
    if spawn_mask is not None:
        # One nonzero (one host sync) and index_select gathers instead of a boolean-mask index per tensor
        # (GR:556-569 indexes each tensor with the mask: same rows, same order; every mask index costs a
        # nonzero + a sync forward and a sort-based index_put backward).
        sidx = torch.nonzero(spawn_mask).view(-1)
        position = position.index_select(0, sidx)
        orientation = orientation.index_select(0, sidx)
        if color is not None:
            color = color.index_select(0, sidx)
        if free_mask is not None:
            free_mask = free_mask.index_select(0, sidx)
        if stability is not None:
            stability = stability.index_select(0, sidx)
        # index list padded with -1 = the trailing padding row of the feature tables, GR:563-569 / quirk q19
        visible_idx = torch.cat((sidx, sidx.new_full((1,), geo_features.shape[0] - 1)))
        geo_features = geo_features.index_select(0, visible_idx)
        color_features = color_features.index_select(
            0, torch.cat((sidx, sidx.new_full((1,), color_features.shape[0] - 1))))
    if position.shape[0] < 10:
        return None

    xyz_mlp, scale_mlp, rot_mlp = decoders["gauss_xyz"], decoders["gauss_scale"], decoders["gauss_rot"]
    alpha_mlp, color_mlp = decoders["gauss_alpha"], decoders["gauss_color"]
    n_pts = position.shape[0]

    view_direction = view_distance = None
    if cam_origin is not None:
        cam_origin = cam_origin.float()
        view_direction = position - cam_origin
        if view_direction_xy_only:
            view_direction[:, -1] = 0
        view_distance = view_direction.norm(dim=1, keepdim=True)
        view_direction = view_direction / view_distance

    geo_in = geo_features[:-1]
    disp_range = displacement_range_ratio * resolution * torch.ones((n_pts, 1)).to(position)
    xyz_disp = disp_range * torch.tanh(xyz_mlp.mlp_batch(geo_in))
    K = xyz_mlp.out_k
    NK = n_pts * K

    shifted_position = None
    if record_shifted:
        disp_all = xyz_disp.view(n_pts, 3, K)
        max_disp, max_idx = torch.max(torch.norm(disp_all, dim=1), dim=1)
        disp_max = torch.gather(disp_all, dim=2, index=max_idx.unsqueeze(1).expand(-1, 3).unsqueeze(2)).squeeze(2)
        flag = max_disp > 2.0 * resolution
        shifted_position = position[flag] + disp_max[flag]

    point_quat = orientation.repeat(1, K).view(NK, -1)
    xyz_disp = apply_quaternion_rotation(point_quat, xyz_disp.view(NK, -1))
    gaussian_xyz = position.repeat(1, K).view(NK, -1) + xyz_disp

    gaussian_rot = rot_mlp.mlp_batch(geo_in).view(NK, -1)
    gaussian_rot = torch.nan_to_num(torch.nn.functional.normalize(gaussian_rot), 0, 0)
    gaussian_rot = quat_multiply(point_quat, gaussian_rot)

    dist_ratio = 0.0
    if view_distance is not None and dist_adaptive_scale:
        dist_ratio = (view_distance / z_far).repeat(1, scale_mlp.mlp_out_dim)
    gaussian_scale = unit_scale_ratio * resolution * torch.exp(scale_mlp.mlp_batch(geo_in) + dist_ratio)
    gaussian_scale = torch.clamp(gaussian_scale, max=max_scale_ratio * resolution).view(NK, -1)
    if gs_type == "gaussian_surfel":
        thin = torch.full((NK, 1), 1e-7).to(gaussian_scale)
        gaussian_scale = torch.cat((gaussian_scale[:, :2], thin), dim=1)
    elif gs_type == "2d_gs":
        gaussian_scale = gaussian_scale[:, :2]

    if dist_concat_on and view_distance is not None:
        geo_in = torch.concat((geo_in, view_distance), dim=1)
    gaussian_alpha = torch.tanh(alpha_mlp.mlp_batch(geo_in)).view(NK, -1)

    color_in = color_features[:-1]
    if view_concat_on and view_direction is not None:
        view_direction = apply_quaternion_rotation(quat_inverse(orientation), view_direction)
        color_in = torch.concat((color_in, view_direction), dim=1)
    if learn_color_residual and color is not None:
        residual = 0.1 * torch.tanh(color_mlp.mlp_batch(color_in))
        gaussian_color = torch.clamp(color.repeat(1, K) + residual, 0.0, 1.0)
    else:
        gaussian_color = torch.sigmoid(color_mlp.mlp_batch(color_in))
    gaussian_color = gaussian_color.view(NK, -1)

    alpha_all = gaussian_alpha.clone()
    if free_mask is not None:
        gaussian_free_mask = free_mask.repeat(1, K).view(-1)

    def _keep(mask):
        nonlocal gaussian_xyz, gaussian_scale, gaussian_rot, gaussian_alpha, gaussian_color, gaussian_free_mask
        kidx = torch.nonzero(mask).view(-1)   # one compaction index for all six tensors (see above)
        gaussian_xyz, gaussian_scale = gaussian_xyz.index_select(0, kidx), gaussian_scale.index_select(0, kidx)
        gaussian_rot, gaussian_alpha = gaussian_rot.index_select(0, kidx), gaussian_alpha.index_select(0, kidx)
        gaussian_color = gaussian_color.index_select(0, kidx)
        if gaussian_free_mask is not None:
            gaussian_free_mask = gaussian_free_mask.index_select(0, kidx)

    if alpha_filter_on:
        _keep(gaussian_alpha.squeeze(-1) > 0.0)
    if scale_filter_on:
        _keep(torch.any(gaussian_scale > scale_filter_ratio * resolution, dim=1))

    return {
        "gaussian_xyz": gaussian_xyz, "gaussian_scale": gaussian_scale, "gaussian_rot": gaussian_rot,
        "gaussian_alpha": gaussian_alpha, "gaussian_color": gaussian_color, "alpha_all": alpha_all,
        "gaussian_free_mask": gaussian_free_mask, "local_view_gaussian_count": gaussian_xyz.shape[0],
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
    assert not bool(torch.isnan(rotations).any()), "NaN in rotation"

    if gs_type == "gaussian_surfel":
        image, normal, depth, alpha, radii, contributions = rasterizer(
            means3D=means3D, means2D=screenspace_points, colors_precomp=colors, opacities=opacity, scales=scales,
            rotations=rotations, theta=viewpoint_camera.cam_rot_delta, rho=viewpoint_camera.cam_trans_delta)
        alpha_d = alpha.detach()
        mask_vis = alpha_d > min_alpha
        d2n = depth2normal(depth, mask_vis, viewpoint_camera, img_scale=img_scale) * alpha_d if d2n_on else None
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
        d2n = depth2normal(depth, mask_vis, viewpoint_camera, img_scale=img_scale) * alpha_d if d2n_on else None
        depth = torch.where(mask_vis, depth, torch.zeros_like(depth))
        results.update({"rend_normal": None, "surf_depth": depth, "rend_alpha": alpha, "surf_normal": d2n,
                        "rend_dist": None, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
                        "radii": radii})

    if correct_exposure:   # GR:449-462
        if correct_exposure_affine:
            shp = image.shape
            flat = image.permute(1, 2, 0).reshape(-1, 3)
            flat = flat @ viewpoint_camera.exposure_mat.T + viewpoint_camera.exposure_offset
            image = flat.view(shp[1], shp[2], 3).permute(2, 0, 1)
        else:
            image = torch.exp(viewpoint_camera.exposure_a) * image + viewpoint_camera.exposure_b
    results.update({"render": image})
    return results
