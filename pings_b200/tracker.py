"""Mirror of the reference tracker's registration path (utils/tracker.py) over the new ops - SURVEY.md §8f row 4:
`Tracker.query_source_points` (:211-351), `registration_step` (:353-607), `tracking` (:42-209) and the functions
`implicit_reg` (:610-688), `implicit_color_reg` (:692-737), `expmap` / `skew` / `rotation_matrix_to_axis_angle`
(:765-795).  Same names, arguments, return tuples and thresholds.

What changes underneath: the SDF / colour queries run on `NeuralPoints.query_feature` + `Decoder` (CUDA kernels, with
the analytical gradient w.r.t. the query through their backward kernels), the per-point validity mask and robust
weights are one kernel (`pg_registration_weights`; invalid points get weight 0 instead of four boolean-mask
compactions), the 6x6 normal equations are one fp64-accumulating reduction (`pg_normal_equations`) instead of two
[N,6] matmuls, and the 6x6 solve happens on the host in fp64 after ONE 256-byte read-back per iteration.
There is no CPU fallback.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib as L

_TRI = [(p, q) for p in range(6) for q in range(p, 6)]


def skew(v):
    S = torch.zeros(3, 3, device=v.device, dtype=v.dtype)
    S[0, 1] = -v[2]
    S[0, 2] = v[1]
    S[1, 2] = -v[0]
    return S - S.T


def expmap(axis_angle: torch.Tensor):
    """utils/tracker.py:774-783 (Rodrigues; a zero vector gives NaN exactly as there)."""
    angle = axis_angle.norm()
    axis = axis_angle / angle
    eye = torch.eye(3, device=axis_angle.device, dtype=axis_angle.dtype)
    S = skew(axis)
    return eye + S * torch.sin(angle) + (S @ S) * (1.0 - torch.cos(angle))


def rotation_matrix_to_axis_angle(R):
    assert torch.is_tensor(R) and R.shape == (3, 3), "Invalid rotation matrix"
    return torch.acos((torch.trace(R) - 1) / 2)


def color_to_intensity(colors: torch.Tensor):
    """utils/tools.py:723-725."""
    return (0.144 * colors[:, 0] + 0.299 * colors[:, 1] + 0.587 * colors[:, 2]).unsqueeze(1)


def _f32(t):
    return t.detach().to(torch.float32).contiguous()


def _normal_eq(points, grad, residual, weight):
    """-> host float64 (N [6,6], b [6] = sum w J r, sum w r^2) from one kernel + one read-back."""
    lib = L.lib()
    dev = points.device
    if not points.is_cuda:
        raise RuntimeError("pings_b200.tracker: expected CUDA tensors (there is no CPU fallback)")
    n = points.shape[0]
    p, g, r = _f32(points), _f32(grad), _f32(residual).reshape(-1)
    w = None if weight is None else _f32(weight).reshape(-1)
    out = torch.empty(28, dtype=torch.float64, device=dev)
    with L.device_guard(dev):
        L.check(lib.pg_normal_equations(n, p.data_ptr(), g.data_ptr(), r.data_ptr(), L.ptr(w), out.data_ptr(),
                                        L.raw_stream(dev)), "normal equations")
    return _unpack(out.cpu().numpy())


def _unpack(h):
    N = np.zeros((6, 6))
    for k, (a, b) in enumerate(_TRI):
        N[a, b] = N[b, a] = h[k]
    return N, h[21:27].copy(), float(h[27])


def _solve(N, g, lm_lambda, device):
    """N += lambda diag(N); t = N^-1 g in fp64; T = [expmap(t[:3]) | t[3:]] (tracker.py:654-668)."""
    N = N + lm_lambda * np.diag(np.diag(N))
    t = np.linalg.inv(N) @ g
    tt = torch.from_numpy(t).to(device=device, dtype=torch.float64)
    T = torch.eye(4, device=device, dtype=torch.float64)
    T[:3, :3] = expmap(tt[:3])
    T[:3, 3] = tt[3:]
    return T


def implicit_reg(points, sdf_grad, sdf_residual, weight, lm_lambda=0.0, require_cov=False, require_eigen=False):
    """utils/tracker.py:610-688: one step of point-to-implicit-model registration.  `weight` [N,1] (or a float)."""
    wt = weight if torch.is_tensor(weight) else torch.full((points.shape[0],), float(weight), device=points.device)
    N, b, wr2 = _normal_eq(points, sdf_grad, sdf_residual, wt)
    # the reference forms N and g in fp32 and solves in fp64
    N32, g32 = N.astype(np.float32).astype(np.float64), (-b).astype(np.float32).astype(np.float64)
    T = _solve(N32, g32, lm_lambda, points.device)
    eigenvalues = cov_mat = None
    if require_eigen:
        eigenvalues = torch.from_numpy(np.linalg.eigvals(N32[3:, 3:].astype(np.float32)).real.astype(np.float32)).to(points.device)
    if require_cov:
        mse = wr2 / max(points.shape[0], 1)
        cov_mat = torch.from_numpy((np.linalg.inv(N32.astype(np.float32)) * np.float32(mse)).astype(np.float32)).to(points.device)
    return T, cov_mat, eigenvalues


def implicit_color_reg(points, sdf_grad, sdf_residual, colors, color_grad, color_residual, weight, w_photo_loss=0.1,
                       lm_lambda=0.0):
    """utils/tracker.py:692-737 (colour channels are intensities by then: usually one)."""
    N, b, _ = _normal_eq(points, sdf_grad, sdf_residual, weight)
    g = -b
    for i in range(colors.shape[1]):
        Nc, bc, _ = _normal_eq(points, color_grad[:, i, :], color_residual[:, i], weight)
        N = N + w_photo_loss * Nc
        g = g - w_photo_loss * bc
    return _solve(N.astype(np.float32).astype(np.float64), g.astype(np.float32).astype(np.float64), lm_lambda, points.device)


def transform_torch(points: torch.Tensor, transformation: torch.Tensor):
    """utils/tools.py transform_torch: homogeneous transform of [N,3] points (result in the points' dtype)."""
    T = transformation.to(points)
    return points @ T[:3, :3].T + T[:3, 3]


class Tracker:
    def __init__(self, config, neural_points, decoders):
        self.config = config
        self.silence = True
        self.neural_points = neural_points
        self.sdf_mlp = decoders["sdf"]
        self.sem_mlp = decoders.get("semantic")
        self.color_mlp = decoders.get("color")
        self.device = config.device
        self.dtype = config.dtype
        self.sdf_scale = config.logistic_gaussian_ratio * config.sigma_sigmoid_m

    # ---- utils/tracker.py:211-351 -------------------------------------------------------------------------------------
    def query_source_points(self, coord, bs, query_sdf=True, query_sdf_grad=True, query_color=False,
                            query_color_grad=False, query_sem=False, query_mask=True, query_certainty=True,
                            query_locally=True, mask_min_nn_count: int = 4):
        n = coord.shape[0]
        dev = coord.device
        z = lambda *s, **k: torch.zeros(*s, device=dev, **k)
        sdf_pred, sdf_std = (z(n), z(n)) if query_sdf else (None, None)
        sem_pred = z(n) if query_sem else None
        cc = getattr(self.config, "color_channel", 0)
        color_pred = z(n, cc) if query_color else None
        mc_mask = z(n, dtype=torch.bool) if query_mask else None
        sdf_grad = z(n, 3) if query_sdf_grad else None
        color_grad = z(n, cc, 3) if query_color_grad else None
        certainty = z(n) if query_certainty else None
        for it in range(math.ceil(n / bs)):
            head, tail = it * bs, min((it + 1) * bs, n)
            batch_coord = coord[head:tail, :].detach()
            want_grad = query_sdf_grad or query_color_grad
            if want_grad:
                batch_coord = batch_coord.clone().requires_grad_(True)
            with torch.set_grad_enabled(want_grad):
                geo, col, weight_knn, nn_count, batch_certainty = self.neural_points.query_feature(
                    batch_coord, accumulate_stability=False, query_locally=query_locally,
                    query_color_feature=query_color, use_only_valid_points=True)
                if query_sdf:
                    batch_sdf = self.sdf_mlp.sdf(geo)
                    if not self.config.weighted_first:
                        mean = torch.sum(batch_sdf * weight_knn, dim=1)
                        var = torch.sum(weight_knn * (batch_sdf - mean.unsqueeze(-1)) ** 2, dim=1)
                        sdf_std[head:tail] = torch.sqrt(var).squeeze(1).detach()
                        batch_sdf = mean.squeeze(1)
                    if query_sdf_grad:
                        g = torch.autograd.grad(batch_sdf, batch_coord, torch.ones_like(batch_sdf), retain_graph=query_color_grad)[0]
                        sdf_grad[head:tail, :] = g.detach()
                    sdf_pred[head:tail] = batch_sdf.detach()
                if query_sem:
                    prob = self.sem_mlp.sem_label_prob(geo)
                    if not self.config.weighted_first:
                        prob = torch.sum(prob * weight_knn, dim=1)
                    sem_pred[head:tail] = torch.argmax(prob, dim=1).detach()
                if query_color:
                    batch_color = self.color_mlp.regress_color(col)
                    if not self.config.weighted_first:
                        batch_color = torch.sum(batch_color * weight_knn, dim=1)
                    if query_color_grad:
                        for i in range(cc):
                            gi = torch.autograd.grad(batch_color[:, i], batch_coord, torch.ones_like(batch_color[:, i]),
                                                     retain_graph=i + 1 < cc)[0]
                            color_grad[head:tail, i, :] = gi.detach()
                    color_pred[head:tail] = batch_color.detach()
            if query_mask:
                mc_mask[head:tail] = nn_count >= mask_min_nn_count
            if query_certainty:
                certainty[head:tail] = batch_certainty.detach()
        return sdf_pred, sdf_grad, color_pred, color_grad, sem_pred, mc_mask, certainty, sdf_std

    # ---- utils/tracker.py:353-607 -------------------------------------------------------------------------------------
    def registration_step(self, points, normals, sdf_labels, colors, min_grad_norm, max_grad_norm, GM_dist=None,
                          GM_grad=None, lm_lambda=0.0, vis_weight_pc=False):
        cfg = self.config
        lib = L.lib()
        dev = points.device
        colors_on = colors is not None and cfg.color_on
        photo_loss_on = bool(getattr(cfg, "photometric_loss_on", False)) and colors_on
        sdf_pred, sdf_grad, color_pred, color_grad, _, mask, certainty, sdf_std = self.query_source_points(
            points, cfg.infer_bs, True, True, colors_on, photo_loss_on, query_locally=True,
            mask_min_nn_count=cfg.track_mask_query_nn_k)
        max_sdf_std = cfg.surface_sample_range_m * cfg.max_sdf_std_ratio
        n = points.shape[0]
        cc = getattr(cfg, "color_channel", 0)
        w_extra = None
        if colors_on:
            col_i, pred_i = colors[:, :cc], color_pred[:, :cc]
            if cc == 3 and cfg.color_on:
                col_i, pred_i = color_to_intensity(col_i), color_to_intensity(pred_i)
            if photo_loss_on:
                cg = color_grad[:, :cc]
                if cc == 3 and cfg.color_on:     # color_to_intensity on [N,3,3]: the same channel mix of the gradients
                    cg = (0.144 * cg[:, 0] + 0.299 * cg[:, 1] + 0.587 * cg[:, 2]).unsqueeze(1)
            elif getattr(cfg, "consist_wieght_on", False):
                w_extra = _f32(torch.exp(-torch.mean(torch.abs(col_i - pred_i), dim=-1)))
        residual = torch.empty(n, dtype=torch.float32, device=dev)
        weight = torch.empty(n, dtype=torch.float32, device=dev)
        valid = torch.empty(n, dtype=torch.uint8, device=dev)
        sums = torch.empty(4, dtype=torch.float64, device=dev)
        out = torch.empty(28, dtype=torch.float64, device=dev)
        p32, g32 = _f32(points), _f32(sdf_grad)
        with L.device_guard(dev):
            st = L.raw_stream(dev)
            L.check(lib.pg_registration_weights(
                n, _f32(sdf_pred).data_ptr(), g32.data_ptr(), _f32(sdf_labels).data_ptr(),
                L.ptr(None if normals is None else _f32(normals)), mask.contiguous().view(torch.uint8).data_ptr(),
                _f32(sdf_std).data_ptr(), L.ptr(w_extra), float(min_grad_norm), float(max_grad_norm), float(max_sdf_std),
                float(GM_dist) if GM_dist is not None else -1.0, float(GM_grad) if GM_grad is not None else -1.0,
                int(bool(cfg.reg_dist_div_grad_norm)), residual.data_ptr(), weight.data_ptr(), valid.data_ptr(),
                sums.data_ptr(), st), "registration weights")
            L.check(lib.pg_normal_equations(n, p32.data_ptr(), g32.data_ptr(), residual.data_ptr(), weight.data_ptr(),
                                            out.data_ptr(), st), "normal equations")
        host = torch.cat([sums, out]).cpu().numpy()          # the one read-back of the step
        count, sum_w, sum_abs = int(round(host[0])), float(host[1]), float(host[2])
        valid_idx = valid.view(torch.bool)
        valid_points = points[valid_idx]
        if count < 10:
            return torch.eye(4, device=dev, dtype=torch.float64), None, None, None, valid_points, 0.0, 0.0
        sdf_residual_mean_cm = sum_abs / count * 100.0
        N, b, wr2 = _unpack(host[4:])
        # w /= 2 mean(w) over the valid points (tracker.py:512-513): the normal equations are linear in w
        no_weights = GM_dist is None and GM_grad is None and normals is None and w_extra is None
        s = 1.0 if no_weights else count / (2.0 * sum_w)
        N, g = N * s, -b * s
        color_residual_mean = None
        if photo_loss_on:
            cres = (pred_i - col_i)
            color_residual_mean = float(torch.mean(torch.abs(cres[valid_idx])))
            for i in range(col_i.shape[1]):
                Nc, bc, _ = _normal_eq(points, cg[:, i, :], cres[:, i], weight)
                N = N + cfg.photometric_loss_weight * s * Nc
                g = g - cfg.photometric_loss_weight * s * bc
            T = _solve(N.astype(np.float32).astype(np.float64), g.astype(np.float32).astype(np.float64), lm_lambda, dev)
            return T, None, None, None, valid_points, sdf_residual_mean_cm, color_residual_mean
        N32, g32h = N.astype(np.float32).astype(np.float64), g.astype(np.float32).astype(np.float64)
        T = _solve(N32, g32h, lm_lambda, dev)
        cov_mat = eigenvalues = None
        if vis_weight_pc:                    # require_cov / require_eigen of implicit_reg (only the last iteration)
            eigenvalues = torch.from_numpy(np.linalg.eigvals(N32[3:, 3:]).real.astype(np.float32)).to(dev)
            mse = wr2 * s / count
            cov_mat = torch.from_numpy((np.linalg.inv(N32) * mse).astype(np.float32)).to(dev)
        return T, cov_mat, eigenvalues, None, valid_points, sdf_residual_mean_cm, color_residual_mean

    # ---- utils/tracker.py:42-209 --------------------------------------------------------------------------------------
    def tracking(self, source_points, init_pose=None, source_colors=None, source_normals=None, source_semantics=None,
                 source_sdf=None, cur_ts=None, loop_reg: bool = False, vis_result: bool = False):
        cfg = self.config
        T = torch.eye(4, dtype=torch.float64, device=self.device) if init_pose is None else init_pose
        cov_mat = None
        GM_dist = cfg.reg_GM_dist_m if cfg.reg_GM_dist_m > 0 else None
        GM_grad = cfg.reg_GM_grad if cfg.reg_GM_grad > 0 else None
        max_valid_final_sdf_residual_cm = cfg.surface_sample_range_m * 0.6 * 100.0
        min_valid_ratio, max_increment_sdf_residual_ratio, eigenvalue_ratio_thre, min_valid_points = 0.05, 1.1, 0.005, 10
        converged, valid_flag, last_sdf_residual_cm = False, True, 1e5
        n_src = source_points.shape[0]
        if source_sdf is None:
            source_sdf = torch.zeros(n_src, device=self.device)
        eigenvalues = weight_point_cloud = None
        valid_point_count, sdf_residual_cm, i = 0, 0.0, 0
        for i in range(cfg.reg_iter_n):
            cur_points = transform_torch(source_points, T)
            delta_T, cov_mat, eigenvalues, weight_point_cloud, valid_points_torch, sdf_residual_cm, _ = self.registration_step(
                cur_points, source_normals, source_sdf, source_colors, cfg.reg_min_grad_norm, cfg.reg_max_grad_norm, GM_dist,
                GM_grad, cfg.reg_lm_lambda, (vis_result and converged))
            T = delta_T @ T
            if (sdf_residual_cm - last_sdf_residual_cm) / last_sdf_residual_cm > max_increment_sdf_residual_ratio:
                valid_flag = False
            else:
                last_sdf_residual_cm = sdf_residual_cm
            valid_point_count = valid_points_torch.shape[0]
            if valid_point_count < min_valid_points or 1.0 * valid_point_count / n_src < min_valid_ratio:
                valid_flag = False
            if not valid_flag or converged:
                break
            rot_angle_deg = float(rotation_matrix_to_axis_angle(delta_T[:3, :3])) * 180.0 / np.pi
            tran_m = float(delta_T[:3, 3].norm())
            if abs(rot_angle_deg) < cfg.reg_term_thre_deg and tran_m < cfg.reg_term_thre_m or i == cfg.reg_iter_n - 2:
                converged = True
        if sdf_residual_cm > max_valid_final_sdf_residual_cm:
            valid_flag = False
        if eigenvalues is not None:
            if cfg.eigenvalue_check and float(torch.min(eigenvalues)) < valid_point_count * eigenvalue_ratio_thre:
                valid_flag = False
        if cov_mat is not None:
            cov_mat = cov_mat.detach().cpu().numpy()
        if not valid_flag and i < 10:
            T = init_pose
            cov_mat = None
        return T, cov_mat, weight_point_cloud, valid_flag
