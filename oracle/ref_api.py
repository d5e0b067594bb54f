"""TEST INFRASTRUCTURE — ctypes access to the reference's own CUDA code built by
oracle/build_ref.py into oracle/_ref/.  Only tests/, __graft_entry__.smoke() and bench.py's
comparator leg import this module; the product (pings_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
vp = C.c_void_p


def available(name: str) -> bool:
    return os.path.exists(os.path.join(REF_DIR, f"lib{name}.so"))


_libs: dict = {}


def _load(name: str) -> C.CDLL:
    if name not in _libs:
        _libs[name] = C.CDLL(os.path.join(REF_DIR, f"lib{name}.so"))
    return _libs[name]


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


class RefRaster3:
    """The unmodified reference 3DGS rasterizer (R3) behind oracle/ref_r3_capi.cu."""

    def __init__(self):
        L = _load("ref_r3")
        L.ref_r3_create.restype = vp
        L.ref_r3_destroy.argtypes = [vp]
        L.ref_r3_forward.restype = C.c_int
        L.ref_r3_forward.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp,
                                     C.c_float, vp, vp, vp, vp, vp, C.c_float, C.c_float, C.c_int, vp, vp, vp, vp, vp]
        L.ref_r3_views.argtypes = [vp] + [C.POINTER(vp)] * 10
        L.ref_r3_backward.restype = C.c_int
        L.ref_r3_backward.argtypes = [vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_float, vp, vp, vp, vp, vp, vp,
                                      C.c_float, C.c_float, vp] + [vp] * 13
        L.ref_r3_mark_visible.argtypes = [C.c_int, vp, vp, vp, vp]
        L.ref_r3_copy.argtypes = [vp, vp, C.c_size_t]
        self.L = L
        self.h = L.ref_r3_create()

    def __del__(self):
        try:
            self.L.ref_r3_destroy(self.h)
        except Exception:
            pass

    def forward(self, scene: dict, cam_mats, W, H, tanx, tany, bg, mod=1.0, sh=None, degree=0, campos=None,
                cov3D=None, want_views=True):
        """scene tensors must be CUDA fp32 contiguous. Returns dict of outputs + internal views."""
        torch.cuda.synchronize()
        view, proj, proj_raw = cam_mats
        means = scene["means3D"]
        dev = means.device
        P = means.shape[0]
        M = 0 if sh is None else sh.shape[1]
        out_color = torch.zeros(3, H, W, device=dev)
        out_depth = torch.zeros(1, H, W, device=dev)
        out_opacity = torch.zeros(1, H, W, device=dev)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        n_touched = torch.zeros(P, dtype=torch.int32, device=dev)
        colors = scene.get("colors_precomp") if sh is None else None
        self._keep = (scene, cam_mats, bg, sh, campos, cov3D)
        R = self.L.ref_r3_forward(self.h, P, degree, M, _p(bg), W, H, _p(means), _p(sh), _p(colors),
                                  _p(scene["opacities"]), _p(scene.get("scales")) if cov3D is None else None, mod,
                                  _p(scene.get("rotations")) if cov3D is None else None, _p(cov3D), _p(view), _p(proj),
                                  _p(campos), tanx, tany, 0, _p(out_color), _p(out_depth), _p(out_opacity),
                                  _p(radii), _p(n_touched))
        if R < 0:
            raise RuntimeError("reference forward failed")
        self.P, self.W, self.H, self.R = P, W, H, R
        res = dict(num_rendered=R, color=out_color, depth=out_depth, opacity=out_opacity, radii=radii,
                   n_touched=n_touched)
        if want_views:
            res.update(self.views())
        return res

    def views(self) -> dict:
        ptrs = [vp() for _ in range(10)]
        self.L.ref_r3_views(self.h, *[C.byref(p) for p in ptrs])
        P, W, H, R = self.P, self.W, self.H, self.R
        T = ((W + 15) // 16) * ((H + 15) // 16)

        def copy(p, n, dtype):
            t = torch.empty(n, dtype=dtype, device="cuda")
            if n and p.value:
                self.L.ref_r3_copy(vp(t.data_ptr()), p, C.c_size_t(n * t.element_size()))
            return t

        return dict(
            point_list=copy(ptrs[0], R, torch.int32),
            keys=copy(ptrs[1], R, torch.int64),
            ranges=copy(ptrs[2], 2 * T, torch.int32).view(T, 2),
            n_contrib=copy(ptrs[3], W * H, torch.int32),
            final_T=copy(ptrs[4], W * H, torch.float32),
            means2D=copy(ptrs[5], 2 * P, torch.float32).view(P, 2),
            depths=copy(ptrs[6], P, torch.float32),
            conic_opacity=copy(ptrs[7], 4 * P, torch.float32).view(P, 4),
            tiles_touched=copy(ptrs[8], P, torch.int32),
            cov3D=copy(ptrs[9], 6 * P, torch.float32).view(P, 6),
        )

    def backward(self, scene: dict, cam_mats, tanx, tany, bg, radii, dL_dcolor, dL_ddepth, mod=1.0, sh=None,
                 degree=0, campos=None, cov3D=None):
        view, proj, proj_raw = cam_mats
        means = scene["means3D"]
        dev = means.device
        P = means.shape[0]
        M = 0 if sh is None else sh.shape[1]
        z = lambda *s: torch.zeros(*s, device=dev)
        g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), depths=z(P, 1), means3D=z(P, 3),
                 cov3D=z(P, 6), sh=z(P, max(M, 1), 3), scales=z(P, 3), rotations=z(P, 4), tau=z(P, 6))
        colors = scene.get("colors_precomp") if sh is None else None
        rc = self.L.ref_r3_backward(self.h, degree, M, _p(bg), _p(means), _p(sh), _p(colors),
                                    _p(scene.get("scales")) if cov3D is None else None, mod,
                                    _p(scene.get("rotations")) if cov3D is None else None, _p(cov3D), _p(view), _p(proj),
                                    _p(proj_raw), _p(campos), tanx, tany, _p(radii), _p(dL_dcolor), _p(dL_ddepth),
                                    _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["colors"]), _p(g["depths"]),
                                    _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]),
                                    _p(g["tau"]))
        if rc != 0:
            raise RuntimeError("reference backward failed")
        return g

    def mark_visible(self, means, view, proj):
        present = torch.zeros(means.shape[0], dtype=torch.bool, device=means.device)
        self.L.ref_r3_mark_visible(means.shape[0], _p(means), _p(view), _p(proj), _p(present))
        return present


class RefRasterS:
    """The unmodified reference Gaussian-surfel rasterizer (RS) behind oracle/ref_rs_capi.cu."""

    def __init__(self):
        L = _load("ref_rs")
        L.ref_rs_create.restype = vp
        L.ref_rs_destroy.argtypes = [vp]
        L.ref_rs_forward.restype = C.c_int
        L.ref_rs_forward.argtypes = ([vp, C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp,
                                      C.c_float, vp, vp, vp, vp, vp, vp, vp, C.c_float, C.c_float, C.c_int, vp]
                                     + [vp] * 6)
        L.ref_rs_views.argtypes = [vp] + [C.POINTER(vp)] * 11
        L.ref_rs_backward.restype = C.c_int
        L.ref_rs_backward.argtypes = ([vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_float, vp, vp, vp, vp, vp, vp,
                                       vp, vp, C.c_float, C.c_float, vp] + [vp] * 4 + [vp] * 12 + [vp])
        L.ref_rs_mark_visible.argtypes = [C.c_int, vp, vp, vp, vp]
        L.ref_rs_copy.argtypes = [vp, vp, C.c_size_t]
        self.L = L
        self.h = L.ref_rs_create()

    def __del__(self):
        try:
            self.L.ref_rs_destroy(self.h)
        except Exception:
            pass

    def forward(self, scene: dict, cam_mats, W, H, tanx, tany, bg, prcppoint, patchbbox, config, mod=1.0, sh=None,
                degree=0, campos=None, want_views=True):
        torch.cuda.synchronize()
        view, proj, proj_raw = cam_mats
        means = scene["means3D"]
        dev = means.device
        P = means.shape[0]
        M = 0 if sh is None else sh.shape[1]
        z = lambda *s: torch.zeros(*s, device=dev)
        out_color, out_normal, out_depth, out_opac = z(3, H, W), z(3, H, W), z(1, H, W), z(1, H, W)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        contributions = z(P)
        colors = scene.get("colors_precomp") if sh is None else None
        self._keep = (scene, cam_mats, bg, sh, campos, prcppoint, patchbbox, config)
        R = self.L.ref_rs_forward(self.h, P, degree, M, _p(bg), W, H, _p(means), _p(sh), _p(colors),
                                  _p(scene["opacities"]), _p(scene["scales"]), mod, _p(scene["rotations"]), None,
                                  _p(view), _p(proj), _p(prcppoint), _p(patchbbox), _p(campos), tanx, tany, 0,
                                  _p(config), _p(out_color), _p(out_normal), _p(out_depth), _p(out_opac), _p(radii),
                                  _p(contributions))
        if R < 0:
            raise RuntimeError("reference surfel forward failed")
        self.P, self.W, self.H, self.R = P, W, H, R
        res = dict(num_rendered=R, color=out_color, normal=out_normal, depth=out_depth, opac=out_opac, radii=radii,
                   contributions=contributions)
        if want_views:
            res.update(self.views())
        return res

    def views(self) -> dict:
        ptrs = [vp() for _ in range(11)]
        self.L.ref_rs_views(self.h, *[C.byref(p) for p in ptrs])
        P, W, H, R = self.P, self.W, self.H, self.R
        T = ((W + 15) // 16) * ((H + 15) // 16)

        def copy(p, n, dtype):
            t = torch.empty(n, dtype=dtype, device="cuda")
            if n and p.value:
                self.L.ref_rs_copy(vp(t.data_ptr()), p, C.c_size_t(n * t.element_size()))
            return t

        return dict(
            point_list=copy(ptrs[0], R, torch.int32),
            keys=copy(ptrs[1], R, torch.int64),
            ranges=copy(ptrs[2], 2 * T, torch.int32).view(T, 2),
            n_contrib=copy(ptrs[3], W * H, torch.int32),
            final_T=copy(ptrs[4], W * H, torch.float32),
            means2D=copy(ptrs[5], 2 * P, torch.float32).view(P, 2),
            depths=copy(ptrs[6], P, torch.float32),
            conic_opacity=copy(ptrs[7], 4 * P, torch.float32).view(P, 4),
            tiles_touched=copy(ptrs[8], P, torch.int32),
            gnormal=copy(ptrs[9], 3 * P, torch.float32).view(P, 3),
            Jinv=copy(ptrs[10], 10 * P, torch.float32).view(P, 10),
        )

    def backward(self, scene: dict, cam_mats, tanx, tany, bg, prcppoint, patchbbox, config, radii, dL_dcolor,
                 dL_dnormal, dL_ddepth, dL_dopac, mod=1.0, sh=None, degree=0, campos=None):
        view, proj, proj_raw = cam_mats
        means = scene["means3D"]
        dev = means.device
        P = means.shape[0]
        M = 0 if sh is None else sh.shape[1]
        z = lambda *s: torch.zeros(*s, device=dev)
        g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), normal=z(P, 3), depths=z(P, 1),
                 means3D=z(P, 3), cov3D=z(P, 6), sh=z(P, max(M, 1), 3), scales=z(P, 3), rotations=z(P, 4), tau=z(P, 6))
        colors = scene.get("colors_precomp") if sh is None else None
        rc = self.L.ref_rs_backward(self.h, degree, M, _p(bg), _p(means), _p(sh), _p(colors), _p(scene["scales"]), mod,
                                    _p(scene["rotations"]), None, _p(view), _p(proj), _p(proj_raw), _p(campos),
                                    _p(prcppoint), _p(patchbbox), tanx, tany, _p(radii), _p(dL_dcolor), _p(dL_dnormal),
                                    _p(dL_ddepth), _p(dL_dopac), _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]),
                                    _p(g["colors"]), _p(g["normal"]), _p(g["depths"]), _p(g["means3D"]), _p(g["cov3D"]),
                                    _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]), _p(g["tau"]), _p(config))
        if rc != 0:
            raise RuntimeError("reference surfel backward failed")
        return g

    def mark_visible(self, means, view, proj):
        present = torch.zeros(means.shape[0], dtype=torch.bool, device=means.device)
        self.L.ref_rs_mark_visible(means.shape[0], _p(means), _p(view), _p(proj), _p(present))
        return present


def ref_ssim_module():
    """The unmodified reference fused-ssim extension (oracle/_ref/fused_ssim_cuda_ref.so), or None."""
    path = os.path.join(REF_DIR, "fused_ssim_cuda_ref.so")
    if not os.path.exists(path):
        return None
    if "fused_ssim_cuda_ref" not in _libs:
        import importlib.util
        spec = importlib.util.spec_from_file_location("fused_ssim_cuda_ref", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _libs["fused_ssim_cuda_ref"] = mod
    return _libs["fused_ssim_cuda_ref"]
