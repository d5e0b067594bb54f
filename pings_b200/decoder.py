"""Mirror of the reference's Decoder (model/decoder.py:14-135): same constructor arguments, same
parameter layout (`layers.{i}.weight/bias`, `lout.weight/bias`, so reference state_dicts load
unchanged), same methods — with the MLP evaluated by the fused sm_100a kernels behind the C-ABI
(pg_mlp_forward / pg_mlp_backward) instead of a cuBLAS sgemm + activation kernel per layer.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib as L


class _MlpFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dec, x, *params):
        lib = L.lib()
        if not x.is_cuda:
            raise RuntimeError("pings_b200.Decoder: expected CUDA tensors (there is no CPU fallback)")
        lead = x.shape[:-1]
        x2 = x.detach()
        if x2.dim() != 2:
            x2 = x2.reshape(-1, x.shape[-1])
        if x2.dtype is not torch.float32 or not x2.is_contiguous():
            x2 = x2.to(torch.float32).contiguous()
        N, dev = x2.shape[0], x2.device
        mlp, keep = dec._mlp_struct()
        y = torch.empty(N, dec.mlp_out_dim, dtype=torch.float32, device=dev)
        need = any(ctx.needs_input_grad[1:])
        hs = torch.empty(dec.hidden_level, N, dec.hidden_dim, dtype=torch.float32, device=dev) if need else None
        with L.device_guard(dev):
            L.check(lib.pg_mlp_forward(C.byref(mlp), N, x2.data_ptr(), y.data_ptr(), L.ptr(hs),
                                       L.raw_stream(dev)), "mlp forward")
        if need:
            ctx.dec, ctx.lead = dec, lead
            # the weights used by THIS forward go through save_for_backward: an optimizer step between forward
            # and backward then trips autograd's version check instead of silently differentiating against
            # different weights (eager torch in the reference raises the same error)
            ctx.save_for_backward(x2, hs, *params)
        return y.reshape(*lead, dec.mlp_out_dim)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy):
        # once_differentiable: the kernels are opaque to autograd, so a create_graph=True pass through this node
        # (the reference's analytic eikonal path, utils/tools.py:409-415 with numerical_grad_on=False) raises
        # instead of silently dropping the second-order term; the default numerical-gradient path is unaffected.
        lib = L.lib()
        dec = ctx.dec
        x2, hs, *saved = ctx.saved_tensors
        N, dev = x2.shape[0], x2.device
        dy2 = dy.reshape(N, dec.mlp_out_dim)
        if dy2.dtype is not torch.float32 or not dy2.is_contiguous():
            dy2 = dy2.to(torch.float32).contiguous()
        mlp, keep = dec._mlp_struct(saved)
        # the kernel ACCUMULATES dW / dB: one flat zero-filled buffer (one allocation + one fill per call instead of one
        # per tensor), returned to autograd as views in _params() order
        shapes = dec._grad_layout()
        flat = torch.zeros(shapes[-1][2], dtype=torch.float32, device=dev)
        grads = [flat[o:e].view(shp) for shp, o, e in shapes[:-1]]
        base = flat.data_ptr()
        nl = len(dec.layers) + 1
        pw, pb = (L.fp * nl)(), (L.fp * nl)()
        k = 0
        for i in range(nl):
            pw[i] = base + 4 * shapes[k][1]
            k += 1
            if dec._has_bias:
                pb[i] = base + 4 * shapes[k][1]
                k += 1
        dX = torch.empty_like(x2) if ctx.needs_input_grad[1] else None
        with L.device_guard(dev):
            L.check(lib.pg_mlp_backward(C.byref(mlp), N, x2.data_ptr(), hs.data_ptr(), dy2.data_ptr(), L.ptr(dX), pw, pb,
                                        L.raw_stream(dev)), "mlp backward")
        return (None, None if dX is None else dX.reshape(*ctx.lead, x2.shape[1]), *grads)


class _MultiMlpFn(torch.autograd.Function):
    """Several decoders on ONE shared input as one autograd node (spawn_gaussians runs its xyz / rot / scale / alpha
    heads on the same gathered feature matrix, GR:609-676): the same kernels as _MlpFn per head, but one
    Function.apply and one backward node instead of one per head - the C4 step is bound by exactly that host work."""

    @staticmethod
    def forward(ctx, decs, x, *params):
        lib = L.lib()
        if not x.is_cuda:
            raise RuntimeError("pings_b200.Decoder: expected CUDA tensors (there is no CPU fallback)")
        lead = x.shape[:-1]
        x2 = x.detach()
        if x2.dim() != 2:
            x2 = x2.reshape(-1, x.shape[-1])
        if x2.dtype is not torch.float32 or not x2.is_contiguous():
            x2 = x2.to(torch.float32).contiguous()
        N, dev = x2.shape[0], x2.device
        need = any(ctx.needs_input_grad[1:])
        ys, hss = [], []
        with L.device_guard(dev):
            st = L.raw_stream(dev)
            for dec in decs:
                mlp, keep = dec._mlp_struct()
                y = torch.empty(N, dec.mlp_out_dim, dtype=torch.float32, device=dev)
                hs = torch.empty(dec.hidden_level, N, dec.hidden_dim, dtype=torch.float32, device=dev) if need else None
                L.check(lib.pg_mlp_forward(C.byref(mlp), N, x2.data_ptr(), y.data_ptr(), L.ptr(hs), st), "mlp forward")
                ys.append(y.reshape(*lead, dec.mlp_out_dim))
                hss.append(hs)
        if need:
            ctx.decs, ctx.lead = decs, lead
            ctx.save_for_backward(x2, *hss, *params)
        return tuple(ys)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, *dys):
        lib = L.lib()
        decs = ctx.decs
        x2, *rest = ctx.saved_tensors
        hss, saved = rest[:len(decs)], rest[len(decs):]
        N, dev = x2.shape[0], x2.device
        want_dx = ctx.needs_input_grad[1]
        dX_total = None
        grads = []
        k = 0
        with L.device_guard(dev):
            st = L.raw_stream(dev)
            for dec, hs, dy in zip(decs, hss, dys):
                np_ = len(dec._params())
                mine = saved[k:k + np_]
                k += np_
                if dy is None:
                    grads += [None] * np_
                    continue
                dy2 = dy.reshape(N, dec.mlp_out_dim)
                if dy2.dtype is not torch.float32 or not dy2.is_contiguous():
                    dy2 = dy2.to(torch.float32).contiguous()
                mlp, keep = dec._mlp_struct(mine)
                shapes = dec._grad_layout()
                flat = torch.zeros(shapes[-1][2], dtype=torch.float32, device=dev)
                base = flat.data_ptr()
                nl = len(dec.layers) + 1
                pw, pb = (L.fp * nl)(), (L.fp * nl)()
                j = 0
                for i in range(nl):
                    pw[i] = base + 4 * shapes[j][1]
                    j += 1
                    if dec._has_bias:
                        pb[i] = base + 4 * shapes[j][1]
                        j += 1
                dX = torch.empty_like(x2) if want_dx else None
                L.check(lib.pg_mlp_backward(C.byref(mlp), N, x2.data_ptr(), hs.data_ptr(), dy2.data_ptr(), L.ptr(dX), pw, pb,
                                            st), "mlp backward")
                grads += [flat[o:e].view(shp) for shp, o, e in shapes[:-1]]
                if dX is not None:
                    dX_total = dX if dX_total is None else dX_total.add_(dX)
        return (None, None if dX_total is None else dX_total.reshape(*ctx.lead, x2.shape[1]), *grads)


def mlp_multi(decs, x):
    """[dec.mlp(x) for dec in decs] as one autograd node (all decoders take the same input)."""
    decs = tuple(decs)
    params = [p for d in decs for p in d._params()]
    return list(_MultiMlpFn.apply(decs, x, *params))


MLP_AUTO, MLP_SIMT, MLP_TCGEN05 = 0, 1, 2


def mlp_forward_impl(dec, x, impl: int = MLP_AUTO, save_hidden: bool = False):
    """Forward only, with an explicit kernel choice (pg_mlp_forward_impl): returns (y, hidden|None)."""
    lib = L.lib()
    if not x.is_cuda:
        raise RuntimeError("pings_b200.Decoder: expected CUDA tensors (there is no CPU fallback)")
    x2 = x.detach().reshape(-1, x.shape[-1]).to(torch.float32).contiguous()
    N, dev = x2.shape[0], x2.device
    mlp, keep = dec._mlp_struct()
    y = torch.empty(N, dec.mlp_out_dim, dtype=torch.float32, device=dev)
    hs = torch.empty(dec.hidden_level, N, dec.hidden_dim, dtype=torch.float32, device=dev) if save_hidden else None
    with L.device_guard(dev):
        L.check(lib.pg_mlp_forward_impl(C.byref(mlp), N, x2.data_ptr(), y.data_ptr(), L.ptr(hs), impl,
                                        L.raw_stream(dev)), "mlp forward")
    return y, hs


class Decoder(nn.Module):
    def __init__(self, config, input_feature_dim, hidden_dim, hidden_level, out_dim, out_k=1, pos_dim=3,
                 is_time_conditioned=False):
        super().__init__()
        self.mlp_out_dim = out_dim * out_k
        self.out_dim_per = out_dim
        self.out_k = out_k
        self.use_leaky_relu = config.mlp_leaky_relu
        self.bs = config.infer_bs
        self.hidden_dim, self.hidden_level = hidden_dim, hidden_level
        bias_on = config.mlp_bias_on
        input_dim = input_feature_dim + pos_dim + (1 if is_time_conditioned else 0)
        self.input_dim = input_dim
        self.layers = nn.ModuleList([nn.Linear(input_dim if i == 0 else hidden_dim, hidden_dim, bias_on)
                                     for i in range(hidden_level)])
        self.lout = nn.Linear(hidden_dim, self.mlp_out_dim, bias_on)
        self.sdf_scale = 1.0
        if config.main_loss_type == "bce":
            self.sdf_scale = config.logistic_gaussian_ratio * config.sigma_sigmoid_m
        self.to(config.device)

    def _linears(self):
        return list(self.layers) + [self.lout]

    def _mlp_struct(self, params=None):
        """PgMlp over the live parameters, or over `params` (weights/biases in _params() order, e.g. the tensors
        a forward saved for its backward).  The ctypes struct is cached per tuple of data pointers (the parameters
        of a training loop keep their storage: FlatAdamW updates them in place), so a call costs a few
        `data_ptr()`s instead of rebuilding the struct (28 us, 80 times per C4 step)."""
        src = list(params) if params is not None else self._params()
        ts = [t if t.is_contiguous() else t.detach().contiguous() for t in src]
        key = tuple(t.data_ptr() for t in ts)
        hit = self.__dict__.get("_struct_cache")
        if hit is not None and hit[0] == key:
            return hit[1], ts
        m = L.PgMlp()
        m.in_dim, m.hidden_dim, m.hidden_level, m.out_dim = self.input_dim, self.hidden_dim, self.hidden_level, self.mlp_out_dim
        m.leaky_relu = int(bool(self.use_leaky_relu))
        it = iter(ts)
        for i, lin in enumerate(self._linears()):
            m.weights[i] = next(it).data_ptr()
            if lin.bias is not None:
                m.biases[i] = next(it).data_ptr()
        self.__dict__["_struct_cache"] = (key, m)
        return m, ts

    @property
    def _has_bias(self) -> bool:
        return self.lout.bias is not None

    def _grad_layout(self):
        """[(shape, begin, end) per parameter in _params() order] + [(None, 0, total)] in elements of one flat buffer."""
        lay = self.__dict__.get("_grad_layout_cache")
        if lay is None:
            lay, off = [], 0
            for p in self._params():
                lay.append((tuple(p.shape), off, off + p.numel()))
                off += p.numel()
            lay.append((None, 0, off))
            self.__dict__["_grad_layout_cache"] = lay
        return lay

    def _params(self):
        ps = []
        for lin in self._linears():
            ps.append(lin.weight)
            if lin.bias is not None:
                ps.append(lin.bias)
        return ps

    def mlp(self, features):
        """DEC:62-82."""
        return _MlpFn.apply(self, features, *self._params())

    def mlp_batch(self, features):
        """DEC:84-98. One fused launch covers any row count, so no infer_bs chunking is needed."""
        return self.mlp(features)

    def sdf(self, features):
        return self.mlp(features).squeeze(1) * self.sdf_scale

    def occupancy(self, features):
        return torch.sigmoid(self.sdf(features) / -self.sdf_scale)

    def sem_label_prob(self, features):
        return F.log_softmax(self.mlp(features), dim=-1)

    def sem_label(self, features):
        return torch.argmax(self.sem_label_prob(features), dim=1)

    def regress_color(self, features):
        return torch.sigmoid(self.mlp(features))
