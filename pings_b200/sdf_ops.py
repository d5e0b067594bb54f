"""The composed "SDF query" op of the reference mapper (utils/mapper.py:2273-2369):
`sdf`, `sdf_batch`, `get_numerical_gradient`, over pings_b200 NeuralPoints + Decoder.

* Under `torch.no_grad()` (meshing, evaluation, numerical gradients, tracking residuals) the query runs as
  ONE fused kernel (pg_sdf_query: kNN -> top-k -> IDW -> decoder -> reduction over k).
* With gradients enabled it runs query_feature + Decoder.sdf through their autograd Functions, exactly the
  reference's composition (`utils/mapper.py:2274-2287`).
"""
from __future__ import annotations

import torch


class SdfQuery:
    """Holds what Mapper.sdf reads: `neural_points`, `sdf_mlp`, `config.weighted_first`."""

    def __init__(self, neural_points, sdf_mlp, silence: bool = True):
        self.neural_points = neural_points
        self.sdf_mlp = sdf_mlp
        self.config = neural_points.config
        self.silence = silence

    def sdf(self, x, get_std=False, min_nn_count=1, accumulate_stability=False):
        """utils/mapper.py:2273-2289 -> (sdf_pred [N], sdf_std [N]|None, valid_mask [N] bool)."""
        npm = self.neural_points
        if not torch.is_grad_enabled() and not accumulate_stability:
            sdf_pred, sdf_std, nn_count = npm.sdf_fused(x, self.sdf_mlp, get_std=get_std)
            return sdf_pred, sdf_std, nn_count >= min_nn_count
        geo_feature, _, weight_knn, nn_count, _ = npm.query_feature(x, accumulate_stability=accumulate_stability)
        sdf_pred = self.sdf_mlp.sdf(geo_feature)
        sdf_std = None
        if not self.config.weighted_first:
            sdf_pred_mean = torch.sum(sdf_pred * weight_knn, dim=1)
            if get_std:
                sdf_var = torch.sum(weight_knn * (sdf_pred - sdf_pred_mean.unsqueeze(-1)) ** 2, dim=1)
                sdf_std = torch.sqrt(sdf_var).squeeze(1)
            sdf_pred = sdf_pred_mean.squeeze(1)
        return sdf_pred, sdf_std, nn_count >= min_nn_count

    def sdf_batch(self, x, bs, get_std=False, min_nn_count=1, accumulate_stability=False):
        """utils/mapper.py:2292-2316.  The fused kernel is persistent (a warp strides over the queries), so any
        row count is one launch; `bs` only bounds the chunking of the autograd path."""
        if not torch.is_grad_enabled() and not accumulate_stability:
            return self.sdf(x, get_std, min_nn_count, accumulate_stability)
        count = x.shape[0]
        sdf_pred = torch.zeros(count, dtype=torch.float32, device=x.device)
        sdf_std = torch.zeros(count, dtype=torch.float32, device=x.device) if get_std else None
        valid_mask = torch.ones(count, dtype=torch.bool, device=x.device)
        for head in range(0, count, bs):
            tail = min(head + bs, count)
            b_sdf, b_std, b_valid = self.sdf(x[head:tail], get_std, min_nn_count, accumulate_stability)
            sdf_pred[head:tail] = b_sdf
            if b_std is not None and sdf_std is not None:
                sdf_std[head:tail] = b_std
            valid_mask[head:tail] = b_valid
        return sdf_pred, sdf_std, valid_mask

    def get_numerical_gradient(self, x, sdf_x=None, eps=0.02, two_side=True):
        """utils/mapper.py:2319-2369: fixed-step finite differences of the SDF, [N,3].  All 6N (3N) offset
        queries go through one launch."""
        N = x.shape[0]
        offs = torch.eye(3, dtype=x.dtype, device=x.device) * eps                      # [3,3]
        if two_side:
            pts = torch.cat([x + offs[0], x - offs[0], x + offs[1], x - offs[1], x + offs[2], x - offs[2]], dim=0)
            s = self.sdf(pts)[0]
            return torch.stack([(s[0:N] - s[N:2 * N]), (s[2 * N:3 * N] - s[3 * N:4 * N]),
                                (s[4 * N:5 * N] - s[5 * N:6 * N])], dim=1) / (2 * eps)
        pts = torch.cat([x + offs[0], x + offs[1], x + offs[2]], dim=0)
        s = self.sdf(pts)[0]
        return torch.stack([s[0:N] - sdf_x, s[N:2 * N] - sdf_x, s[2 * N:3 * N] - sdf_x], dim=1) / eps
