"""Mirror of the grid-query half of the reference's Mesher (utils/mesher.py:40-166):
`query_points` evaluates the SDF (one fused kNN -> IDW -> decoder kernel per batch), the colour head and the
marching-cubes mask on arbitrary query coordinates.  Marching cubes itself (skimage) and the Open3D mesh assembly
stay with the caller."""
from __future__ import annotations

import math

import numpy as np
import torch


class Mesher:
    def __init__(self, config, neural_points, decoders):
        self.config = config
        self.neural_points = neural_points
        self.sdf_mlp = decoders.get("sdf")
        self.sem_mlp = decoders.get("semantic")
        self.color_mlp = decoders.get("color")
        self.device = torch.device(config.device)

    def query_points(self, coord, bs, query_sdf=True, query_sem=False, query_color=False, query_mask=True,
                     query_locally=False, mask_min_nn_count: int = 4, out_torch: bool = False):
        """Same arguments and return tuple (sdf_pred, sem_pred, color_pred, mc_mask) as utils/mesher.py:40-166."""
        n = coord.shape[0]
        dev = self.device
        npm = self.neural_points
        wf = bool(self.config.weighted_first)
        sdf_pred = torch.zeros(n, device=dev) if query_sdf else None
        sem_pred = torch.zeros(n, device=dev) if query_sem else None
        color_pred = torch.zeros((n, getattr(self.config, "color_channel", 3)), device=dev) if query_color else None
        mc_mask = torch.zeros(n, device=dev) if query_mask else None
        with torch.no_grad():
            for b in range(math.ceil(n / bs)):
                head, tail = b * bs, min((b + 1) * bs, n)
                q = coord[head:tail].to(dev)
                nn_count = None
                if query_sdf:
                    # the reference leaves rows without any neighbour at 0 (utils/mesher.py:118-131); the fused kernel
                    # evaluates the decoder on the all-zero feature row instead, so those rows are zeroed here
                    sdf, _, nn_count = npm.sdf_fused(q, self.sdf_mlp, query_locally=query_locally, use_only_valid_points=True)
                    sdf_pred[head:tail] = torch.where(nn_count >= 1, sdf, torch.zeros_like(sdf))
                if query_sem or query_color:
                    geo, col, w, nn_count, _ = npm.query_feature(
                        q, accumulate_stability=False, query_locally=query_locally, query_geo_feature=query_sem,
                        query_color_feature=query_color, use_only_valid_points=True)
                    if query_sem:
                        prob = self.sem_mlp.sem_label_prob(geo)
                        if not wf:
                            prob = torch.sum(prob * w, dim=1)
                        sem_pred[head:tail] = torch.argmax(prob, dim=1).to(sem_pred.dtype)
                    if query_color:
                        c = self.color_mlp.regress_color(col)
                        if not wf:
                            c = torch.sum(c * w, dim=1)
                        color_pred[head:tail] = c
                if query_mask:
                    if nn_count is None:
                        _, _, _, nn_count, _ = npm.query_feature(q, accumulate_stability=False, query_locally=query_locally,
                                                                 use_only_valid_points=True)
                    mc_mask[head:tail] = (nn_count >= mask_min_nn_count).to(mc_mask.dtype)
        if out_torch:
            return sdf_pred, sem_pred, color_pred, mc_mask
        f = lambda t: None if t is None else t.cpu().numpy()
        c = None if color_pred is None else color_pred.cpu().numpy().astype(np.float64)
        return f(sdf_pred), f(sem_pred), c, f(mc_mask)
