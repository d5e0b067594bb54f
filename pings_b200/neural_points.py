"""Host-side mirror of the reference's neural-point map interface for the query path
(model/neural_gaussians.py, class NeuralPoints): same attribute names, method names, argument
meaning and return tuples, so callers written against the reference (utils/mapper.py:2273-2289,
utils/tracker.py:282-320, utils/mesher.py:40-166) keep working.

What runs where:
  * radius_neighborhood_search / query_feature / query_certainty / sdf  -> hand-written sm_100a
    kernels behind the C-ABI (pings_b200/csrc/pg_neural.cuh, pg_sdf.cuh). No CPU fallback: CPU
    tensors raise.
  * update / reset_local_map / assign_local_to_global (map maintenance) are
    plain torch tensor ops here; they only define the state the kernels read.
"""
from __future__ import annotations

import ctypes as C
from types import SimpleNamespace

import torch
import torch.nn as nn

from . import _lib as L

PRIMES = (73856093, 19349669, 83492791)  # NP:80-82


def _need_cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise RuntimeError(f"pings_b200.{what}: expected a CUDA tensor (there is no CPU fallback)")


def default_config(**kw) -> SimpleNamespace:
    """The reference's Config fields this path reads (utils/config.py:94-105,118,390 + run_pin_slam.yaml)."""
    c = SimpleNamespace(device="cuda", dtype=torch.float32, voxel_size_m=0.4, num_nei_cells=2, search_alpha=0.5,
                        query_nn_k=6, buffer_size=int(1e8), feature_dim=8, color_feature_dim=8, feature_std=0.0,
                        color_on=False, weighted_first=True, layer_norm_on=False, local_map_radius=84.0,
                        sorrounding_map_radius=84.0 * 1.4, local_map_travel_dist_ratio=5.0, diff_ts_local=400.0,
                        track_on=True, temporal_local_map_off=False, use_mid_ts=False,
                        range_filter_2d=True,   # utils/config.py:60
                        infer_bs=80000, mlp_bias_on=True, mlp_leaky_relu=False, main_loss_type="bce",
                        logistic_gaussian_ratio=0.55, sigma_sigmoid_m=0.1)
    for k, v in kw.items():
        setattr(c, k, v)
    return c


class _QueryFeature(torch.autograd.Function):
    """Autograd node of query_feature: forward = fused kNN/top-k/IDW kernel, backward = one kernel that
    scatters into the feature tables and differentiates through the neighbour vectors and IDW weights."""

    @staticmethod
    def forward(ctx, npm, query_points, geo_table, color_table, flags):
        out = npm._query_forward(query_points, geo_table, color_table, flags)
        geo_vec, color_vec, weight, nn_counts, certainty, knn_idx, knn_d = out
        ctx.npm, ctx.flags = npm, flags
        # the tables and the point/orientation arrays read by THIS forward are saved, so the backward differentiates
        # against them even if reset_local_map() / assign_local_to_global() re-binds the attributes in between, and
        # an in-place edit trips autograd's version check (as eager torch would in the reference)
        local = flags["query_locally"]
        pts = npm.local_neural_points if local else npm.neural_points
        quat = npm.local_point_orientations if local else npm.point_orientations
        e0 = query_points.new_empty(0)
        ctx.save_for_backward(query_points, knn_idx, knn_d, weight, nn_counts, pts, quat,
                              geo_table if geo_table is not None else e0, color_table if color_table is not None else e0)
        ctx.shapes = (None if geo_table is None else geo_table.shape, None if color_table is None else color_table.shape)
        ctx.mark_non_differentiable(nn_counts, certainty, knn_idx)
        e = query_points.new_empty(0)
        return (geo_vec if geo_vec is not None else e, color_vec if color_vec is not None else e, weight, nn_counts,
                certainty, knn_idx)

    @staticmethod
    @torch.autograd.function.once_differentiable   # opaque kernels: a create_graph=True pass raises (see decoder._MlpFn)
    def backward(ctx, d_geo, d_color, d_weight, *_):
        query_points, knn_idx, knn_d, weight, nn_counts, pts, quat, geo_t, color_t = ctx.saved_tensors
        gshape, cshape = ctx.shapes
        dq, dg, dc = ctx.npm._query_backward(query_points, knn_idx, knn_d, weight, nn_counts,
                                             d_geo if gshape is not None else None,
                                             d_color if cshape is not None else None, d_weight, gshape, cshape,
                                             ctx.flags, ctx.needs_input_grad[1],
                                             saved=(pts, quat, geo_t if gshape is not None else None,
                                                    color_t if cshape is not None else None))
        return None, dq, dg, dc, None


class NeuralPoints(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.config = config
        self.geo_feature_dim = config.feature_dim
        self.color_feature_dim = config.color_feature_dim
        self.geo_feature_std = config.feature_std
        self.color_feature_std = config.feature_std
        self.device = torch.device(config.device)
        self.dtype = config.dtype
        self.idx_dtype = torch.int64
        self.resolution = config.voxel_size_m
        self.buffer_size = int(config.buffer_size)
        self.sorrounding_map_radius = config.sorrounding_map_radius
        self.temporal_local_map_on = not ((not config.track_on) and config.temporal_local_map_off)
        self.diff_travel_dist_local = config.local_map_radius * config.local_map_travel_dist_ratio
        self.cur_ts = 0
        self.max_ts = 0
        self.travel_dist = None
        self.after_pgo = False
        dev, dt = self.device, self.dtype
        self.primes = torch.tensor(PRIMES, dtype=torch.int64, device=dev)
        self.buffer_pt_index = torch.full((self.buffer_size,), -1, dtype=torch.int64, device=dev)
        self.neural_points = torch.empty((0, 3), dtype=dt, device=dev)
        self.point_orientations = torch.empty((0, 4), dtype=dt, device=dev)
        self.geo_features = torch.empty((1, self.geo_feature_dim), dtype=dt, device=dev)
        self.color_on = bool(config.color_on)
        self.color_features = torch.empty((1, self.color_feature_dim), dtype=dt, device=dev) if self.color_on else None
        self.point_colors = torch.empty((0, 3), dtype=dt, device=dev) if self.color_on else None
        self.point_ts_create = torch.empty(0, dtype=torch.int32, device=dev)
        self.point_ts_update = torch.empty(0, dtype=torch.int32, device=dev)
        self.point_certainties = torch.empty(0, dtype=dt, device=dev)
        self.valid_color_mask = torch.empty(0, dtype=torch.bool, device=dev)
        self.valid_gs_mask = torch.empty(0, dtype=torch.bool, device=dev)
        self.free_gs_mask = torch.empty(0, dtype=torch.bool, device=dev)
        self.local_neural_points = torch.empty((0, 3), dtype=dt, device=dev)
        self.local_point_orientations = torch.empty((0, 4), dtype=dt, device=dev)
        self.local_geo_features = nn.Parameter()
        self.local_color_features = nn.Parameter()
        self.local_point_certainties = torch.empty(0, dtype=dt, device=dev)
        self.local_point_ts_update = torch.empty(0, dtype=torch.int32, device=dev)
        self.local_mask = None
        self.global2local = None
        self.sorrounding_mask = None
        self.local_valid_gs_mask = torch.empty(0, dtype=torch.bool, device=dev)
        self.local_free_gs_mask = torch.empty(0, dtype=torch.bool, device=dev)
        self.set_search_neighborhood(config.num_nei_cells, config.search_alpha)

    # ------------------------------------------------------------------ state helpers
    def is_empty(self):
        return self.neural_points.shape[0] == 0

    def count(self):
        return self.neural_points.shape[0]

    def local_count(self):
        return self.local_neural_points.shape[0]

    @classmethod
    def from_reference(cls, ref) -> "NeuralPoints":
        """Wrap a live reference `NeuralPoints` (model/neural_gaussians.py:20-120) without copying:
        every tensor attribute (slot table, points, features, masks, local map) is shared, so the
        reference's own `update` / `reset_local_map` keep maintaining the map while queries run here."""
        self = cls.__new__(cls)
        nn.Module.__init__(self)
        for k, v in vars(ref).items():
            if k.startswith("_"):
                continue
            setattr(self, k, v)
        for k, v in ref._parameters.items():   # local_geo_features / local_color_features
            setattr(self, k, v)
        self.device = torch.device(ref.config.device)
        if not hasattr(self, "neighbor_dx"):
            self.set_search_neighborhood(ref.config.num_nei_cells, ref.config.search_alpha)
        return self

    @classmethod
    def from_state(cls, config, state: dict) -> "NeuralPoints":
        """Adopt tensors saved from a reference NeuralPoints (fixtures, pin_map.pth contents).
        `state["hash_slots"], state["hash_values"]` hold the occupied slot-table entries."""
        self = cls(config)
        dev = self.device
        for k, v in state.items():
            if k in ("hash_slots", "hash_values") or v is None:
                continue
            v = v.to(dev) if isinstance(v, torch.Tensor) else v
            if k in ("local_geo_features", "local_color_features"):
                setattr(self, k, nn.Parameter(v.clone()))
            else:
                setattr(self, k, v)
        self.buffer_pt_index[state["hash_slots"].to(dev)] = state["hash_values"].to(dev)
        return self

    def set_search_neighborhood(self, num_nei_cells: int = 1, search_alpha: float = 1.0):
        """NP:1026-1059."""
        d = torch.arange(-num_nei_cells, num_nei_cells + 1, device=self.device, dtype=torch.int64)
        dx = torch.stack(torch.meshgrid(d, d, d, indexing="ij"), dim=-1).reshape(-1, 3)
        self.neighbor_dx = dx[(dx ** 2).sum(-1) < (num_nei_cells + search_alpha) ** 2].contiguous()
        self.neighbor_K = self.neighbor_dx.shape[0]
        self.max_valid_dist2 = 3 * ((num_nei_cells + 1) * self.resolution) ** 2

    # ------------------------------------------------------------------ C-ABI plumbing
    def _slot_mirror(self):
        """int32 mirror of the int64 slot table (the reference layout, NP:86-88, is kept as the source of truth so a
        live reference map can be shared): 4-byte probes, half the footprint (400 MB at buffer_size 1e8).  Rebuilt
        when the table tensor was replaced or modified in place (torch's version counter); `use_slot_mirror = False`
        turns it off."""
        if not getattr(self, "use_slot_mirror", True) or not self.buffer_pt_index.is_cuda:
            return None
        key = (self.buffer_pt_index.data_ptr(), self.buffer_pt_index._version, int(self.buffer_size))
        cache = getattr(self, "_mirror_cache", None)
        if cache is None or cache[0] != key:
            t32 = torch.empty(int(self.buffer_size), dtype=torch.int32, device=self.buffer_pt_index.device)
            with L.device_guard(t32.device):
                L.check(L.lib().pg_np_build_slot_mirror(int(self.buffer_size), self.buffer_pt_index.data_ptr(), t32.data_ptr(),
                                                        L.raw_stream(t32.device)), "slot mirror")
            self._mirror_cache = cache = (key, t32)
        return cache[1].data_ptr()

    def _map_struct(self, time_filtering: bool):
        m = L.PgNeuralMap()
        self._keep = [self.buffer_pt_index, self.neural_points, self.neighbor_dx]
        m.buffer_size = int(self.buffer_size)
        m.buffer_pt_index = self.buffer_pt_index.data_ptr()
        m.buffer_pt_index32 = self._slot_mirror()
        m.neural_points = self.neural_points.data_ptr()
        m.n_points = self.neural_points.shape[0]
        m.neighbor_dx = self.neighbor_dx.data_ptr()
        m.K = int(self.neighbor_K)
        for i in range(3):
            m.primes[i] = PRIMES[i]
        m.resolution = float(self.resolution)
        m.max_valid_dist2 = float(self.max_valid_dist2)
        if time_filtering:
            td = self.travel_dist.to(self.device, torch.float32).contiguous()
            self._keep.append(td)
            m.point_ts_create = self.point_ts_create.data_ptr()
            m.travel_dist = td.data_ptr()
            m.cur_ts = int(self.cur_ts)
            m.diff_travel_dist_local = float(self.diff_travel_dist_local)
        return m

    def _query_struct(self, flags, geo_table, color_table):
        a = L.PgQueryArgs()
        local = flags["query_locally"]
        a.nn_k = int(self.config.query_nn_k)
        a.time_filtering = int(flags["time_filtering"])
        a.query_locally = int(local)
        a.use_only_measured = int(flags["use_only_measured_points"])
        a.use_only_valid = int(flags["use_only_valid_points"])
        a.weighted_first = int(self.config.weighted_first)
        a.after_pgo = int(self.after_pgo)
        a.F_geo = 0 if geo_table is None else geo_table.shape[1]
        a.F_color = 0 if color_table is None else color_table.shape[1]
        self._fm = self.free_gs_mask.view(torch.uint8)
        self._vm = self.valid_gs_mask.view(torch.uint8)
        a.free_gs_mask = L.ptr(self._fm)
        a.valid_gs_mask = L.ptr(self._vm)
        a.global2local = L.ptr(self.global2local) if local else None
        pts = self.local_neural_points if local else self.neural_points
        quat = self.local_point_orientations if local else self.point_orientations
        cert = self.local_point_certainties if local else self.point_certainties
        a.points, a.orientations, a.certainties = L.ptr(pts), L.ptr(quat), L.ptr(cert)
        a.geo_features = L.ptr(geo_table)
        a.color_features = L.ptr(color_table)
        return a

    def _query_forward(self, q, geo_table, color_table, flags):
        lib = L.lib()
        _need_cuda(q, "query_feature")
        if self.config.layer_norm_on:
            raise NotImplementedError("layer_norm_on=True is not supported by the fused query kernel")
        q = q.detach().to(torch.float32).contiguous()
        N, k = q.shape[0], int(self.config.query_nn_k)
        dev = q.device
        wf = bool(self.config.weighted_first)
        mk = lambda F: torch.empty((N, F + 3) if wf else (N, k, F + 3), dtype=torch.float32, device=dev)
        geo_vec = mk(geo_table.shape[1]) if geo_table is not None else None
        color_vec = mk(color_table.shape[1]) if color_table is not None else None
        weight = torch.empty(N, k, dtype=torch.float32, device=dev)
        nn_counts = torch.empty(N, dtype=torch.int64, device=dev)
        certainty = torch.empty(N, dtype=torch.float32, device=dev)
        knn_idx = torch.empty(N, k, dtype=torch.int32, device=dev)
        knn_d = torch.empty(N, k, dtype=torch.float32, device=dev)
        m = self._map_struct(flags["time_filtering"])
        a = self._query_struct(flags, geo_table, color_table)
        o = L.PgQueryOut()
        o.geo_vec, o.color_vec, o.weight = L.ptr(geo_vec), L.ptr(color_vec), weight.data_ptr()
        o.nn_counts, o.certainty = nn_counts.data_ptr(), certainty.data_ptr()
        o.knn_idx, o.knn_dist2 = knn_idx.data_ptr(), knn_d.data_ptr()
        with L.device_guard(dev):
            L.check(lib.pg_np_query_feature(C.byref(m), C.byref(a), N, q.data_ptr(), C.byref(o),
                                            L.raw_stream(dev)), "query_feature")
        return geo_vec, color_vec, weight, nn_counts, certainty, knn_idx, knn_d

    def _query_backward(self, q, knn_idx, knn_d, weight, nn_counts, d_geo, d_color, d_weight, gshape, cshape, flags,
                        need_dq, saved=None):
        lib = L.lib()
        dev = q.device
        N = q.shape[0]
        qc = q.detach().to(torch.float32).contiguous()
        if saved is not None:
            pts, quat, geo_table, color_table = saved
            geo_table = geo_table.detach() if geo_table is not None else None
            color_table = color_table.detach() if color_table is not None else None
        else:
            pts = quat = None
            geo_table = self.local_geo_features if flags["query_locally"] else self.geo_features
            color_table = self.local_color_features if flags["query_locally"] else self.color_features
            geo_table = geo_table.detach() if gshape is not None else None
            color_table = color_table.detach() if cshape is not None else None
        a = self._query_struct(flags, geo_table, color_table)
        if pts is not None:   # the arrays the forward read (see _QueryFeature.forward)
            a.points, a.orientations = L.ptr(pts), L.ptr(quat)
        dq = torch.zeros(N, 3, dtype=torch.float32, device=dev) if need_dq else None
        dg = torch.zeros(gshape, dtype=torch.float32, device=dev) if gshape is not None else None
        dc = torch.zeros(cshape, dtype=torch.float32, device=dev) if cshape is not None else None
        c = lambda t: None if t is None else t.contiguous()
        d_geo, d_color, d_weight = c(d_geo), c(d_color), c(d_weight)
        with L.device_guard(dev):
            L.check(lib.pg_np_query_feature_backward(
                C.byref(a), N, qc.data_ptr(), knn_idx.data_ptr(), knn_d.data_ptr(), weight.data_ptr(),
                nn_counts.data_ptr(), L.ptr(d_geo), L.ptr(d_color), L.ptr(d_weight), L.ptr(dg), L.ptr(dc), L.ptr(dq),
                L.raw_stream(dev)), "query_feature backward")
        return dq, dg, dc

    # ------------------------------------------------------------------ public query API
    def radius_neighborhood_search(self, points: torch.Tensor, time_filtering: bool = False):
        """NP:1061-1115 -> (dist2 [N,K] f32, idx [N,K] i64)."""
        lib = L.lib()
        _need_cuda(points, "radius_neighborhood_search")
        q = points.detach().to(torch.float32).contiguous()
        N, K = q.shape[0], int(self.neighbor_K)
        dist2 = torch.empty(N, K, dtype=torch.float32, device=q.device)
        idx = torch.empty(N, K, dtype=torch.int64, device=q.device)
        m = self._map_struct(time_filtering)
        with L.device_guard(q.device):
            L.check(lib.pg_np_radius_search(C.byref(m), N, q.data_ptr(), int(bool(time_filtering)), dist2.data_ptr(),
                                            idx.data_ptr(), L.raw_stream(q.device)),
                    "radius_neighborhood_search")
        return dist2, idx

    def query_feature(self, query_points, query_ts=None, accumulate_stability=True, query_locally=True,
                      query_geo_feature=True, query_color_feature=False, use_only_measured_points=True,
                      use_only_valid_points=False):
        """NP:506-725 -> (geo_features_vector, color_features_vector, weight [N,k,1], nn_counts, certainty)."""
        if not query_geo_feature and not query_color_feature:
            raise SystemExit("you need to at least query one kind of feature")
        flags = dict(time_filtering=bool(self.temporal_local_map_on and query_locally), query_locally=bool(query_locally),
                     use_only_measured_points=bool(use_only_measured_points),
                     use_only_valid_points=bool(use_only_valid_points))
        geo_table = (self.local_geo_features if query_locally else self.geo_features) if query_geo_feature else None
        color_table = None
        if query_color_feature and self.color_features is not None:
            color_table = self.local_color_features if query_locally else self.color_features
        geo_vec, color_vec, weight, nn_counts, certainty, knn_idx = _QueryFeature.apply(
            self, query_points, geo_table, color_table, flags)
        if accumulate_stability:
            with torch.no_grad():
                cert = self.local_point_certainties if query_locally else self.point_certainties
                ts = None
                if query_locally and query_ts is not None:
                    ts = query_ts.to(torch.int32).contiguous()
                N, k = knn_idx.shape
                with L.device_guard(knn_idx.device):
                    L.check(L.lib().pg_np_accumulate_stability(
                        N, k, knn_idx.data_ptr(), weight.detach().data_ptr(), L.ptr(ts), cert.data_ptr(),
                        self.local_point_ts_update.data_ptr() if ts is not None else None,
                        L.raw_stream(knn_idx.device)), "accumulate_stability")
        return (geo_vec if geo_table is not None else None, color_vec if color_table is not None else None,
                weight.unsqueeze(-1), nn_counts, certainty)

    def query_certainty(self, query_points: torch.Tensor):
        """NP:1117-1133 (max certainty over the current search neighbourhood, global map)."""
        _, idx = self.radius_neighborhood_search(query_points)
        certainty = self.point_certainties[idx]
        certainty[idx < 0] = 0.0
        return torch.max(certainty, dim=-1)[0]

    @torch.no_grad()
    def sdf_fused(self, query_points, decoder, get_std=False, query_locally=True, use_only_measured_points=True,
                  use_only_valid_points=False):
        """Inference-only Mapper.sdf (utils/mapper.py:2273-2289) as ONE kernel: returns (sdf, std|None, nn_counts)."""
        lib = L.lib()
        _need_cuda(query_points, "sdf_fused")
        q = query_points.detach().to(torch.float32).contiguous()
        N, dev = q.shape[0], q.device
        flags = dict(time_filtering=bool(self.temporal_local_map_on and query_locally), query_locally=bool(query_locally),
                     use_only_measured_points=bool(use_only_measured_points),
                     use_only_valid_points=bool(use_only_valid_points))
        geo_table = (self.local_geo_features if query_locally else self.geo_features).detach()
        m = self._map_struct(flags["time_filtering"])
        a = self._query_struct(flags, geo_table, None)
        mlp, keep = decoder._mlp_struct()
        sdf = torch.empty(N, dtype=torch.float32, device=dev)
        want_std = get_std and not self.config.weighted_first
        std = torch.empty(N, dtype=torch.float32, device=dev) if want_std else None
        nn_counts = torch.empty(N, dtype=torch.int64, device=dev)
        with L.device_guard(dev):
            L.check(lib.pg_sdf_query(C.byref(m), C.byref(a), C.byref(mlp), float(decoder.sdf_scale), N, q.data_ptr(),
                                     sdf.data_ptr(), L.ptr(std), nn_counts.data_ptr(),
                                     L.raw_stream(dev)), "sdf_query")
        return sdf, std, nn_counts

    # ------------------------------------------------------------------ map maintenance (torch ops)
    def _hash(self, grid_coords: torch.Tensor) -> torch.Tensor:
        return torch.fmod((grid_coords * self.primes).sum(-1), int(self.buffer_size))

    def update(self, points, colors=None, normals=None, sensor_position=None, sensor_orientation=None, cur_ts=0,
               is_reliable=True):
        """NP:214-375: voxel-downsample the scan, insert unseen (or hash-colliding / out-of-window) voxels."""
        res = self.resolution
        sample_idx = voxel_down_sample(points, res)
        sp = points[sample_idx]
        sc = colors[sample_idx] if colors is not None else None
        hv = self._hash((sp / res).floor().to(torch.int64))
        hidx = self.buffer_pt_index[hv]
        if not self.is_empty():
            d2 = ((self.neural_points[hidx] - sp) ** 2).sum(-1)
            upd = (hidx == -1) | (d2 > 3 * res ** 2)
            if sc is not None:
                cm = (hidx > -1) & (self.valid_color_mask[hidx] == 0) & (sc[:, 0] >= 0.0)
                self.point_colors[hidx[cm]] = sc[cm]
                self.valid_color_mask[hidx[cm]] = True
            if self.temporal_local_map_on:
                dtd = self.travel_dist[cur_ts] - self.travel_dist[self.point_ts_update[hidx].long()]
                upd = upd | (dtd > self.diff_travel_dist_local)
        else:
            upd = torch.ones_like(hidx, dtype=torch.bool)
        added = sp[upd]
        n_new, n_old = added.shape[0], self.neural_points.shape[0]
        cur = self.buffer_pt_index[hv]
        cur[upd] = torch.arange(n_new, dtype=torch.int64, device=self.device) + n_old
        self.buffer_pt_index[hv] = cur
        dev, dt = self.device, self.dtype
        self.neural_points = torch.cat((self.neural_points, added), 0)
        ident = torch.zeros(n_new, 4, dtype=dt, device=dev)
        ident[:, 0] = 1.0
        self.point_orientations = torch.cat((self.point_orientations, ident), 0)
        ts = torch.full((n_new,), int(cur_ts), dtype=torch.int32, device=dev)
        self.point_ts_create = torch.cat((self.point_ts_create, ts), 0)
        self.point_ts_update = torch.cat((self.point_ts_update, ts), 0)
        self.geo_features = torch.cat((self.geo_features[:-1], self.geo_feature_std * torch.randn(
            n_new + 1, self.geo_feature_dim, device=dev, dtype=dt)), 0)
        if self.color_features is not None:
            self.color_features = torch.cat((self.color_features[:-1], self.color_feature_std * torch.randn(
                n_new + 1, self.color_feature_dim, device=dev, dtype=dt)), 0)
        self.point_certainties = torch.cat((self.point_certainties, torch.zeros(n_new, dtype=dt, device=dev)), 0)
        self.free_gs_mask = torch.cat((self.free_gs_mask, torch.full((n_new,), not is_reliable, dtype=torch.bool, device=dev)), 0)
        if sc is not None:
            self.point_colors = torch.cat((self.point_colors, sc[upd]), 0)
            self.valid_color_mask = torch.cat((self.valid_color_mask, sc[upd][:, 0] >= 0.0), 0)
        else:
            self.valid_color_mask = torch.cat((self.valid_color_mask, torch.ones(n_new, dtype=torch.bool, device=dev)), 0)
        self.valid_gs_mask = torch.cat((self.valid_gs_mask, torch.ones(n_new, dtype=torch.bool, device=dev)), 0)
        if sensor_position is not None:
            self.reset_local_map(sensor_position, sensor_orientation, cur_ts)
        return n_new / max(sp.shape[0], 1)

    def reset_local_map(self, sensor_position, sensor_orientation=None, cur_ts=0, use_travel_dist=True, diff_ts_local=50):
        """NP:378-478: select the local (trainable) window and the frozen surrounding ring."""
        self.cur_ts = cur_ts
        self.max_ts = max(self.max_ts, cur_ts)
        dev = self.device
        n = self.count()
        time_mask = torch.ones(n, dtype=torch.bool, device=dev)
        if self.temporal_local_map_on:
            used = ((self.point_ts_create + self.point_ts_update) / 2).int() if self.config.use_mid_ts else self.point_ts_create
            if use_travel_dist:
                tm = torch.abs(self.travel_dist[cur_ts] - self.travel_dist[used.long()]) < self.diff_travel_dist_local
            else:
                tm = torch.abs(cur_ts - used) < diff_ts_local
            if tm.sum() >= 100:
                time_mask = tm
        vec = self.neural_points[time_mask] - sensor_position
        d2 = (vec[:, :2] ** 2).sum(-1) if self.config.range_filter_2d else (vec ** 2).sum(-1)
        in_local = d2 < self.config.local_map_radius ** 2
        in_ring = d2 < self.sorrounding_map_radius ** 2
        tidx = torch.nonzero(time_mask).flatten()
        local_mask = torch.zeros(n, dtype=torch.bool, device=dev)
        local_mask[tidx[in_local]] = True
        ring = torch.zeros(n, dtype=torch.bool, device=dev)
        ring[tidx[~in_local & in_ring]] = True
        pad = torch.tensor([True], device=dev)
        self.sorrounding_mask = torch.cat((ring, pad))
        self.local_neural_points = self.neural_points[local_mask]
        self.local_point_orientations = self.point_orientations[local_mask]
        self.local_point_certainties = self.point_certainties[local_mask]
        self.local_point_ts_update = self.point_ts_update[local_mask]
        if self.point_colors is not None:
            self.local_point_colors = self.point_colors[local_mask]
        self.local_valid_color_mask = self.valid_color_mask[local_mask]
        self.local_valid_gs_mask = self.valid_gs_mask[local_mask]
        self.local_free_gs_mask = self.free_gs_mask[local_mask]
        local_mask = torch.cat((local_mask, pad))
        self.local_mask = local_mask
        # NP:455-457: `torch.full_like(local_mask, -1).long()` on a BOOL mask yields 1, not -1: points outside the
        # local window map to local index 1 in the reference.  Kept (parity quirk; the kernels read the table as is).
        g2l = torch.full((n + 1,), 1, dtype=torch.int64, device=dev)
        li = torch.nonzero(local_mask).flatten()
        g2l[li] = torch.arange(li.numel(), device=dev)
        g2l[-1] = -1
        self.global2local = g2l
        self.local_geo_features = nn.Parameter(self.geo_features[local_mask])
        if self.color_features is not None:
            self.local_color_features = nn.Parameter(self.color_features[local_mask])
        self.local_orientation = sensor_orientation
        self.local_position = sensor_position.float()

    def recreate_hash(self, sensor_position, sensor_orientation=None, kept_points: bool = True, with_ts: bool = True,
                      cur_ts=0):
        """NP:939-1024: rebuild the slot table from the stored points, one point per voxel (the one created
        closest to cur_ts, or the most certain one); kept_points=False also drops the other points."""
        res, dev = self.resolution, self.device
        self.buffer_pt_index = torch.full((self.buffer_size,), -1, dtype=torch.int64, device=dev)
        if with_ts:
            used = ((self.point_ts_create + self.point_ts_update) / 2).int() if self.config.use_mid_ts else self.point_ts_create
            value = torch.abs(used - cur_ts).float()
        else:
            value = self.point_certainties.max() - self.point_certainties
        sample_idx = voxel_down_sample_min_value(self.neural_points, res, value)
        if kept_points:
            sp = self.neural_points[sample_idx]
            self.buffer_pt_index[self._hash((sp / res).floor().to(torch.int64))] = sample_idx
        else:
            self.neural_points = self.neural_points[sample_idx]
            self.point_orientations = self.point_orientations[sample_idx]
            self.point_ts_create = self.point_ts_create[sample_idx]
            self.point_ts_update = self.point_ts_update[sample_idx]
            self.point_certainties = self.point_certainties[sample_idx]
            if self.point_colors is not None:
                self.point_colors = self.point_colors[sample_idx]
            pad = torch.cat((sample_idx, sample_idx.new_full((1,), -1)))   # -1 = the trailing padding row
            self.geo_features = self.geo_features[pad]
            if self.color_features is not None:
                self.color_features = self.color_features[pad]
            self.valid_color_mask = self.valid_color_mask[sample_idx]
            self.valid_gs_mask = self.valid_gs_mask[sample_idx]
            self.free_gs_mask = self.free_gs_mask[sample_idx]
            n = self.neural_points.shape[0]
            self.buffer_pt_index[self._hash((self.neural_points / res).floor().to(torch.int64))] = torch.arange(
                n, dtype=torch.int64, device=dev)
        if sensor_position is not None:
            self.reset_local_map(sensor_position, sensor_orientation, cur_ts)

    def assign_local_to_global(self):
        """NP:482-494."""
        lm = self.local_mask
        self.point_certainties[lm[:-1]] = self.local_point_certainties
        self.point_ts_update[lm[:-1]] = self.local_point_ts_update
        self.geo_features[lm] = self.local_geo_features.data
        if self.color_features is not None:
            self.color_features[lm] = self.local_color_features.data


def voxel_down_sample_min_value(points: torch.Tensor, voxel_size: float, value: torch.Tensor) -> torch.Tensor:
    """Index of the point with the smallest `value` in each occupied voxel (utils/tools.py:970-1010: value
    quantised to 1000 levels relative to its maximum, ties -> lowest index)."""
    g = torch.floor(points / voxel_size).long() - torch.floor(points.min(dim=0)[0] / voxel_size).long()
    vs = g.max()
    key = g[:, 0] + g[:, 1] * vs + g[:, 2] * vs * vs
    uniq, inv = torch.unique(key, return_inverse=True)
    ar = torch.arange(inv.numel(), dtype=torch.int64, device=points.device)
    off = 10 ** len(str(inv.numel() - 1))
    vq = (value / value.max() * 999).long()
    best = torch.empty(uniq.shape, dtype=torch.int64, device=points.device).scatter_reduce_(
        0, inv, ar + vq * off, reduce="amin", include_self=False)
    return best % off


def voxel_down_sample(points: torch.Tensor, voxel_size: float) -> torch.Tensor:
    """Index of the point closest to each occupied voxel's centre (utils/tools.py:924-967 semantics:
    distance quantised to 1000 levels, ties -> lowest index)."""
    g = torch.floor(points / voxel_size)
    dist = ((points - (g + 0.5) * voxel_size) ** 2).sum(1) ** 0.5
    dq = (dist / dist.max() * 999).long()
    g = g.long() - torch.floor(points.min(dim=0)[0] / voxel_size).long()
    vs = g.max()
    key = g[:, 0] + g[:, 1] * vs + g[:, 2] * vs * vs
    uniq, inv = torch.unique(key, return_inverse=True)
    ar = torch.arange(inv.numel(), dtype=torch.int64, device=points.device)
    off = 10 ** len(str(inv.numel() - 1))
    best = torch.empty(uniq.shape, dtype=torch.int64, device=points.device).scatter_reduce_(
        0, inv, ar + dq * off, reduce="amin", include_self=False)
    return best % off
