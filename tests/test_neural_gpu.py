"""GPU parity of the neural-point kernels (through the C-ABI + the Python mirror) against
(a) golden vectors produced by the reference itself (tests/golden/neural_*.npz) and
(b) the CPU oracle (oracle/neural_cpu.py) on larger seeded inputs.
kNN index lists / nn_counts bit-exact; dist2 bit-exact; interpolated features, weights, SDF and all
gradients within rtol 1e-4 / atol 1e-6 (fp32 summation order differs from torch's)."""
import os

import numpy as np
import pytest
import torch

from oracle import neural_cpu as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["pin", "gs", "window"]
RTOL, ATOL = 1e-4, 2e-6


def build(name):
    from pings_b200.decoder import Decoder
    from pings_b200.neural_points import NeuralPoints, default_config
    z = np.load(os.path.join(GOLD, f"neural_{name}.npz"))
    m = z["meta"]
    cfg = default_config(voxel_size_m=float(m[0]), num_nei_cells=int(m[1]), search_alpha=float(m[2]), query_nn_k=int(m[3]),
                         buffer_size=int(m[4]), feature_dim=int(m[5]), color_feature_dim=int(m[6]), color_on=int(m[6]) > 0,
                         weighted_first=bool(m[9]))
    state = {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("state.")}
    npm = NeuralPoints.from_state(cfg, state)
    npm.after_pgo = bool(m[10])
    npm.cur_ts = int(m[11])
    npm.diff_travel_dist_local = float(m[12])
    npm.temporal_local_map_on = bool(m[13])
    dec = Decoder(cfg, int(m[5]), int(m[7]), int(m[8]), 1)
    dec.load_state_dict({k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("dec.")})
    dec.sdf_scale = float(m[17])
    return z, m, npm, dec.cuda()


@pytest.mark.parametrize("name", CASES)
def test_radius_search_bitexact(name):
    z, m, npm, dec = build(name)
    d2, idx = npm.radius_neighborhood_search(torch.from_numpy(z["q"]).cuda(), time_filtering=bool(m[13]))
    assert np.array_equal(idx.cpu().numpy().astype(np.int32), z["idx"])
    assert np.array_equal(d2.cpu().numpy(), z["dist2"])


@pytest.mark.parametrize("name", CASES)
def test_query_feature_sdf_and_gradients(name):
    z, m, npm, dec = build(name)
    wf, color_q = bool(m[9]), bool(m[14])
    q = torch.from_numpy(z["q"]).cuda().requires_grad_(True)
    q_ts = torch.from_numpy(z["q_ts"]).cuda()
    geo_vec, color_vec, weight, nn_counts, certainty = npm.query_feature(
        q, q_ts, accumulate_stability=True, query_color_feature=color_q, use_only_valid_points=bool(m[15]),
        use_only_measured_points=bool(m[16]))
    assert np.array_equal(nn_counts.cpu().numpy(), z["nn_counts"])
    c = lambda t: t.detach().cpu().numpy()
    np.testing.assert_allclose(c(weight), z["weight"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(c(geo_vec), z["geo_vec"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(c(certainty), z["certainty"], rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(c(npm.local_point_certainties), z["cert_after"], rtol=RTOL, atol=1e-5)
    assert np.array_equal(c(npm.local_point_ts_update), z["ts_after"])
    sdf = dec.sdf(geo_vec)
    if not wf:
        sdf = torch.sum(sdf * weight, dim=1).squeeze(1)
    np.testing.assert_allclose(c(sdf), z["sdf"], rtol=RTOL, atol=ATOL)
    loss = (sdf * torch.from_numpy(z["r1"]).cuda()).sum() + (weight * torch.from_numpy(z["r2"]).cuda()).sum()
    if color_q:
        np.testing.assert_allclose(c(color_vec), z["color_vec"], rtol=RTOL, atol=ATOL)
        loss = loss + (color_vec * torch.from_numpy(z["r3"]).cuda()).sum()
    loss.backward()

    def close(a, b, what):
        scale = np.abs(b).max() + 1e-30
        assert np.abs(a - b).max() / scale < 2e-4, (what, np.abs(a - b).max(), scale)

    close(c(npm.local_geo_features.grad), z["d_geo"], "d_geo")
    close(c(q.grad), z["d_q"], "d_q")
    if color_q:
        close(c(npm.local_color_features.grad), z["d_color"], "d_color")
    for k, p in dec.named_parameters():
        close(c(p.grad), z[f"dgrad.{k}"], k)


@pytest.mark.parametrize("name", CASES)
def test_fused_sdf_query_matches(name):
    z, m, npm, dec = build(name)
    sdf, std, cnt = npm.sdf_fused(torch.from_numpy(z["q"]).cuda(), dec, get_std=True,
                                  use_only_valid_points=bool(m[15]), use_only_measured_points=bool(m[16]))
    assert np.array_equal(cnt.cpu().numpy(), z["nn_counts"])
    np.testing.assert_allclose(sdf.cpu().numpy(), z["sdf"], rtol=RTOL, atol=ATOL)
    if std is not None:
        S = O.state_from_golden(z)
        nl = int(m[8])
        Ws = [torch.from_numpy(z[f"dec.layers.{i}.weight"]) for i in range(nl)] + [torch.from_numpy(z["dec.lout.weight"])]
        Bs = [torch.from_numpy(z[f"dec.layers.{i}.bias"]) for i in range(nl)] + [torch.from_numpy(z["dec.lout.bias"])]
        _, std_ref, _ = O.sdf_query(S, torch.from_numpy(z["q"]), int(m[3]), False, Ws, Bs, float(m[17]),
                                    time_filtering=bool(m[13]), use_only_valid=bool(m[15]), use_only_measured=bool(m[16]),
                                    after_pgo=bool(m[10]))
        np.testing.assert_allclose(std.cpu().numpy(), std_ref.numpy(), rtol=1e-3, atol=1e-5)


def test_large_scan_vs_oracle_and_properties():
    """C1-shaped inputs (64x1800 synthetic scan) against the CPU oracle + size-independent properties:
    weights sum to 1 where any neighbour exists, nn_counts == #(idx >= 0), queries are independent (a
    permutation of the queries permutes the outputs)."""
    from pings_b200 import synth_neural as SN
    npm, dec, q = SN.c1_map_and_queries(device="cuda", n_az=600)
    S = SN.oracle_state(npm)
    qc = q.cpu()
    d2, idx = npm.radius_neighborhood_search(q, time_filtering=False)
    d2o, idxo = O.radius_search(S, qc, False)
    assert torch.equal(idx.cpu(), idxo) and torch.equal(d2.cpu(), d2o)
    sdf, _, cnt = npm.sdf_fused(q, dec)
    Ws, Bs = [l.weight.detach().cpu() for l in dec._linears()], [l.bias.detach().cpu() for l in dec._linears()]
    so, _, co = O.sdf_query(S, qc, 6, True, Ws, Bs, dec.sdf_scale, time_filtering=False)
    assert torch.equal(cnt.cpu(), co)
    np.testing.assert_allclose(sdf.cpu().numpy(), so.numpy(), rtol=RTOL, atol=ATOL)
    geo, _, w, cnt2, _ = npm.query_feature(q, accumulate_stability=False)
    ws = w.sum(1).squeeze(-1)
    assert torch.allclose(ws[cnt2 > 0], torch.ones_like(ws[cnt2 > 0]), atol=1e-5)
    assert float(ws[cnt2 == 0].abs().sum()) == 0.0
    perm = torch.randperm(q.shape[0], device="cuda")
    sdf_p, _, _ = npm.sdf_fused(q[perm], dec)
    assert torch.equal(sdf_p, sdf[perm])


def test_sdf_query_mirror_paths_agree():
    """pings_b200.sdf_ops.SdfQuery (Mapper.sdf / sdf_batch / get_numerical_gradient, utils/mapper.py:2273-2369):
    the fused no_grad kernel, the autograd composition and the chunked batch variant give the same SDF; the
    numerical gradient equals central differences of sdf() evaluated point by point."""
    from pings_b200 import synth_neural as SN
    from pings_b200.sdf_ops import SdfQuery
    npm, dec, q = SN.c1_map_and_queries(device="cuda", n_az=200)
    sq = SdfQuery(npm, dec)
    q = q[:20000]
    with torch.no_grad():
        s_fused, _, valid_fused = sq.sdf(q)
    s_grad, _, valid_grad = sq.sdf(q.clone())                     # grad mode: query_feature + decoder
    s_batch, _, valid_batch = sq.sdf_batch(q, 4096)
    assert torch.equal(valid_fused, valid_grad) and torch.equal(valid_fused, valid_batch)
    np.testing.assert_allclose(s_fused.cpu().numpy(), s_grad.detach().cpu().numpy(), rtol=RTOL, atol=ATOL)
    # chunks below 4096 rows take the SIMT decoder, larger ones the tcgen05 one: fp32-level differences only
    np.testing.assert_allclose(s_batch.detach().cpu().numpy(), s_grad.detach().cpu().numpy(), rtol=RTOL, atol=ATOL)
    with torch.no_grad():
        g2 = sq.get_numerical_gradient(q[:512], eps=0.02, two_side=True)
        e = torch.tensor([0.02, 0.0, 0.0], device="cuda")
        gx = (sq.sdf(q[:512] + e)[0] - sq.sdf(q[:512] - e)[0]) / 0.04
        assert torch.equal(g2[:, 0], gx)
        g1 = sq.get_numerical_gradient(q[:512], sdf_x=s_fused[:512], eps=0.02, two_side=False)
        assert torch.equal(g1[:, 0], (sq.sdf(q[:512] + e)[0] - s_fused[:512]) / 0.02)
        assert g2.shape == (512, 3) and g1.shape == (512, 3)


def test_query_certainty_matches_reference_golden():
    """NeuralPoints.query_certainty (NP:1117-1133) with the single-cell neighbourhood the mapper switches to
    (utils/mapper.py:462-476) and with the default one, on the map the mirror's own update() builds on the GPU
    from the three scans of tests/golden/maintenance.npz; the expected values come from the reference
    (tests/golden/make_maintenance_golden.py).  A max over stored certainties: exact."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("t_maint", os.path.join(os.path.dirname(__file__), "test_maintenance_cpu.py"))
    tm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tm)
    G = tm.G
    npm = tm.build_map("cuda")
    tm.insert_scans(npm)
    tm.apply_edits(npm)
    assert np.array_equal(npm.neural_points.cpu().numpy(), G["assigned.neural_points"])   # GPU-built map == reference map
    occ = torch.nonzero(npm.buffer_pt_index >= 0).flatten()
    assert np.array_equal(occ.cpu().numpy(), G["assigned.hash_slots"])
    assert np.array_equal(npm.buffer_pt_index[occ].cpu().numpy(), G["assigned.hash_values"])
    q = torch.from_numpy(G["certainty.q"]).cuda()
    npm.set_search_neighborhood(num_nei_cells=1, search_alpha=0.0)
    assert npm.neighbor_K == int(G["certainty.K"][0]) == 1   # dx^2 < (1 + 0)^2 keeps the query's own voxel only
    c1 = npm.query_certainty(q)
    npm.set_search_neighborhood(num_nei_cells=npm.config.num_nei_cells, search_alpha=npm.config.search_alpha)
    c2 = npm.query_certainty(q)
    assert np.array_equal(c1.cpu().numpy(), G["certainty.value"])
    assert np.array_equal(c2.cpu().numpy(), G["certainty.value_default_nbh"])
    assert float(c1[:20].abs().max()) == 0.0 and float(c1.max()) > 0.0


def test_fused_sdf_dense_volume_many_candidates():
    """A volume filled at the voxel size: nearly every one of the 81 probed cells holds a point, so a query has far
    more than 16 occupied slots - the half-warp SDF kernel's multi-pass distance / rank-counting path.  nn_counts
    bit-exact against the CPU oracle, SDF within the usual tolerance."""
    from pings_b200.decoder import Decoder
    from pings_b200.neural_points import NeuralPoints, default_config
    from pings_b200 import synth_neural as SN
    g = torch.Generator().manual_seed(5)
    cfg = default_config(device="cuda", buffer_size=int(2e6), track_on=False, temporal_local_map_off=True)
    res = cfg.voxel_size_m
    n = 14
    ii = torch.stack(torch.meshgrid(*[torch.arange(n, dtype=torch.float32)] * 3, indexing="ij"), -1).reshape(-1, 3)
    pts = ((ii + 0.5 + 0.3 * (torch.rand(ii.shape, generator=g) - 0.5)) * res).contiguous()
    npm = NeuralPoints(cfg)
    npm.travel_dist = torch.zeros(1, device="cuda")
    npm.update(pts.cuda(), None, None, torch.zeros(3, device="cuda"), None, 0)
    npm.geo_features = (0.1 * torch.randn(npm.count() + 1, cfg.feature_dim, generator=g)).cuda()
    npm.reset_local_map(torch.zeros(3, device="cuda"), None, 0)
    torch.manual_seed(3)
    dec = Decoder(cfg, cfg.feature_dim, 64, 1, 1).cuda()
    q = (torch.rand(6001, 3, generator=g) * (n * res)).contiguous()      # odd count: the idle half of the last pair
    S = SN.oracle_state(npm)
    d2o, idxo = O.radius_search(S, q, False)
    assert int((idxo >= 0).sum(1).max()) > 16                              # more candidates than a half-warp has lanes
    sdf, _, cnt = npm.sdf_fused(q.cuda(), dec)
    Ws, Bs = [l.weight.detach().cpu() for l in dec._linears()], [l.bias.detach().cpu() for l in dec._linears()]
    so, _, co = O.sdf_query(S, q, cfg.query_nn_k, True, Ws, Bs, dec.sdf_scale, time_filtering=False)
    assert torch.equal(cnt.cpu(), co)
    np.testing.assert_allclose(sdf.cpu().numpy(), so.numpy(), rtol=RTOL, atol=ATOL)
