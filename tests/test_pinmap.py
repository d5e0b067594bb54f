"""pin_map.pth reader (pings_b200/pin_map.py) and the Mesher.query_points mirror (pings_b200/mesher.py) against a file
WRITTEN BY THE REFERENCE (utils/tools.py:469-491) and the outputs of the reference's Mesher.query_points
(utils/mesher.py:40-166) on it - tests/golden/make_pinmap_golden.py.  The loader part runs on the CPU; no reference
module is importable here (the loader resolves the pickled class names itself)."""
import os
import sys

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(device):
    from pings_b200.decoder import Decoder
    from pings_b200.neural_points import default_config
    from pings_b200.pin_map import load_implicit_map
    z = np.load(os.path.join(GOLD, "pin_map_small.npz"))
    cfg = default_config(device=device, color_on=True, feature_dim=8, color_feature_dim=4, buffer_size=int(z["meta"][1]))
    decs = {"sdf": Decoder(cfg, 8, 64, 1, 1), "semantic": None, "color": Decoder(cfg, 4, 64, 1, 3)}
    npm, loaded = load_implicit_map(os.path.join(GOLD, "pin_map_small.pth"), decs, config=cfg, device=device)
    return z, cfg, npm, decs, loaded


def test_load_reference_pin_map_cpu():
    assert "model" not in sys.modules or not hasattr(sys.modules["model"], "neural_gaussians")
    z, cfg, npm, decs, loaded = _load("cpu")
    assert set(loaded.keys()) == {"neural_points", "sdf", "semantic", "color"} and loaded["semantic"] is None
    assert np.array_equal(npm.neural_points.numpy(), z["neural_points"])
    assert np.array_equal(npm.geo_features.detach().numpy(), z["geo_features"])
    assert np.array_equal(npm.color_features.detach().numpy(), z["color_features"])
    assert np.array_equal(npm.valid_gs_mask.numpy(), z["valid_gs_mask"])
    occ = torch.nonzero(npm.buffer_pt_index >= 0).flatten()
    assert np.array_equal(occ.numpy(), z["hash_slots"])
    assert np.array_equal(npm.buffer_pt_index[occ].numpy(), z["hash_values"])
    assert npm.temporal_local_map_on is False and npm.neighbor_K == 81
    for name in ("sdf", "color"):
        for k, v in decs[name].state_dict().items():
            assert np.array_equal(v.numpy(), z[f"dec.{name}.{k}"]), (name, k)
        assert all(not p.requires_grad for p in decs[name].parameters())


@pytest.mark.gpu
def test_mesher_query_points_matches_reference():
    from pings_b200.mesher import Mesher
    z, cfg, npm, decs, _ = _load("cuda")
    cfg.color_channel = 3
    m = Mesher(cfg, npm, decs)
    coord = torch.from_numpy(z["coord"])
    sdf, sem, color, mask = m.query_points(coord, 3000, query_sdf=True, query_color=True, query_mask=True,
                                           mask_min_nn_count=4, out_torch=True)
    assert sem is None
    assert np.array_equal(mask.cpu().numpy(), z["mask"])
    torch.testing.assert_close(sdf.cpu(), torch.from_numpy(z["sdf"]).float(), rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(color.cpu(), torch.from_numpy(z["color"]).float(), rtol=1e-4, atol=1e-6)
    s2, _, c2, m2 = m.query_points(coord, 5000, query_sdf=True, query_color=True, out_torch=False)
    assert isinstance(s2, np.ndarray) and c2.dtype == np.float64 and np.array_equal(m2, z["mask"])
