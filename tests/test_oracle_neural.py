"""CPU: the oracle restatement (oracle/neural_cpu.py) against the golden vectors produced by running the
reference itself (tests/golden/make_neural_golden.py). Index lists bit-exact, floats to 1e-6."""
import os

import numpy as np
import pytest
import torch

from oracle import neural_cpu as O

GOLD = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["pin", "gs", "window"]


def load(name):
    return np.load(os.path.join(GOLD, f"neural_{name}.npz"))


@pytest.mark.parametrize("name", CASES)
def test_radius_search_matches_reference(name):
    z = load(name)
    S = O.state_from_golden(z)
    d2, idx = O.radius_search(S, torch.from_numpy(z["q"]), bool(z["meta"][13]))
    assert np.array_equal(idx.numpy().astype(np.int32), z["idx"])
    assert np.array_equal(d2.numpy(), z["dist2"])


@pytest.mark.parametrize("name", CASES)
def test_query_feature_and_sdf_match_reference(name):
    z = load(name)
    m = z["meta"]
    S = O.state_from_golden(z)
    q = torch.from_numpy(z["q"])
    wf, after_pgo, color_q = bool(m[9]), bool(m[10]), bool(m[14])
    kw = dict(time_filtering=bool(m[13]), use_only_valid=bool(m[15]), use_only_measured=bool(m[16]), after_pgo=after_pgo)
    r = O.query_feature(S, q, int(m[3]), wf, geo=S["local_geo_features"],
                        color=S.get("local_color_features") if color_q else None, **kw)
    assert np.array_equal(r["nn_counts"].numpy(), z["nn_counts"])
    np.testing.assert_allclose(r["weight"].numpy(), z["weight"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(r["geo"].numpy(), z["geo_vec"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(r["certainty"].numpy(), z["certainty"], rtol=1e-5, atol=1e-6)
    if color_q:
        np.testing.assert_allclose(r["color"].numpy(), z["color_vec"], rtol=1e-5, atol=1e-6)
    cert, ts = O.stability_side_effects(r["knn_idx"], r["weight"], torch.from_numpy(z["q_ts"]),
                                        S["local_point_certainties"], S["local_point_ts_update"])
    np.testing.assert_allclose(cert.numpy(), z["cert_after"], rtol=1e-5, atol=1e-6)
    assert np.array_equal(ts.numpy(), z["ts_after"])
    nl = int(m[8])
    Ws = [torch.from_numpy(z[f"dec.layers.{i}.weight"]) for i in range(nl)] + [torch.from_numpy(z["dec.lout.weight"])]
    Bs = [torch.from_numpy(z[f"dec.layers.{i}.bias"]) for i in range(nl)] + [torch.from_numpy(z["dec.lout.bias"])]
    sdf, std, cnt = O.sdf_query(S, q, int(m[3]), wf, Ws, Bs, float(m[17]), **kw)
    np.testing.assert_allclose(sdf.numpy(), z["sdf"], rtol=1e-5, atol=1e-6)
