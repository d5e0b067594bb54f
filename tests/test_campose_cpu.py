"""pings_b200.campose_utils against golden vectors generated from the reference
(tests/golden/make_campose_golden.py -> utils/campose_utils.py:28-76) and update_pose semantics."""
import os
import types

import numpy as np
import torch

from pings_b200 import campose_utils as CU

GOLD = os.path.join(os.path.dirname(__file__), "golden", "campose.npz")


def test_se3_exp_matches_reference_golden_batched_and_single():
    d = np.load(GOLD)
    tau, T = torch.from_numpy(d["tau"]), torch.from_numpy(d["T"])
    got = CU.SE3_exp(tau)                      # batched
    assert torch.allclose(got, T, rtol=0, atol=1e-12)
    for i in range(tau.shape[0]):              # single, like the reference call
        assert torch.allclose(CU.SE3_exp(tau[i]), T[i], rtol=0, atol=1e-12)
    R = got[:, :3, :3]
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3, dtype=torch.float64).expand_as(R), atol=1e-12)


def test_update_pose_left_multiplies_and_zeroes_deltas():
    cam = types.SimpleNamespace()
    cam.R = torch.eye(3)
    cam.T = torch.tensor([0.1, -0.2, 0.3])
    cam.cam_trans_delta = torch.nn.Parameter(torch.tensor([0.01, 0.02, -0.03]))
    cam.cam_rot_delta = torch.nn.Parameter(torch.tensor([0.002, -0.001, 0.004]))
    got = {}
    cam.set_pose = lambda T_c2w: got.setdefault("T", T_c2w)
    tau = torch.cat([cam.cam_trans_delta, cam.cam_rot_delta]).detach().clone()
    assert CU.update_pose(cam) is False
    T_w2c = torch.eye(4)
    T_w2c[:3, 3] = cam.T
    assert torch.allclose(torch.linalg.inv(got["T"]), CU.SE3_exp(tau) @ T_w2c, atol=1e-6)
    assert float(cam.cam_trans_delta.abs().sum()) == 0.0 and float(cam.cam_rot_delta.abs().sum()) == 0.0
    assert CU.update_pose(cam) is True          # zero step: pose optimisation off
