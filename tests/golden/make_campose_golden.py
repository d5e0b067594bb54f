"""Golden vectors for pings_b200.campose_utils from the reference (run in the build container only:
imports /root/reference/utils/campose_utils.py, which needs nothing but numpy + torch)."""
import importlib.util
import numpy as np
import torch

spec = importlib.util.spec_from_file_location("ref_campose", "/root/reference/utils/campose_utils.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)
g = torch.Generator().manual_seed(5)
taus = torch.randn(12, 6, generator=g, dtype=torch.float64) * torch.tensor([1.0, 1, 1, 0.7, 0.7, 0.7], dtype=torch.float64)
taus[0] = 0.0
taus[1, 3:] = torch.tensor([1e-7, -2e-7, 3e-7], dtype=torch.float64)    # small-angle branch
taus[2, 3:] = torch.tensor([0.0, 0.0, 3.1], dtype=torch.float64)
out = torch.stack([ref.SE3_exp(t) for t in taus])
np.savez("tests/golden/campose.npz", tau=taus.numpy(), T=out.numpy())
print("wrote", out.shape)
