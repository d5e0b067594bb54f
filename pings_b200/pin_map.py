"""Reader of the reference's on-disk map format `pin_map.pth`.

The reference writes `{"neural_points": <the live NeuralPoints nn.Module>, "sdf": state_dict | None, "color": ...,
"gauss_xyz": ..., ...}` with torch.save (utils/tools.py:469-491 save_implicit_map) and reads it back with a plain
torch.load (utils/tools.py:494-520, inspect_pings.py:184-200) - i.e. the file is a pickle that names the classes
`model.neural_gaussians.NeuralPoints` and `utils.config.Config`.  This loader resolves those two names to light shells
(no reference code needs to be importable), adopts every tensor of the pickled module into a
pings_b200.neural_points.NeuralPoints without copying, and loads the decoders' state_dicts into pings_b200 Decoders
(same parameter layout).
"""
from __future__ import annotations

import pickle
from types import SimpleNamespace
from typing import Dict, Optional

import torch
import torch.nn as nn

from .neural_points import NeuralPoints, default_config


class _RefNeuralPointsShell(nn.Module):
    """Stands in for model.neural_gaussians.NeuralPoints while unpickling (nn.Module.__setstate__ restores __dict__)."""


class _RefConfigShell(SimpleNamespace):
    """Stands in for utils.config.Config (a plain attribute bag)."""

    def __setstate__(self, state):
        self.__dict__.update(state)


class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module == "model.neural_gaussians" and name == "NeuralPoints":
            return _RefNeuralPointsShell
        if module == "utils.config" and name == "Config":
            return _RefConfigShell
        return super().find_class(module, name)


class _PickleModule:
    """The `pickle_module` interface torch.load expects."""
    __name__ = "pings_b200.pin_map"
    Unpickler = _Unpickler
    load = staticmethod(lambda f, **kw: _Unpickler(f, **kw).load())
    loads = staticmethod(pickle.loads)
    dump = staticmethod(pickle.dump)
    dumps = staticmethod(pickle.dumps)


def load_implicit_map(model_path: str, mlp_dict: Optional[Dict[str, nn.Module]] = None, config=None, device="cuda",
                      freeze_decoders: bool = True):
    """-> (NeuralPoints, loaded dict).  Mirrors inspect_pings.py:184-200 + utils/tools.py:522-531 load_decoders:
    the map's tensors move to `device`, `temporal_local_map_on` is switched off as the inspector does, and every
    decoder of `mlp_dict` whose key is present in the file gets its state_dict (and is frozen)."""
    loaded = torch.load(model_path, map_location="cpu", pickle_module=_PickleModule, weights_only=False)
    shell = loaded["neural_points"]
    ref_cfg = getattr(shell, "config", None)
    if config is None:
        kw = {}
        for k in vars(default_config()):
            if ref_cfg is not None and hasattr(ref_cfg, k):
                kw[k] = getattr(ref_cfg, k)
        kw["device"] = device
        kw["dtype"] = torch.float32
        config = default_config(**kw)
    dev = torch.device(device)
    for k, v in list(vars(shell).items()):
        if isinstance(v, torch.Tensor):
            setattr(shell, k, v.to(dev))
    for k, v in list(shell._parameters.items()):
        if v is not None:
            shell._parameters[k] = nn.Parameter(v.data.to(dev), requires_grad=v.requires_grad)
    shell.config = config
    npm = NeuralPoints.from_reference(shell)
    npm.device = dev
    npm.temporal_local_map_on = False
    if npm.travel_dist is not None and isinstance(npm.travel_dist, torch.Tensor):
        npm.travel_dist = npm.travel_dist.to(dev)
    if mlp_dict is not None:
        for key, state in loaded.items():
            if key == "neural_points" or state is None or mlp_dict.get(key) is None:
                continue
            mlp_dict[key].load_state_dict(state)
            mlp_dict[key].to(dev)
            if freeze_decoders:
                for p in mlp_dict[key].parameters():
                    p.requires_grad_(False)
    return npm, loaded
