"""TEST INFRASTRUCTURE — imports the reference's own Python modules (model/neural_gaussians.py,
model/decoder.py, gaussian_splatting/...) from /root/reference on CPU, with import-time stubs for
the GUI/IO packages that are absent here.  Only used by the golden-fixture
generators under tests/golden/ and by tests that run in the build container; the GPU box has no
/root/reference, so everything that must travel is committed as fixtures.
"""
from __future__ import annotations

import os
import sys
from unittest.mock import MagicMock

REF = os.environ.get("PINGS_REFERENCE", "/root/reference")
if not os.path.isdir(os.path.join(REF, "model")):
    # the GPU box has no /root/reference: use the modules oracle/build_ref.py staged into oracle/_ref/py/
    _staged = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "py")
    if os.path.isdir(os.path.join(_staged, "model")):
        REF = _staged


def available() -> bool:
    return os.path.isdir(os.path.join(REF, "model"))


def staged_only() -> bool:
    return REF.endswith(os.path.join("_ref", "py"))


def _stub():
    for m in ["open3d", "matplotlib", "matplotlib.cm", "matplotlib.pyplot", "roma", "imageio", "wandb", "natsort",
              "open3d.visualization", "open3d.visualization.gui", "open3d.visualization.rendering", "pypose", "kornia",
              "cv2", "plyfile", "rich", "gtsam", "pytorch3d", "skimage", "skimage.metrics", "skimage.measure", "lpips"]:
        if m not in sys.modules:
            try:
                __import__(m)
            except Exception:
                sys.modules[m] = MagicMock()


def load_config(yaml_name: str = "run_pin_slam.yaml", **overrides):
    """Reference Config with device=cpu."""
    _stub()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from utils.config import Config  # type: ignore
    import torch
    cfg = Config()
    cfg.load(os.path.join(REF, "config", yaml_name))
    cfg.device = "cpu"
    cfg.dtype = torch.float32
    cfg.tran_dtype = torch.float64
    for k, v in overrides.items():
        setattr(cfg, k, v)
    return cfg


def neural_points_cls():
    _stub()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from model.neural_gaussians import NeuralPoints  # type: ignore
    return NeuralPoints


def decoder_cls():
    _stub()
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from model.decoder import Decoder  # type: ignore
    return Decoder
