"""Top-level alias so `from fused_ssim import fused_ssim` (utils/mapper.py, inspect_pings.py) resolves to pings_b200."""
from pings_b200.fused_ssim import FusedSSIMMap, allowed_padding, fused_ssim, fusedssim, fusedssim_backward  # noqa: F401
