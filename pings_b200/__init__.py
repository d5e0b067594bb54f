"""pings_b200 — B200-native (sm_100a) implementation of the PINGS rendering / neural-point hot path.

Sub-packages mirror the reference's import surface:
  pings_b200.diff_gaussian_rasterization          (submodules/diff-gaussian-rasterization-w-pose)
  pings_b200.diff_gaussian_surfel_rasterization   (submodules/diff-gaussian-surfel-rasterization-w-pose)
  pings_b200.fused_ssim                           (submodules/fused-ssim)
Put `<repo>/dropin` on sys.path to get them under the reference's own top-level names.
"""
__version__ = "0.1.0"
