"""diner_b200 -- B200-native implementation of DINER's per-ray rendering hot path.

Public API (mirrors the reference plugin, src/models/nerf_renderer.py): `NeRFRendererDGS`.
"""
from .renderer import NeRFRendererDGS, RenderResult, sample_rays  # noqa: F401

__all__ = ["NeRFRendererDGS", "RenderResult", "sample_rays"]
