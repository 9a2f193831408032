"""diner_b200 -- B200-native implementation of DINER's per-ray rendering hot path.

Public API (mirrors the reference plugin, src/models/nerf_renderer.py): `NeRFRendererDGS`; and the two input
producers next to it with the reference's signatures: `gen_rays`, `depth2normal` (+ `gen_rays_at` for selected pixels);
the caller-side ray handling of `DINER` lives in `diner_b200.callers`, the decoder of the datasets' on-disk uint16 depth
format in `diner_b200.depth_format`.
"""
from .producers import depth2normal, gen_rays, gen_rays_at  # noqa: F401
from .renderer import NeRFRendererDGS, RenderResult, sample_rays  # noqa: F401
from .scene import invalidate  # noqa: F401

__all__ = ["NeRFRendererDGS", "RenderResult", "sample_rays", "gen_rays", "gen_rays_at", "depth2normal", "invalidate"]
