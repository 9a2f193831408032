"""Caller-side ray handling around the renderer: what `DINER` does between a batch and
`renderer.forward` (reference src/models/diner.py), with the ray generation on the device.

    predict_imgs(...)        DINER.predict_imgs_from_batch   diner.py:72-97   (after encode_batch)
    render_sweep(...)        the frame loop of create_cam_sweep   diner.py:178-199
    random_pixels(...)       diner.py:229
    patch_pixels(...)        diner.py:231-247
    training_rays(...)       diner.py:222-227, :256-263   rays + ground-truth colours of the selected pixels

The reference cannot import its own `diner.py` here (pytorch_lightning is absent), so these are
restatements of the cited lines; tests compare them with the numpy restatement in oracle/producers_oracle.py.
"""
from __future__ import annotations

import torch

from .producers import gen_rays, gen_rays_at


def _per_scene(v, SB: int, device) -> torch.Tensor:
    """`self.znear.expand(SB)` (diner.py:77-78): a scalar or one value per scene."""
    t = torch.as_tensor(v, dtype=torch.float32, device=device).reshape(-1)
    if t.numel() == 1:
        t = t.expand(SB)
    if t.numel() != SB:
        raise ValueError("z_near / z_far must be a scalar or hold one value per scene")
    return t.contiguous()


@torch.no_grad()
def predict_imgs(renderer, model, target_extrinsics, target_intrinsics, W, H, z_near, z_far, ray_batch_size=None,
                 return_depth: bool = False):
    """Render the full target frames of SB scenes already encoded in `model`: rgb (SB,3,H,W) [, depth (SB,1,H,W)].

    `ray_batch_size=None` hands the whole frame to `renderer.forward` (it tiles the rays itself); an integer splits the
    rays along dim 1 exactly like the reference loop (diner.py:86-89), which only matters for reproducing its noise
    consumption order."""
    SB = target_extrinsics.shape[0]
    dev = target_extrinsics.device
    rays = gen_rays(target_extrinsics, target_intrinsics, W, H, _per_scene(z_near, SB, dev), _per_scene(z_far, SB, dev))
    rays = rays.view(SB, H * W, 8)
    if ray_batch_size is None:
        out = renderer.forward(model=model, rays=rays)
        rgb, depth = out.fine.rgb, out.fine.depth
    else:
        rgbs, depths = [], []
        for ray_batch in torch.split(rays, int(ray_batch_size), dim=1):
            out = renderer.forward(model=model, rays=ray_batch.contiguous())
            rgbs.append(out.fine.rgb)
            depths.append(out.fine.depth)
        rgb, depth = torch.cat(rgbs, dim=1), torch.cat(depths, dim=1)
    rgb = rgb.view(SB, H, W, 3).permute(0, 3, 1, 2)
    depth = depth.view(SB, H, W, 1).permute(0, 3, 1, 2)
    return (rgb, depth) if return_depth else rgb


@torch.no_grad()
def render_sweep(renderer, model, target_extrinsics, target_intrinsics, W, H, z_near, z_far, frames_per_call: int = 1):
    """Frames of a camera sweep around ONE encoded scene (SB = 1): target_extrinsics (F,4,4), one intrinsics (1,3,3) ->
    rgb (F,3,H,W), depth (F,1,H,W).  The rays of all frames come from one launch; `frames_per_call` frames are handed
    to the renderer as one ray batch (rays are independent, so the images do not depend on it)."""
    F = target_extrinsics.shape[0]
    dev = target_extrinsics.device
    K = target_intrinsics.expand(F, 3, 3)
    rays = gen_rays(target_extrinsics, K, W, H, _per_scene(z_near, 1, dev).expand(F), _per_scene(z_far, 1, dev).expand(F))
    rays = rays.view(1, F * H * W, 8)
    rgbs, depths = [], []
    step = max(1, int(frames_per_call)) * H * W
    for ray_batch in torch.split(rays, step, dim=1):
        out = renderer.forward(model=model, rays=ray_batch)
        rgbs.append(out.fine.rgb.view(-1, H, W, 3))
        depths.append(out.fine.depth.view(-1, H, W, 1))
    return torch.cat(rgbs).permute(0, 3, 1, 2), torch.cat(depths).permute(0, 3, 1, 2)


def random_pixels(SB: int, H: int, W: int, ray_batch_size: int, device=None, generator=None) -> torch.Tensor:
    """Uniform pixel indices (SB, ray_batch_size) -- diner.py:229 (the reference draws them on the CPU)."""
    idx = torch.randint(0, H * W, (SB, ray_batch_size), generator=generator)
    return idx if device is None else idx.to(device)


def patch_pixels(target_alpha: torch.Tensor, patch: int, generator=None):
    """One `patch` x `patch` window per scene centred on a foreground pixel: flat indices (SB, patch*patch) and the
    centres (SB, 2) as (x, y) -- diner.py:231-247.  target_alpha (SB,1,H,W); foreground pixels closer than
    pad = (patch+1)//2 to the border are not eligible (:233-237), so every window lies inside the image."""
    fg = target_alpha[:, 0].clone()
    SB, H, W = fg.shape
    pad = (patch + 1) // 2
    fg[..., :pad] = 0
    fg[..., :pad, :] = 0
    fg[..., -pad:] = 0
    fg[..., -pad:, :] = 0
    centers = torch.multinomial(fg.reshape(SB, H * W), 1, generator=generator)            # (SB,1)
    return patch_pixels_at(centers, W, patch), torch.cat((centers % W, torch.div(centers, W, rounding_mode="floor")), -1)


def patch_pixels_at(centers: torch.Tensor, W: int, patch: int) -> torch.Tensor:
    """Flat indices of the window around flat centre indices (SB,1) or (SB,): x and y run from c - pad to
    c - pad + patch - 1, y-major (diner.py:238-247)."""
    c = centers.reshape(-1, 1, 1).to(torch.int64)
    pad = (patch + 1) // 2
    off = torch.arange(patch, device=c.device, dtype=torch.int64) - pad
    x = (c % W) + off.view(1, 1, patch)
    y = torch.div(c, W, rounding_mode="floor") + off.view(1, patch, 1)
    return (x + y * W).flatten(1)


def training_rays(target_extrinsics, target_intrinsics, W, H, z_near, z_far, pix_idcs, target_rgb=None,
                  check_bounds: bool = False):
    """Rays (SB,B,8) of the selected pixels and, if `target_rgb` (SB,3,H,W) is given, their colours (SB,B,3):
    diner.py:222-227 + :256-263 without the full-frame ray tensor."""
    SB = target_extrinsics.shape[0]
    dev = target_extrinsics.device
    pix = pix_idcs.to(dev)
    rays = gen_rays_at(target_extrinsics, target_intrinsics, W, H, _per_scene(z_near, SB, dev),
                       _per_scene(z_far, SB, dev), pix, check_bounds=check_bounds)
    if target_rgb is None:
        return rays
    helper = torch.arange(SB, device=dev).unsqueeze(-1).expand(-1, pix.shape[1])
    gt = target_rgb.view(SB, 3, -1).permute(0, 2, 1)[helper, pix]
    return rays, gt
