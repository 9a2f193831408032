"""CPU oracle for the DINER rendering hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
import this module; nothing under `diner_b200/` does.  It is a restatement, in plain PyTorch CPU
ops, of the reference renderer's algorithm (tancredeguillou/diner):

    src/models/nerf_renderer.py   NeRFRendererDGS.sample_coarse / sample_depthguided /
                                  fill_up_uniform_samples / composite / forward
    src/models/pixelnerf.py       PixelNeRF.forward
    src/models/resnetfc.py        ResnetFC.forward, ResnetBlockFC.forward, combine
    src/models/image_encoder.py   SpatialEncoder.index / index_depth / index_depth_std / index_normal
    src/models/positional_encoding.py  PositionalEncoding.forward
    src/util/torch_helpers.py     grid_sample(padding_mode="exponential"), exponential_padding,
                                  weighted_mean_n_std

Pinning: the reference has no tests or golden vectors for this path, so the
oracle is pinned against *outputs of the reference itself*: `oracle/make_golden.py` imports the
unmodified reference from /root/reference (build container only), runs it on seeded synthetic
scenes with recorded jitter, and commits the vectors under tests/golden/; tests/test_oracle.py
requires this file to reproduce them bit-for-bit (fp32, CPU).

Every function takes explicit noise so that CUDA kernels can be fed the very same jitter.  The op
sequence deliberately follows the reference one rounding at a time: nearest-neighbour taps are
discontinuous, so a single differently-rounded multiply can move a sample.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import numpy as np
import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------
# scene access (duck-typed PixelNeRF after encode())
# ----------------------------------------------------------------------------------------------

def _cams(model):
    R = model.poses[:, :, :3, :3]
    t = model.poses[:, :, :3, -1]
    return R, t


def project_to_views(model, xyz_world):
    """world points (SB,B,3) -> camera-space points (SB,NV,B,3) and normalised uv (SB,NV,B,2).
    nerf_renderer.py:99-110 == pixelnerf.py:91-93,105-108 (same chain in sampler and field)."""
    SB = xyz_world.shape[0]
    R, t = _cams(model)
    NV = R.shape[1]
    x = xyz_world.unsqueeze(1).expand(-1, NV, -1, -1)
    x = torch.matmul(R, x.transpose(-2, -1)).transpose(-2, -1) + t.unsqueeze(-2)
    uv = x[..., :2] / x[..., 2:]
    uv *= model.focal.unsqueeze(-2)
    uv += model.c.unsqueeze(-2)
    uv = uv / model.image_shape * 2 - 1
    return x, uv


def rotate_to_views(model, dirs_world):
    """(SB,B,3) -> (SB,NV,B,3): directions get the rotation only (nerf_renderer.py:102-103, pixelnerf.py:99-101)."""
    R, _ = _cams(model)
    NV = R.shape[1]
    d = dirs_world.unsqueeze(1).expand(-1, NV, -1, -1)
    return torch.matmul(R, d.transpose(-1, -2)).transpose(-1, -2)


def _tap(img, uv, mode, padding):
    """grid_sample wrapper: img (SB,NV,Ch,H,W), uv (SB,NV,N,2) -> (SB,NV,Ch,N)  (image_encoder.py:108-127)."""
    SB, NV, N, _ = uv.shape
    out = F.grid_sample(img.reshape(SB * NV, *img.shape[-3:]), uv.reshape(SB * NV, N, 1, 2),
                        align_corners=False, mode=mode, padding_mode=padding)
    return out[:, :, :, 0].view(SB, NV, -1, N)


def exp_padded_std(std_maps, pad=100, double_width=12):
    """Replicate-pad by `pad` px and grow the value by 2^(e/double_width), e = ring index counted from
    the image border (torch_helpers.py:100-122).  Closed form of the reference's 100-iteration loop:
    the loop writes ring i (rows AND columns at distance i+1 from the image) for i ascending, so the
    last write at a padded pixel is max(dy,dx)-1 where dy/dx count pixels outside the image."""
    N, C, H, W = std_maps.shape
    base = F.pad(std_maps, [pad] * 4, mode="replicate")
    yy = torch.arange(H + 2 * pad, device=std_maps.device)
    xx = torch.arange(W + 2 * pad, device=std_maps.device)
    dy = torch.clamp(torch.maximum(pad - yy, yy - (H + pad - 1)), min=0)
    dx = torch.clamp(torch.maximum(pad - xx, xx - (W + pad - 1)), min=0)
    e = torch.clamp(torch.maximum(dy[:, None], dx[None, :]) - 1, min=0).to(std_maps.dtype)
    return base * torch.exp(e / double_width * np.log(2))


def tap_depth(model, uv):          # image_encoder.py:129-151
    return _tap(model.encoder.depths, uv, "nearest", "border")


def tap_normal(model, uv):         # image_encoder.py:182-204
    return _tap(model.encoder.normals, uv, "nearest", "zeros")


def tap_depth_std(model, uv, pad=100, double_width=12):
    """image_encoder.py:153-180 -> torch_helpers.grid_sample(padding_mode="exponential", zeros beyond)."""
    std = model.encoder.depths_std
    SB, NV = std.shape[:2]
    H, W = std.shape[-2:]
    padded = exp_padded_std(std.reshape(SB * NV, *std.shape[-3:]), pad, double_width)
    img_size = torch.tensor([W, H], dtype=torch.float, device=uv.device)
    scale = img_size / (img_size + 2 * pad)
    uv_s = uv.reshape(SB * NV, -1, 1, 2) * scale.view(1, 1, 1, 2)
    out = F.grid_sample(padded, uv_s, mode="nearest", padding_mode="zeros", align_corners=False)
    return out[:, :, :, 0].view(SB, NV, -1, uv.shape[2])


def tap_latent(model, uv):
    """Bilinear/border gather of encoder features with the feature-padding rescale (image_encoder.py:97-127)."""
    lat = model.encoder.latent
    size = torch.tensor([lat.shape[-1], lat.shape[-2]], device=uv.device)
    uv = uv * ((size - model.encoder.feature_padding * 2) / size).view(1, 1, 1, 2)
    return _tap(lat, uv, model.encoder.index_interp, model.encoder.index_padding)


def posenc(code, x):
    """[x, sin(x*f + phase)] with phase = 0, pi/2 alternating (positional_encoding.py:33-53)."""
    shp = x.shape
    x = x.reshape(-1, shp[-1])
    e = x.unsqueeze(1).repeat(1, code.num_freqs * 2, 1)
    e = torch.sin(torch.addcmul(code._phases, e, code._freqs)).view(x.shape[0], -1)
    if code.include_input:
        e = torch.cat((x, e), dim=-1)
    return e.reshape(*shp[:-1], -1)


# ----------------------------------------------------------------------------------------------
# sampler  (nerf_renderer.py:39-63, 65-284)
# ----------------------------------------------------------------------------------------------

def coarse_candidates(rays, n_cand, u_coarse=None):
    """Stratified candidates z = near*(1-s) + far*s, s = linspace(0,1-1/C,C) + U/C   (nerf_renderer.py:39-63)."""
    shp = rays.shape
    r = rays.reshape(-1, 8)
    near, far = r[:, -2:-1], r[:, -1:]
    step = 1.0 / n_cand
    s = torch.linspace(0, 1 - step, n_cand, device=r.device).unsqueeze(0).repeat(r.shape[0], 1)
    if u_coarse is None:
        u_coarse = torch.rand_like(s)
    s += u_coarse.reshape(s.shape) * step
    z = near * (1 - s) + far * s
    return z.view(*shp[:-1], n_cand), u_coarse


def candidate_likelihood(model, rays, z_cand, depth_diff_max=0.05, deform=None):
    """Per-candidate surface likelihood, max over views, and its occlusion-aware version
    (nerf_renderer.py:95-132).  `deform` (NOVEL variant, novel/nerf_novel_renderer.py:104-107): a map applied to the
    candidate positions (SB,B,3) before they are projected into the source views."""
    SB, NR, C = z_cand.shape
    NV = model.poses.shape[1]
    step = (rays[..., -1] - rays[..., -2]) / C                                   # SB,NR
    xyz = rays[..., None, :3] + z_cand.unsqueeze(-1) * rays[..., None, 3:6]
    xyz = xyz.reshape(SB, -1, 3)
    if deform is not None:
        xyz = deform(xyz)
    x_cam, uv = project_to_views(model, xyz)                                     # SB,NV,B,3 / 2
    dirs_cam = rotate_to_views(model, rays[..., 3:6]).repeat_interleave(C, dim=-2)

    d_ref = tap_depth(model, uv)                                                 # SB,NV,1,B
    s_ref = tap_depth_std(model, uv)
    n_ref = tap_normal(model, uv)                                                # SB,NV,3,B
    z_ref = x_cam[..., 2:].permute(0, 1, 3, 2)
    step = step.repeat_interleave(C, dim=1).view(SB, 1, 1, NR * C).expand_as(d_ref)

    cosd = (dirs_cam.transpose(-2, -1) * n_ref).sum(dim=-2, keepdim=True)
    mask = (s_ref != 0) & ((d_ref - z_ref).abs() < depth_diff_max) & (cosd <= 0)
    p = torch.zeros_like(d_ref)
    p[mask] = 0.5 * (
        torch.special.erf((z_ref[mask] + step[mask] / 2 - d_ref[mask]) / (s_ref[mask] * np.sqrt(2))) -
        torch.special.erf((z_ref[mask] - step[mask] / 2 - d_ref[mask]) / (s_ref[mask] * np.sqrt(2)))
    ).abs()
    p = torch.max(p, dim=1).values.squeeze(1).reshape(SB, NR, C)
    p_opq = p.clone()
    p_opq[..., 1:] *= torch.cumprod(1. - p, dim=-1)[..., :-1]
    return p, p_opq


def weighted_moments(x, w):
    """torch_helpers.py:294-302 with dim=-1, keepdims=True."""
    wn = w / w.sum(dim=-1, keepdims=True)
    mean = (x * wn).sum(dim=-1, keepdims=True)
    std = ((x - mean).pow(2) * wn).sum(dim=-1, keepdims=True).sqrt()
    return mean, std


def sample_depthguided(model, rays, n_samples, n_cand, n_gaussian, depth_diff_max=0.05, noise=None, deform=None):
    """nerf_renderer.py:65-284.  Returns z (SB,NR,K) with zeros in unfilled slots plus debug info."""
    noise = dict(noise or {})
    SB, NR = rays.shape[:2]
    z_cand, noise["u_coarse"] = coarse_candidates(rays, n_cand, noise.get("u_coarse"))
    p, p_opq = candidate_likelihood(model, rays, z_cand, depth_diff_max, deform=deform)

    order = p.argsort(dim=-1, descending=True)[..., :n_samples]                 # nerf_renderer.py:172
    p_sel = torch.gather(p, -1, order)
    z_sel = torch.gather(z_cand, -1, order)
    z_sel[p_sel == 0.] = 0

    ray_mask = None
    if n_gaussian > 0:                                                          # nerf_renderer.py:181-190
        ray_mask = torch.any(p_opq != 0, dim=-1)
        mean, std = weighted_moments(z_cand[ray_mask], p_opq[ray_mask])
        g = torch.zeros(SB, NR, n_gaussian, dtype=z_cand.dtype, device=z_cand.device)
        if noise.get("n_gauss") is None and noise.get("n_gauss_full") is not None:
            # per-ray (unpacked) draws (SB*NR,G): packing-independent form used to compare torch backends
            noise["n_gauss"] = noise["n_gauss_full"].view(SB, NR, n_gaussian)[ray_mask]
        if noise.get("n_gauss") is None:
            noise["n_gauss"] = torch.randn_like(g[ray_mask])
        g[ray_mask] = noise["n_gauss"] * std + mean
        z_sel[..., -n_gaussian:] = g
    dbg = SimpleNamespace(z_cand=z_cand, p=p, p_opq=p_opq, order=order, ray_mask=ray_mask)
    return z_sel, noise, dbg


def fill_up_uniform(z, rays, u_fill=None, u_fill_full=None):
    """nerf_renderer.py:367-397: sort, fill the zero slots (now in front) with stratified uniform samples
    using the *sorted position* as stratum index, sort again.  `u_fill_full` (SB*NR,K) is the unpacked,
    per-(ray, sorted slot) form of the draws."""
    shp = rays.shape
    z = z.sort(dim=-1).values
    z = z.view(-1, z.shape[-1])
    r = rays.view(-1, 8)
    missing = z == 0
    iray, islot = torch.where(missing)
    n_missing = missing.int().sum(dim=-1)[iray]
    near, far = r[iray, -2], r[iray, -1]
    step = (far - near) / n_missing
    zm = near + islot * step
    if u_fill is None and u_fill_full is not None:
        u_fill = u_fill_full[iray, islot]
    if u_fill is None:
        u_fill = torch.rand_like(zm)
    zm += u_fill * step
    z[iray, islot] = zm
    z = z.view(*shp[:2], z.shape[-1])
    return z.sort(dim=-1).values, u_fill


# ----------------------------------------------------------------------------------------------
# field  (pixelnerf.py:55-145, resnetfc.py:129-159)
# ----------------------------------------------------------------------------------------------

def field_inputs(model, xyz, viewdirs):
    """(SB,B,3) x2 -> MLP input (SB,NV,B,567) = [latent 512 | poscode(x_cam) 39 | dir_cam 3 | depthcode 13]."""
    x_cam, uv = project_to_views(model, xyz)
    feat = posenc(model.poscode, x_cam)
    feat = torch.cat((feat, rotate_to_views(model, viewdirs)), dim=-1)
    latent = tap_latent(model, uv).transpose(-1, -2)
    d_ref = tap_depth(model, uv)
    dd = d_ref.squeeze(-2) - x_cam[..., -1]
    dfeat = posenc(model.depthcode, dd.unsqueeze(-1))
    return torch.cat((latent, feat, dfeat), dim=-1)


def resnetfc(mlp, zx):
    """zx (SB,NV,B,d_latent+d_in) -> (SB,B,d_out); mean over views (dim 1) before block `combine_layer`."""
    z, x = zx[..., :mlp.d_latent], zx[..., mlp.d_latent:]
    x = mlp.lin_in(x)
    for b in range(mlp.n_blocks):
        if b == mlp.combine_layer:
            x = torch.mean(x, dim=1)
        if b < mlp.combine_layer:
            x = x + mlp.lin_z[b](z)
        blk = mlp.blocks[b]
        net = blk.fc_0(torch.relu(x))
        x = x + blk.fc_1(torch.relu(net))
    return mlp.lin_out(torch.relu(x))


def field(model, xyz, viewdirs):
    """PixelNeRF.forward: (SB,B,3) x2 -> (SB,B,4) = [sigmoid(rgb), relu(sigma)]."""
    SB, B, _ = xyz.shape
    out = resnetfc(model.mlp_fine, field_inputs(model, xyz, viewdirs)).reshape(SB, B, 4)
    return torch.cat([torch.sigmoid(out[..., :3]), torch.relu(out[..., 3:4])], dim=-1)


# ----------------------------------------------------------------------------------------------
# compositing  (nerf_renderer.py:286-365)
# ----------------------------------------------------------------------------------------------

def alpha_composite(rgbsigma, z, rays, white_bkgd):
    """rgbsigma (SB,B,K,4), z (SB,B,K) sorted -> weights (SB,B,K), rgb (SB,B,3), depth (SB,B)."""
    deltas = torch.cat([z[..., 1:] - z[..., :-1], rays[..., -1:] - z[..., -1:]], -1)
    rgbs, sigmas = rgbsigma[..., :3], rgbsigma[..., 3]
    alphas = 1 - torch.exp(-deltas * torch.relu(sigmas))
    shifted = torch.cat([torch.ones_like(alphas[..., :1]), 1 - alphas + 1e-10], -1)
    T = torch.cumprod(shifted, -1)
    weights = alphas * T[..., :-1]
    rgb = torch.sum(weights.unsqueeze(-1) * rgbs, -2)
    depth = torch.sum(weights * z, -1)
    if white_bkgd:
        rgb = rgb + 1 - weights.sum(dim=-1).unsqueeze(-1)
    return weights, rgb, depth


def composite(model, rays, z, white_bkgd=True, eval_batch_size=100000):
    """Evaluate the field at the samples in chunks and blend (nerf_renderer.py:286-365)."""
    SB, B, K = z.shape
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, B * K, 3)
    dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, B * K, 3)
    chunk = (eval_batch_size - 1) // SB + 1
    outs = [field(model, p, d) for p, d in zip(torch.split(pts, chunk, dim=1), torch.split(dirs, chunk, dim=1))]
    rgbsigma = torch.cat(outs, dim=1).reshape(SB, B, K, -1)
    w, rgb, depth = alpha_composite(rgbsigma, z, rays, white_bkgd)
    return w, rgb, depth, rgbsigma


def render(model, rays, n_samples=40, n_cand=1000, n_gaussian=15, white_bkgd=True,
           eval_batch_size=100000, noise=None):
    """NeRFRendererDGS.forward (nerf_renderer.py:399-424).  Returns a namespace with every stage's result."""
    assert rays.dim() == 3
    with torch.no_grad():
        z0, noise, dbg = sample_depthguided(model, rays, n_samples, n_cand, n_gaussian, noise=noise)
        z, noise["u_fill"] = fill_up_uniform(z0, rays, noise.get("u_fill"))
    w, rgb, depth, rgbsigma = composite(model, rays, z, white_bkgd, eval_batch_size)
    return SimpleNamespace(rgb=rgb, depth=depth, weights=w, z=z, z_unfilled=z0, rgbsigma=rgbsigma,
                           noise=noise, dbg=dbg)
