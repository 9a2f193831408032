"""Precision-contract oracle for the tensor-core field MLP -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

`oracle/dgs_oracle.py` restates the reference in fp32.  The CUDA path evaluates ResnetFC with bf16 operands
and fp32 accumulation (DESIGN.md "Precision contract"):

    * encoder.latent is stored bf16 (channels-last); the bilinear gather interpolates in fp32 and rounds the
      gathered feature row to bf16; the 55-wide geometric code is computed in fp32 and rounded to bf16
    * every nn.Linear weight is rounded to bf16, biases stay fp32, products accumulate in fp32
    * the residual stream x stays fp32; the GEMM inputs relu(x) and relu(fc_0(.)) are rounded to bf16
    * lin_out runs in fp32 on the bf16-rounded relu(x)

This module is the same algorithm as dgs_oracle.field() (reference src/models/pixelnerf.py:55-145,
src/models/resnetfc.py:129-159) with exactly those roundings inserted (straight-through for autograd), in
plain torch ops on any device.  Tests use it in two steps: CUDA kernels vs. this (tight tolerance: only the
accumulation order differs), and this vs. the fp32 oracle (the stated bf16 tolerance).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import dgs_oracle as O


def r16(t: torch.Tensor) -> torch.Tensor:
    """Round to bf16 (nearest-even), keep fp32 storage; identity gradient."""
    return t + (t.to(torch.bfloat16).to(t.dtype) - t).detach()


def _lin(x, layer, bias=True):
    y = F.linear(x, r16(layer.weight))
    return y + layer.bias if bias else y


def field_bf16(model, xyz, viewdirs):
    """(SB,B,3) x2 -> (SB,B,4) = [sigmoid(rgb), relu(sigma)] under the CUDA path's precision contract."""
    SB, B, _ = xyz.shape
    mlp = model.mlp_fine
    lat_fp32 = model.encoder.latent
    try:
        model.encoder.latent = r16(lat_fp32)
        zx = O.field_inputs(model, xyz, viewdirs)               # SB,NV,B,567
    finally:
        model.encoder.latent = lat_fp32
    lat, code = r16(zx[..., :mlp.d_latent]), r16(zx[..., mlp.d_latent:])
    x = _lin(code, mlp.lin_in)
    for b in range(mlp.n_blocks):
        if b == mlp.combine_layer:
            x = torch.mean(x, dim=1)
        if b < mlp.combine_layer:
            x = x + _lin(lat, mlp.lin_z[b])
        blk = mlp.blocks[b]
        net = _lin(r16(torch.relu(x)), blk.fc_0)
        x = x + _lin(r16(torch.relu(net)), blk.fc_1)
    out = F.linear(r16(torch.relu(x)), mlp.lin_out.weight, mlp.lin_out.bias).reshape(SB, B, 4)
    return torch.cat([torch.sigmoid(out[..., :3]), torch.relu(out[..., 3:4])], dim=-1)


def render_bf16(model, rays, z, white_bkgd=True):
    """Field (bf16 contract) + compositing on given sorted sample depths z (SB,NR,K)."""
    SB, NR, K = z.shape
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
    dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
    rgbsigma = field_bf16(model, pts, dirs).reshape(SB, NR, K, 4)
    w, rgb, depth = O.alpha_composite(rgbsigma, z, rays, white_bkgd)
    return w, rgb, depth, rgbsigma
