"""Nearest-vertex deformation of sample points on the device (first piece).

    deform_points(points, target_vertices, offsets)      NeRFRendererDGS.deform_points of the NOVEL renderer variants,
                                                          src/models/novel/nerf_novel_renderer.py:40-50

The reference finds the nearest target-mesh vertex of every point with `pytorch3d.ops.knn_points(K=1)`, gathers that
vertex's offset and adds it.  `csrc/deform.cu` (`dnr_deform_points`) does the three steps in one kernel.  The result is
differentiable like the reference's: d points = d out, d offsets[v] = sum of d out over the points whose nearest vertex is
v (the index itself carries no gradient).  Only this primitive of the NOVEL variants is built; pytorch3d is absent from the
build image, so parity is checked against the numpy restatement (oracle/deform_oracle.py), not against the reference.
"""
from __future__ import annotations

import torch

from . import _lib


def _check(points, target_vertices, offsets):
    for name, t in (("points", points), ("target_vertices", target_vertices), ("offsets", offsets)):
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise ValueError(f"{name} must be a CUDA tensor (diner_b200 has no CPU path)")
        if t.dim() != 3 or t.shape[-1] != 3:
            raise ValueError(f"{name} must have shape (SB, n, 3), got {tuple(t.shape)}")
    SB, B, _ = points.shape
    if target_vertices.shape[0] != SB or offsets.shape != target_vertices.shape:
        raise ValueError("target_vertices and offsets must have shape (SB, NV, 3) with the batch size of points")
    if target_vertices.shape[1] < 1 and B > 0:
        raise ValueError("target_vertices must hold at least one vertex")
    return SB, B, target_vertices.shape[1]


def nearest_vertex(points, target_vertices, offsets=None):
    """(deformed (SB,B,3), idx (SB,B) int32, squared distance (SB,B)) without autograd."""
    if offsets is None:
        offsets = torch.zeros_like(target_vertices)
    SB, B, V = _check(points, target_vertices, offsets)
    f = lambda t: t.detach().to(torch.float32).contiguous()
    pts, vts, off = f(points), f(target_vertices), f(offsets)
    out = torch.empty(SB, B, 3, device=pts.device, dtype=torch.float32)
    idx = torch.empty(SB, B, device=pts.device, dtype=torch.int32)
    dist = torch.empty(SB, B, device=pts.device, dtype=torch.float32)
    with torch.cuda.device(pts.device):
        _lib.call("dnr_deform_points", _lib.ptr(pts), _lib.ptr(vts), _lib.ptr(off), SB, B, V, _lib.ptr(out),
                  _lib.ptr(idx), _lib.ptr(dist), _lib.stream_ptr())
    return out, idx, dist


class _DeformFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, target_vertices, offsets):
        out, idx, _ = nearest_vertex(points, target_vertices, offsets)
        ctx.save_for_backward(idx)
        ctx.n_vert = target_vertices.shape[1]
        return out

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        d_points = g if ctx.needs_input_grad[0] else None
        d_offsets = None
        if ctx.needs_input_grad[2]:
            SB, B, _ = g.shape
            d_offsets = torch.zeros(SB, ctx.n_vert, 3, device=g.device, dtype=g.dtype)
            d_offsets.scatter_add_(1, idx.long().unsqueeze(-1).expand(SB, B, 3), g.contiguous())
        return d_points, None, d_offsets


def deform_points(points, target_vertices, offsets):
    """points (SB,B,3), target_vertices (SB,NV,3), offsets (SB,NV,3) -> (SB,B,3): every point moved by the offset of its
    nearest vertex (the reference's signature and result, nerf_novel_renderer.py:40-50)."""
    _check(points, target_vertices, offsets)
    return _DeformFn.apply(points, target_vertices, offsets)
