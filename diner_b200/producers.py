"""Per-frame / per-scene input producers on the device, with the reference's own function signatures
(SURVEY.md 8f rows 1-2):

    gen_rays(extrinsics, intrinsics, W, H, z_near, z_far)   src/util/cam_geometry.py:36-79
    gen_rays_at(..., pix_idcs)                               the same for selected pixels (src/models/diner.py:222-258)
    depth2normal(dmap, K)                                    src/util/depth2normal.py:6-87

Both call the CUDA kernels in csrc/producers.cu through the C ABI; CPU tensors raise (there is no CPU path).
"""
from __future__ import annotations

import torch

from . import _lib


def _dev_f32(t: torch.Tensor, name: str, shape_tail) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (diner_b200 has no CPU path)")
    if tuple(t.shape[1:]) != tuple(shape_tail):
        raise ValueError(f"{name} must have shape (B, {', '.join(map(str, shape_tail))}), got {tuple(t.shape)}")
    return t.detach().to(torch.float32).contiguous()


def _camera_args(extrinsics, intrinsics, W, H, z_near, z_far):
    E = _dev_f32(extrinsics, "extrinsics", (4, 4))
    K = _dev_f32(intrinsics, "intrinsics", (3, 3))
    B = E.shape[0]
    if K.shape[0] != B:
        raise ValueError("extrinsics and intrinsics must have the same batch size")
    W, H = int(W), int(H)
    if W < 1 or H < 1:
        raise ValueError("W and H must be positive")
    zn = torch.as_tensor(z_near, dtype=torch.float32, device=E.device).reshape(-1).contiguous()
    zf = torch.as_tensor(z_far, dtype=torch.float32, device=E.device).reshape(-1).contiguous()
    if zn.numel() != B or zf.numel() != B:
        raise ValueError("z_near and z_far must hold one value per camera")
    return E, K, B, W, H, zn, zf


def gen_rays(extrinsics, intrinsics, W, H, z_near, z_far) -> torch.Tensor:
    """(B,4,4) world->cam extrinsics, (B,3,3) intrinsics, z_near / z_far (B,) -> rays (B,H,W,8) =
    [origin(3), direction(3), near, far] through the pixel centres (OpenCV convention)."""
    E, K, B, W, H, zn, zf = _camera_args(extrinsics, intrinsics, W, H, z_near, z_far)
    rays = torch.empty(B, H, W, 8, device=E.device, dtype=torch.float32)
    with torch.cuda.device(E.device):
        _lib.call("dnr_gen_rays", _lib.ptr(E), _lib.ptr(K), B, W, H, _lib.ptr(zn), _lib.ptr(zf), _lib.ptr(rays),
                  _lib.stream_ptr())
    return rays


def gen_rays_at(extrinsics, intrinsics, W, H, z_near, z_far, pix_idcs, check_bounds: bool = True) -> torch.Tensor:
    """Rays of selected pixels: `gen_rays(...).view(B, H*W, 8)[arange(B)[:, None], pix_idcs]` (the gather of
    DINER.calc_losses, diner.py:256-257) without materialising the full-frame ray tensor.  pix_idcs (B, n) integer flat
    pixel indices y*W+x.  `check_bounds` reads one flag back (a sync) and raises IndexError like the gather would."""
    E, K, B, W, H, zn, zf = _camera_args(extrinsics, intrinsics, W, H, z_near, z_far)
    if not isinstance(pix_idcs, torch.Tensor) or pix_idcs.dim() != 2 or pix_idcs.shape[0] != B:
        raise ValueError("pix_idcs must have shape (B, n)")
    if pix_idcs.dtype.is_floating_point or pix_idcs.dtype == torch.bool:
        raise IndexError("pix_idcs must be an integer tensor")
    pix = pix_idcs.to(device=E.device, dtype=torch.int64).contiguous()
    n = pix.shape[1]
    rays = torch.empty(B, n, 8, device=E.device, dtype=torch.float32)
    bad = torch.zeros(1, device=E.device, dtype=torch.int32) if check_bounds else None
    with torch.cuda.device(E.device):
        _lib.call("dnr_gen_rays_at", _lib.ptr(E), _lib.ptr(K), B, W, H, _lib.ptr(zn), _lib.ptr(zf), _lib.ptr(pix), n,
                  _lib.ptr(rays), _lib.ptr(bad), _lib.stream_ptr())
    if check_bounds and int(bad.item()):
        raise IndexError(f"pix_idcs out of range for an image of {H}x{W} pixels")
    return rays


@torch.no_grad()
def depth2normal(dmap, K) -> torch.Tensor:
    """dmap (N,1,H,W), K (N,3,3) -> normals (N,3,H,W): central differences of the back-projected points, normals of
    pixels next to a hole taken one step away from it, zero where there is no depth.  Identical to the reference on
    the same inputs -- including N == 3, where the reference's `torch.cross` without `dim` runs over the batch axis."""
    if not isinstance(dmap, torch.Tensor) or dmap.dim() != 4 or dmap.shape[1] != 1:
        raise ValueError("dmap must have shape (N, 1, H, W)")
    N, _, H, W = dmap.shape
    d = _dev_f32(dmap, "dmap", (1, H, W))
    k = _dev_f32(K, "K", (3, 3))
    if k.shape[0] != N:
        raise ValueError("dmap and K must have the same batch size")
    if H == 3 or W == 3:
        raise NotImplementedError("H == 3 or W == 3 changes the axis of the reference's torch.cross; not supported")
    out = torch.empty(N, 3, H, W, device=d.device, dtype=torch.float32)
    with torch.cuda.device(d.device):
        _lib.call("dnr_depth2normal", _lib.ptr(d), _lib.ptr(k), N, H, W, _lib.ptr(out), _lib.stream_ptr())
    return out
