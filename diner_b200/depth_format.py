"""On-disk depth format of the reference's datasets, decoded on the device.

    read_depth(tri_u16, mesh_u16, depth_type)          FacescapeDataSet.read_depth      src/data/facescape.py:69-106
    conf2std laws                                       facescape.py:54-56, dtu.py:68-70, multiface.py:306-311
    decode_facescape / decode_dtu / decode_multiface    read_depth + conf2std of one sample, both maps in one kernel

Inputs are the *decoded* uint16 PNG planes (the host image library inflates the files: `load_u16_png`); the CUDA
kernel (csrc/depth_format.cu, `dnr_decode_depth_u16`) turns them into the fp32 `src_depths` / `src_depth_stds` maps the
encoder consumes, with the reference's fp32 operation order.  CPU tensors raise: there is no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

SCALE_FACTOR = 1e-4                                       # facescape.py:72, dtu.py:105, multiface.py:105
FACESCAPE_CONF2STD = (-1.582e-2, 1.649e-2)                # facescape.py:54-56
DTU_CONF2STD = (-2.5679e-2, 3.2818e-2)                    # dtu.py:68-70
MULTIFACE_CONF2STD = (-1.582e-2, 1.649e-2)                # multiface.py:310
DEPTH_TYPES = {"original": 0, "mesh": 1, "merge": 2}
FLAG_CLIP0, FLAG_ZERO_NO_DEPTH, FLAG_CONF_SCALED, FLAG_STD_CONST = 1, 2, 4, 8


def load_u16_png(path) -> torch.Tensor:
    """Inflate a 16-bit PNG on the host (PIL) -> uint16 tensor (H, W) in pinned memory."""
    from PIL import Image
    a = np.array(Image.open(path))
    if a.dtype != np.uint16:
        if a.dtype.kind not in "iu" or a.min() < 0 or a.max() > 65535:
            raise ValueError(f"{path}: not a 16-bit depth image (dtype {a.dtype})")
        a = a.astype(np.uint16)
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.pin_memory() if torch.cuda.is_available() else t


def _u16(t, name):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (diner_b200 has no CPU path)")
    if t.dtype not in (torch.uint16, torch.int16):
        raise ValueError(f"{name} must hold uint16 pixels (torch.uint16, or torch.int16 with the same bits), got {t.dtype}")
    if t.dim() == 2:
        t = t[None]
    if t.dim() != 3:
        raise ValueError(f"{name} must have shape (N, H, W) or (H, W)")
    return t.contiguous()


def _decode(pred, conf, img_stride, row_stride, mesh, N, H, W, depth_type, div0, mul1, a, b, std_const, flags, dev):
    depth = torch.empty(N, 1, H, W, device=dev, dtype=torch.float32)
    std = torch.empty(N, 1, H, W, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        _lib.call("dnr_decode_depth_u16", pred, conf, img_stride, row_stride, _lib.ptr(mesh), N, H, W, depth_type,
                  SCALE_FACTOR, div0, mul1, a, b, std_const, flags, _lib.ptr(depth), _lib.ptr(std), _lib.stream_ptr())
    return depth, std


def decode_facescape(tri_u16, mesh_u16=None, depth_type: str = "original", conf2std=FACESCAPE_CONF2STD):
    """Triptych images (N,H,3W) uint16 [gt | pred | conf] (+ mesh depth images (N,H,W) for 'mesh' / 'merge') ->
    (src_depths (N,1,H,W), src_depth_stds (N,1,H,W)) = stack(read_depth(...)), conf2std(stack(...)) of
    facescape.py:257-266.  `conf2std=None` returns the confidence itself (what read_depth returns)."""
    if depth_type not in DEPTH_TYPES:
        raise ValueError(f"depth type should be in original, mesh or merge but got: {depth_type}")      # facescape.py:105
    tri = _u16(tri_u16, "tri_u16")
    mesh = _u16(mesh_u16, "mesh_u16")
    N, H, W3 = tri.shape
    if W3 % 3:
        raise ValueError("triptych width must be a multiple of 3")
    W = W3 // 3
    if DEPTH_TYPES[depth_type] and (mesh is None or tuple(mesh.shape) != (N, H, W)):
        raise ValueError(f"depth_type '{depth_type}' needs mesh depth images of shape {(N, H, W)}")
    a, b = conf2std if conf2std is not None else (1.0, 0.0)
    base = tri.data_ptr()
    return _decode(base + 2 * W, base + 4 * W, H * W3, W3, mesh, N, H, W, DEPTH_TYPES[depth_type], 1.0, 1.0, a, b, 0.0,
                   0, tri.device)


def read_depth(tri_u16, mesh_u16=None, depth_type: str = "original"):
    """FacescapeDataSet.read_depth on decoded planes of ONE view: returns (pred (1,H,W), conf (1,H,W))."""
    d, c = decode_facescape(tri_u16, mesh_u16, depth_type, conf2std=None)
    return d[0], c[0]


def decode_dtu(depth_u16, conf_u16, scale_factor: float, conf2std=DTU_CONF2STD):
    """TransMVSNet PNG predictions of DTU (N,H,W) uint16 -> (depths (N,1,H,W), depth_stds (N,1,H,W)):
    `u16 * 1e-4 / (0.7 / 872) * scale_factor` for the depth AND for the confidence image (both go through
    DTUDataSet.read_depth, dtu.py:104-107,119,221), then conf2std (dtu.py:223).  The data loader runs on the CPU, where
    torch divides by the scalar in fp32 (a true division, not a reciprocal multiply); so does the kernel."""
    dep, conf = _u16(depth_u16, "depth_u16"), _u16(conf_u16, "conf_u16")
    if dep.shape != conf.shape:
        raise ValueError("depth and confidence images must have the same shape")
    N, H, W = dep.shape
    a, b = conf2std
    return _decode(dep.data_ptr(), conf.data_ptr(), H * W, W, None, N, H, W, 0, float(np.float32(0.7 / 872.0)), float(np.float32(scale_factor)),
                   a, b, 0.0, FLAG_CONF_SCALED, dep.device)


def decode_multiface(depth_u16, conf_u16=None, conf2std=MULTIFACE_CONF2STD):
    """Multiface depth images (N,H,W) uint16 -> (depths, depth_stds): constant std 1e-3 without a confidence image,
    else clip(a * conf + b, min=0); std = 0 where depth == 0 (multiface.py:305-311)."""
    dep, conf = _u16(depth_u16, "depth_u16"), _u16(conf_u16, "conf_u16")
    N, H, W = dep.shape
    if conf is not None and conf.shape != dep.shape:
        raise ValueError("depth and confidence images must have the same shape")
    a, b = conf2std
    flags = FLAG_ZERO_NO_DEPTH | (FLAG_CLIP0 if conf is not None else FLAG_STD_CONST)
    return _decode(dep.data_ptr(), conf.data_ptr() if conf is not None else None, H * W, W, None, N, H, W, 0, 1.0, 1.0,
                   a, b, 1e-3, flags, dep.device)
