"""TEST INFRASTRUCTURE (oracle): numpy restatement of the reference's on-disk depth decoding.  Only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline leg may import this; the product path never does.

Follows FacescapeDataSet.read_depth + conf2std (src/data/facescape.py:54-56,69-106,266), the PNG branch of
DTUDataSet.read_depth + conf2std (src/data/dtu.py:68-70,104-107,119,221-223) and the Multiface law
(src/data/multiface.py:103-109,305-311).  Every step is one fp32 rounding, as torch does it.  Pinned bit-for-bit to
outputs of the unmodified reference functions (oracle/make_golden_depth_format.py -> tests/golden/depth_format.npz).
"""
import numpy as np

F = np.float32
SCALE = F(1e-4)


def read_depth_facescape(tri_u16, mesh_u16, depth_type):
    """tri (H,3W) uint16, mesh (H,W) uint16 -> (pred (1,H,W), conf (1,H,W)) fp32; facescape.py:69-106."""
    H, W3 = tri_u16.shape
    W = W3 // 3
    pred_mesh = mesh_u16.astype(F) * SCALE if mesh_u16 is not None else None              # :76
    conf_mesh = np.where(pred_mesh == 0, F(0), F(0.8)).astype(F) if mesh_u16 is not None else None   # :77-79
    pred_mvs = tri_u16[:, W:2 * W].astype(F) * SCALE                                      # :88
    conf_mvs = tri_u16[:, 2 * W:].astype(F) * SCALE                                       # :89
    if depth_type == "original":
        return pred_mvs[None], conf_mvs[None]
    if depth_type == "mesh":
        return pred_mesh[None], conf_mesh[None]
    if depth_type == "merge":                                                             # :95-101
        pred = np.where((pred_mesh == 0) & (pred_mvs != 0), pred_mvs, pred_mesh)
        conf = np.where((conf_mesh == 0) & (conf_mvs != 0), conf_mvs, conf_mesh)
        return pred[None].astype(F), conf[None].astype(F)
    raise ValueError(f"depth type should be in original, mesh or merge but got: {depth_type}")


def conf2std(conf, a, b):
    """`a * x + b` on an fp32 tensor with Python-float coefficients: fp32 multiply, then fp32 add."""
    return (F(a) * conf.astype(F)).astype(F) + F(b)


def read_depth_dtu_png(u16, scale_factor):
    """dtu.py:104-107,119: u16 * 1e-4, `/= (0.7 / 872.)` (torch CPU: a true fp32 division), `*= scale_factor`."""
    return ((u16.astype(F) * SCALE) / F(0.7 / 872.0)) * F(scale_factor)


def multiface_std(depth, conf, a=-1.582e-2, b=1.649e-2):
    """multiface.py:306-311: constant 1e-3 without a confidence image, else clip(a * conf + b, min=0); 0 where depth == 0."""
    std = np.full_like(depth, F(1e-3)) if conf is None else np.maximum(conf2std(conf, a, b), F(0))
    std = std.copy()
    std[depth == 0] = 0
    return std
