"""TEST INFRASTRUCTURE (oracle): numpy restatement of the NOVEL variants' deform_points
(src/models/novel/nerf_novel_renderer.py:40-50): brute-force nearest vertex (the published algorithm of
pytorch3d.ops.knn_points with K = 1: squared L2 distance, first minimum), offset gather, add.

PARITY UNPINNED: pytorch3d (reference README.md pins no version; not in the build image) cannot be imported, so this
restatement is checked only by known-answer tests, not against outputs of the reference."""
import numpy as np

F = np.float32


def nearest_vertex(points, vertices):
    """points (B,3), vertices (V,3) fp32 -> (idx (B,), squared distance (B,)); distance accumulated x, y, z with the fused
    multiply-adds the CUDA kernel uses (float64 emulation of fmaf is exact here: products of fp32 fit in fp64)."""
    p, v = points.astype(F), vertices.astype(F)
    d = (p[:, None, :] - v[None, :, :]).astype(F)                       # fp32 differences
    acc = (d[..., 0] * d[..., 0]).astype(F)
    acc = (d[..., 1].astype(np.float64) * d[..., 1].astype(np.float64) + acc.astype(np.float64)).astype(F)   # fmaf
    acc = (d[..., 2].astype(np.float64) * d[..., 2].astype(np.float64) + acc.astype(np.float64)).astype(F)
    idx = acc.argmin(axis=1)                                             # first minimum = lowest index on ties
    return idx.astype(np.int32), acc[np.arange(len(p)), idx]


def deform_points(points, vertices, offsets):
    """(SB,B,3), (SB,V,3), (SB,V,3) -> (SB,B,3)."""
    out = np.empty_like(points, dtype=F)
    for sb in range(points.shape[0]):
        idx, _ = nearest_vertex(points[sb], vertices[sb])
        out[sb] = points[sb].astype(F) + offsets[sb].astype(F)[idx]
    return out
