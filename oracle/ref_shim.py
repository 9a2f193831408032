"""Import the UNMODIFIED reference renderer from /root/reference (build container only).

Test infrastructure: used by oracle/make_golden.py and by the optional `-m "not gpu"` cross-checks
that skip when /root/reference is absent (it never exists on the GPU box).  The hot-path modules
import three packages that are not installed here and never touch arithmetic (`dotmap`,
`matplotlib.pyplot`, `imageio`); they are stubbed in sys.modules before the import.
"""
import os
import sys
import types
from contextlib import contextmanager

REF_ROOT = os.environ.get("DINER_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "src", "models", "nerf_renderer.py"))


class _AttrDict(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def _install_stubs():
    if "dotmap" not in sys.modules:
        m = types.ModuleType("dotmap")
        m.DotMap = _AttrDict
        sys.modules["dotmap"] = m
    for name in ("matplotlib", "matplotlib.pyplot", "imageio"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    if not hasattr(sys.modules["matplotlib"], "pyplot"):
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]


def load():
    """Returns a namespace with the reference classes."""
    if not available():
        raise RuntimeError(f"reference sources not found under {REF_ROOT}")
    _install_stubs()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from src.models.nerf_renderer import NeRFRendererDGS
    from src.models.pixelnerf import PixelNeRF
    from src.models.resnetfc import ResnetFC
    from src.models.image_encoder import SpatialEncoder
    from src.models.positional_encoding import PositionalEncoding
    from src.util.cam_geometry import gen_rays
    return types.SimpleNamespace(NeRFRendererDGS=NeRFRendererDGS, PixelNeRF=PixelNeRF, ResnetFC=ResnetFC,
                                 SpatialEncoder=SpatialEncoder, PositionalEncoding=PositionalEncoding,
                                 gen_rays=gen_rays, DotMap=_AttrDict)


def knn_points_k1(p1, p2, K=1, **_):
    """Stand-in for `pytorch3d.ops.knn.knn_points(p1, p2, K=1)` (pytorch3d is absent from the build image): the published
    semantics -- squared L2 distance, the FIRST minimum (lowest index) on ties -- returning (dists (N,P1,1),
    idx (N,P1,1) int64, None); arithmetic of oracle/novel_oracle.nearest_vertex."""
    import torch
    from oracle.novel_oracle import nearest_vertex
    assert K == 1
    res = [nearest_vertex(a, b) for a, b in zip(p1, p2)]
    return torch.stack([d for _, d in res]).unsqueeze(-1), torch.stack([i for i, _ in res]).unsqueeze(-1), None


def load_novel():
    """The reference's NOVEL renderer variant and its PixelNeRF (src/models/novel/), UNMODIFIED, with
    `pytorch3d.ops.knn.knn_points` replaced by knn_points_k1 (the one dependency that cannot be imported here)."""
    if not available():
        raise RuntimeError(f"reference sources not found under {REF_ROOT}")
    _install_stubs()
    for name in ("pytorch3d", "pytorch3d.ops", "pytorch3d.ops.knn"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["pytorch3d.ops.knn"].knn_points = knn_points_k1
    sys.modules["pytorch3d.ops"].knn = sys.modules["pytorch3d.ops.knn"]
    sys.modules["pytorch3d.ops"].knn_points = knn_points_k1
    sys.modules["pytorch3d"].ops = sys.modules["pytorch3d.ops"]
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from src.models.novel.nerf_novel_renderer import NeRFRendererDGS
    from src.models.novel.novel_pixelnerf import PixelNeRF
    return types.SimpleNamespace(NeRFRendererDGS=NeRFRendererDGS, PixelNeRF=PixelNeRF, DotMap=_AttrDict)


def load_novel_pe():
    """The reference's NOVEL_PE renderer and PixelNeRF (src/models/novel_pe/), UNMODIFIED, knn_points replaced as in
    load_novel()."""
    load_novel()
    from src.models.novel_pe.pe_nerf_novel_renderer import NeRFRendererDGS
    from src.models.novel_pe.pe_novel_pixelnerf import PixelNeRF
    return types.SimpleNamespace(NeRFRendererDGS=NeRFRendererDGS, PixelNeRF=PixelNeRF, DotMap=_AttrDict)


@contextmanager
def record_noise(store: dict):
    """Record the three random tensors the reference draws, in order:
    rand_like (SB*NR,C) -> randn_like (n_masked,G) -> rand_like (N_missing,)."""
    import torch
    rl, rnl = torch.rand_like, torch.randn_like
    calls = []

    def rand_like(x, *a, **k):
        out = rl(x, *a, **k)
        calls.append(("rand", out.clone()))
        return out

    def randn_like(x, *a, **k):
        out = rnl(x, *a, **k)
        calls.append(("randn", out.clone()))
        return out

    torch.rand_like, torch.randn_like = rand_like, randn_like
    try:
        yield store
    finally:
        torch.rand_like, torch.randn_like = rl, rnl
        rands = [t for kind, t in calls if kind == "rand"]
        randns = [t for kind, t in calls if kind == "randn"]
        store["u_coarse"] = rands[0] if rands else None
        store["u_fill"] = rands[1] if len(rands) > 1 else None
        store["n_gauss"] = randns[0] if randns else None
