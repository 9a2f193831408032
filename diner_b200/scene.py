"""Per-scene device buffers for the CUDA renderer.

Reads the attributes a reference `PixelNeRF` exposes after `encode()` (reference
src/models/pixelnerf.py:35-53, src/models/image_encoder.py:206-273), validates them and
repacks what the kernels want in a different layout:

  * encoder.latent  (SB,NV,512,Hl,Wl) fp32 NCHW  ->  (SB,NV,Hl,Wl,512) bf16 channels-last
  * encoder.depths_std                            ->  exponentially padded (+100 px) std map, built once per
                                                     scene instead of on every index_depth_std call
                                                     (image_encoder.py:153-180, torch_helpers.py:100-122)
  * cameras / depth / normal maps                 ->  contiguous fp32

The repack is cached on the tensors' identity and version counters, so it is redone after the next
`encode()` and not before.
"""
from __future__ import annotations

import math
import weakref
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib

STD_PAD = 100            # image_encoder.py:175
STD_DOUBLE_WIDTH = 12    # image_encoder.py:174
DEPTH_DIFF_MAX = 0.05    # nerf_renderer.py:67


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.float32).contiguous()


@dataclass
class SceneBuffers:
    SB: int
    NV: int
    H: int
    W: int
    Hl: int
    Wl: int
    L: int
    poses: torch.Tensor
    focal: torch.Tensor
    cxy: torch.Tensor
    depth: torch.Tensor
    normal: torch.Tensor
    std_padded: torch.Tensor
    latent_nhwc: torch.Tensor
    freqs: torch.Tensor
    phases: torch.Tensor
    num_freqs: int
    img_w: float
    img_h: float
    std_sx: float
    std_sy: float
    lat_sx: float
    lat_sy: float
    key: tuple
    _latent_ref: object = None         # weak reference to encoder.latent (the cache must not keep the scene alive)
    _latent_lo: torch.Tensor = None

    def latent_lo(self) -> torch.Tensor:
        """Verification mode (precision="split"): the bf16 image of latent - bf16(latent), channels-last; built on
        first use."""
        if self._latent_lo is None:
            lat = self._latent_ref() if self._latent_ref is not None else None
            if lat is None:
                raise RuntimeError("diner_b200: encoder.latent was released; call encode() again")
            lat32 = _f32c(lat).view(self.SB * self.NV, self.L, self.Hl, self.Wl)
            res = (lat32 - lat32.to(torch.bfloat16).to(torch.float32)).contiguous()
            out = torch.empty(self.SB, self.NV, self.Hl, self.Wl, self.L, device=lat32.device, dtype=torch.bfloat16)
            _lib.call("dnr_pack_latent_nhwc", _lib.ptr(res), self.SB * self.NV, self.L, self.Hl, self.Wl, _lib.ptr(out),
                      _lib.stream_ptr())
            self._latent_lo = out
        return self._latent_lo


_lin_tabs: dict = {}
_gain_tabs: dict = {}


def lin_table(C: int, device) -> torch.Tensor:
    """torch.linspace(0, 1 - 1/C, C) exactly as sample_coarse builds it (nerf_renderer.py:53-55); computed on the
    CPU so that the table is bit-identical to the reference's CPU path, then uploaded once."""
    k = (C, str(device))
    if k not in _lin_tabs:
        step = 1.0 / C
        _lin_tabs[k] = torch.linspace(0, 1 - step, C).to(device)
    return _lin_tabs[k]


# Where exp(e / 12 * ln 2) of the padded ring is evaluated.  "device": with torch ops on the scene's CUDA device, i.e. the
# bits the reference produces where it is deployed (torch's CUDA exp) -- the sampler then agrees bit for bit with the
# reference algorithm on torch CUDA ops even when the foreground reaches the image border.  "cpu": torch's CPU exp, the
# bits of the reference on the CPU (what tests/golden/exp_padding.npz holds).  The two differ in the last bit of a few
# of the 100 gains; everything else of the padding is exact either way.
GAIN_TABLE_BACKEND = "device"


def gain_table(device, pad=STD_PAD, double_width=STD_DOUBLE_WIDTH, backend=None) -> torch.Tensor:
    """exp(e / double_width * ln 2) for e = 0..pad-1, evaluated like torch_helpers.py:121 (fp32)."""
    backend = backend or GAIN_TABLE_BACKEND
    k = (pad, double_width, str(device), backend)
    if k not in _gain_tabs:
        e = torch.arange(pad, dtype=torch.float32, device=device if backend == "device" else "cpu")
        _gain_tabs[k] = torch.exp(e / double_width * np.log(2)).to(device)
    return _gain_tabs[k]


def validate_model(model) -> None:
    """Raise NotImplementedError for model variants the kernels do not implement."""
    enc, mlp = model.encoder, model.mlp_fine
    if getattr(enc, "index_interp", "bilinear") != "bilinear" or getattr(enc, "index_padding", "border") != "border":
        raise NotImplementedError("diner_b200: encoder.index must be bilinear / border")
    # NOVEL_PE (pe_novel_pixelnerf.py:21,27): d_latent = 518 behind a 518 -> 512 deformation layer (diner_b200.novel only)
    d_latent = 518 if getattr(model, "deformation_layer", None) is not None else 512
    if mlp.d_hidden != 512 or mlp.d_latent != d_latent or mlp.n_blocks != 5 or mlp.combine_layer != 3:
        raise NotImplementedError("diner_b200: ResnetFC must be 5 blocks x 512 with combine_layer 3 and d_latent 512 "
                                  "(518 with the NOVEL_PE deformation layer)")
    if getattr(mlp, "combine_type", "average") != "average":
        raise NotImplementedError("diner_b200: combine_type must be 'average'")
    if not isinstance(mlp.activation, torch.nn.ReLU):
        raise NotImplementedError("diner_b200: only ReLU activations (beta=0) are implemented")
    if mlp.d_in != 55 or mlp.d_out != 4:
        raise NotImplementedError("diner_b200: MLP must map 512+55 -> 4")
    for code, d_in in ((model.poscode, 3), (model.depthcode, 1)):
        if code.num_freqs != 6 or not code.include_input or code.d_in != d_in:
            raise NotImplementedError("diner_b200: positional encodings must use 6 frequencies with include_input")
    for blk in mlp.blocks:
        if getattr(blk, "shortcut", None) is not None:
            raise NotImplementedError("diner_b200: ResnetBlockFC shortcuts are not implemented")


_cache: dict = {}


def _key(model):
    enc = model.encoder
    ts = (enc.latent, enc.depths, enc.depths_std, enc.normals, model.poses, model.focal, model.c)
    return tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in ts)


def invalidate() -> None:
    """Forget the cached per-scene repacks AND the packed MLP weights.  The caches key on (data_ptr, _version, shape) of
    the source tensors; writers that do not bump the version counter -- `t.data.copy_`, c10d broadcasts into existing
    storage, external kernels, optimizers that step on `.data` -- leave stale repacks behind, so they must call this
    (`diner_b200.sharding.broadcast_scene` does)."""
    from . import field, novel
    _cache.clear()
    field.invalidate_caches()
    novel._gen_cache.clear()


def get_scene(model) -> SceneBuffers:
    """Device buffers for the scene currently encoded in `model` (cached until the next encode(); see invalidate())."""
    key = _key(model)
    ck = id(model)
    hit = _cache.get(ck)
    if hit is not None and hit.key == key:
        return hit
    validate_model(model)
    enc = model.encoder
    lat = enc.latent
    if not lat.is_cuda:
        raise RuntimeError("diner_b200: the scene must live on a CUDA device (there is no CPU path)")
    dev = lat.device
    SB, NV, L, Hl, Wl = lat.shape
    H, W = enc.depths.shape[-2:]
    if L != 512:
        raise NotImplementedError("diner_b200: latent must have 512 channels")
    if NV > 8:
        raise NotImplementedError("diner_b200: at most 8 source views")
    stream = _lib.stream_ptr()

    depth = _f32c(enc.depths).view(SB, NV, H, W)
    normal = _f32c(enc.normals).view(SB, NV, 3, H, W)
    std = _f32c(enc.depths_std).view(SB * NV, H, W)
    std_padded = torch.empty(SB, NV, H + 2 * STD_PAD, W + 2 * STD_PAD, device=dev, dtype=torch.float32)
    _lib.call("dnr_pad_depth_std", _lib.ptr(std), _lib.ptr(gain_table(dev)), SB * NV, H, W, STD_PAD,
              _lib.ptr(std_padded), stream)

    lat32 = _f32c(lat).view(SB * NV, L, Hl, Wl)
    latent_nhwc = torch.empty(SB, NV, Hl, Wl, L, device=dev, dtype=torch.bfloat16)
    _lib.call("dnr_pack_latent_nhwc", _lib.ptr(lat32), SB * NV, L, Hl, Wl, _lib.ptr(latent_nhwc), stream)

    # scale factors, evaluated with the reference's own tensor expressions (one fp32 rounding each)
    img_size = torch.tensor([W, H], dtype=torch.float)
    std_scale = img_size / (img_size + 2 * STD_PAD)                      # torch_helpers.py:155-158
    lat_size = torch.tensor([Wl, Hl])
    lat_scale = (lat_size - float(enc.feature_padding) * 2) / lat_size    # image_encoder.py:113-114
    image_shape = _f32c(model.image_shape).cpu()

    sc = SceneBuffers(
        SB=SB, NV=NV, H=H, W=W, Hl=Hl, Wl=Wl, L=L,
        poses=_f32c(model.poses), focal=_f32c(model.focal), cxy=_f32c(model.c),
        depth=depth, normal=normal, std_padded=std_padded, latent_nhwc=latent_nhwc,
        freqs=_f32c(model.poscode._freqs).view(-1).to(dev), phases=_f32c(model.poscode._phases).view(-1).to(dev),
        num_freqs=int(model.poscode.num_freqs),
        img_w=float(image_shape[0]), img_h=float(image_shape[1]),
        std_sx=float(std_scale[0]), std_sy=float(std_scale[1]),
        lat_sx=float(lat_scale[0].float()), lat_sy=float(lat_scale[1].float()),
        key=key, _latent_ref=weakref.ref(lat))
    if sc.poses.shape != (SB, NV, 4, 4):
        raise ValueError(f"model.poses must be (SB,NV,4,4), got {tuple(sc.poses.shape)}")
    _cache.clear()
    _cache[ck] = sc
    return sc
