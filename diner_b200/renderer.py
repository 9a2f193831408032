"""Drop-in replacement for the reference renderer plugin.

Select it by changing one line of the reference's YAML config
(configs/train_diner_facescape.yaml:61-66, loaded by src/models/diner.py:48 through import_obj):

    renderer:
      module: diner_b200.NeRFRendererDGS          # was: src.models.nerf_renderer.NeRFRendererDGS
      kwargs: {n_samples: 40, n_depth_candidates: 1000, n_gaussian: 15}

Same constructor, same `forward(model, rays, want_weights=False)` contract and result object
(`out.fine.rgb`, `out.fine.depth`, `out.fine.weights`) as reference src/models/nerf_renderer.py:12-430, same
autograd contract (differentiable w.r.t. `model.mlp_fine` parameters and `model.encoder.latent`; the sampler is
not differentiated, nerf_renderer.py:65).  All arithmetic runs in the hand-written CUDA kernels behind the C ABI
of include/diner_b200.h; there is no CPU or torch-op fallback.
"""
from __future__ import annotations

import torch

from . import _lib, field as F_
from .scene import DEPTH_DIFF_MAX, STD_PAD, SceneBuffers, get_scene, lin_table


class RenderResult(dict):
    """Attribute-access dict (stands in for dotmap.DotMap, nerf_renderer.py:421-430)."""
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def sample_rays(sc: SceneBuffers, rays: torch.Tensor, K: int, C: int, G: int, noise: dict | None = None,
                seed: int | None = None, want_debug: bool = False):
    """Depth-guided sampling + fill-up + sort for rays (SB,NR,8) -> z (SB*NR, K).

    noise=None: in-kernel Philox jitter (no host sync).  noise=dict(u_coarse (SB*NR,C), n_gauss (n_masked,G)
    packed, u_fill (n_missing,) packed): parity mode, consumes the reference's own random draws in the reference's
    packing order."""
    p, st = _lib.ptr, _lib.stream_ptr()
    SB, NR, _ = rays.shape
    dev = rays.device
    n_rays = SB * NR
    z = torch.empty(n_rays, K, device=dev, dtype=torch.float32)
    dbg = {}
    ray_mask = n_missing = sel = p_cand = None
    if want_debug:
        p_cand = torch.empty(n_rays, C, device=dev, dtype=torch.float32)
        ray_mask = torch.empty(n_rays, device=dev, dtype=torch.int32)
        n_missing = torch.empty(n_rays, device=dev, dtype=torch.int32)
        sel = torch.empty(n_rays, K - G, device=dev, dtype=torch.int32)
        dbg = dict(ray_mask=ray_mask, n_missing=n_missing, sel_idx=sel, p=p_cand)
    lin = lin_table(C, dev)
    if seed is None:
        seed = int(torch.empty((), dtype=torch.int64).random_().item())

    def launch(u_coarse, n_gauss, u_fill, goff, foff, count_only, rm, nm, sl, pc=None, unpacked=False):
        _lib.call("dnr_sampler_fwd", p(rays), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth), p(sc.std_padded),
                  p(sc.normal), p(lin), p(u_coarse), p(n_gauss), p(u_fill), p(goff), p(foff), seed & (2 ** 64 - 1),
                  SB, sc.NV, NR, sc.H, sc.W, C, K, G, STD_PAD, DEPTH_DIFF_MAX, sc.img_w, sc.img_h, sc.std_sx, sc.std_sy,
                  1 if count_only else 0, 1 if unpacked else 0, p(z), p(rm), p(nm), p(sl), p(pc), st)

    if noise is None:
        launch(None, None, None, None, None, False, ray_mask, n_missing, sel, p_cand)
    else:
        f = lambda t: None if t is None else t.to(device=dev, dtype=torch.float32).contiguous()
        u_coarse, n_gauss, u_fill = f(noise.get("u_coarse")), f(noise.get("n_gauss")), f(noise.get("u_fill"))
        if noise.get("n_gauss_full") is not None or noise.get("u_fill_full") is not None:
            # unpacked draws: (SB*NR, G) and (SB*NR, K), indexed by ray and sorted slot (packing-independent)
            launch(u_coarse, f(noise.get("n_gauss_full")), f(noise.get("u_fill_full")), None, None, False, ray_mask,
                   n_missing, sel, p_cand, unpacked=True)
            return (z, dbg) if want_debug else z
        i32 = lambda t: None if t is None else t.to(device=dev, dtype=torch.int32).contiguous()
        goff, foff = i32(noise.get("gauss_off")), i32(noise.get("fill_off"))   # explicit slices (tests)
        if goff is None and foff is None and (n_gauss is not None or u_fill is not None):
            # packed draws: a counting pass tells where each ray's slice starts
            rm = torch.empty(n_rays, device=dev, dtype=torch.int32)
            nm = torch.empty(n_rays, device=dev, dtype=torch.int32)
            launch(u_coarse, None, None, None, None, True, rm, nm, None)
            goff = (torch.cumsum(rm, 0) - rm).to(torch.int32)
            foff = (torch.cumsum(nm, 0) - nm).to(torch.int32)
        # the reference's packing depends on its own ray_mask / missing counts; should a ray disagree (a flipped
        # texel), the slices shift -- pad so that no slice can run past the end of the recorded draws
        if n_gauss is not None:
            n_gauss = torch.cat([n_gauss.view(-1, max(G, 1)), torch.zeros(n_rays + 1, max(G, 1), device=dev)], 0)
        if u_fill is not None:
            u_fill = torch.cat([u_fill.view(-1), torch.zeros((n_rays + 1) * K, device=dev)], 0)
        launch(u_coarse, n_gauss, u_fill, goff, foff, False, ray_mask, n_missing, sel, p_cand)
    return (z, dbg) if want_debug else z


class _FieldFn(torch.autograd.Function):
    """rgbsigma = field(rays, z); differentiable w.r.t. encoder.latent and the 30 MLP parameter tensors."""

    @staticmethod
    def forward(ctx, sc, mlp, names, rays, z, NR, ray0, latent, *params):
        need_grad = any(ctx.needs_input_grad[7:])
        pk = F_.get_packed(mlp, need_transposed=need_grad)
        out, saved = F_.field_forward(sc, pk, rays, z, NR, ray0, save=need_grad)
        ctx.sc, ctx.mlp, ctx.names, ctx.pk, ctx.saved = sc, mlp, names, pk, saved
        ctx.rays, ctx.z, ctx.NR, ctx.ray0 = rays, z, NR, ray0
        ctx.latent_shape = latent.shape
        return out

    @staticmethod
    def backward(ctx, d_out):
        sc, mlp = ctx.sc, ctx.mlp
        dev = d_out.device
        params = dict(mlp.named_parameters())
        # one zero fill for all parameter gradients (views into a flat buffer, 16-byte aligned starts)
        offs, total = {}, 0
        for n in ctx.names:
            offs[n] = total
            total += (params[n].numel() + 3) // 4 * 4
        flat = torch.zeros(total, device=dev, dtype=torch.float32)
        grads = {n: flat[offs[n]:offs[n] + params[n].numel()].view(params[n].shape) for n in ctx.names}
        d_lat_nhwc = None
        if ctx.needs_input_grad[7]:
            d_lat_nhwc = torch.zeros(sc.SB, sc.NV, sc.Hl, sc.Wl, sc.L, device=dev, dtype=torch.float32)
        F_.field_backward(sc, ctx.pk, mlp, ctx.saved, ctx.rays, ctx.z, ctx.NR, ctx.ray0, d_out, grads, d_lat_nhwc)
        d_latent = None
        if d_lat_nhwc is not None:
            d_latent = torch.empty(ctx.latent_shape, device=dev, dtype=torch.float32)
            _lib.call("dnr_unpack_dlatent_nchw", _lib.ptr(d_lat_nhwc), sc.SB * sc.NV, sc.L, sc.Hl, sc.Wl,
                      _lib.ptr(d_latent), _lib.stream_ptr())
        ctx.saved = None
        pg = tuple(grads[n].to(params[n].dtype) if ctx.needs_input_grad[8 + i] else None
                   for i, n in enumerate(ctx.names))
        return (None, None, None, None, None, None, None, d_latent, *pg)


class _CompositeFn(torch.autograd.Function):
    """(rgb, depth, weights) = composite(rgbsigma, z, rays)   -- nerf_renderer.py:344-365."""

    @staticmethod
    def forward(ctx, rgbsigma, z, rays, white_bkgd):
        n, K = z.shape
        dev = z.device
        weights = torch.empty(n, K, device=dev, dtype=torch.float32)
        rgb = torch.empty(n, 3, device=dev, dtype=torch.float32)
        depth = torch.empty(n, device=dev, dtype=torch.float32)
        p = _lib.ptr
        _lib.call("dnr_composite_fwd", p(rgbsigma), p(z), p(rays), n, K, int(white_bkgd), p(weights), p(rgb), p(depth),
                  _lib.stream_ptr())
        ctx.save_for_backward(rgbsigma, z, rays)
        ctx.white = int(white_bkgd)
        return rgb, depth, weights

    @staticmethod
    def backward(ctx, d_rgb, d_depth, d_weights):
        rgbsigma, z, rays = ctx.saved_tensors
        n, K = z.shape
        p = _lib.ptr
        c = lambda t: None if t is None else t.contiguous().float()
        if d_rgb is None:
            d_rgb = torch.zeros(n, 3, device=z.device)
        d = torch.empty_like(rgbsigma)
        _lib.call("dnr_composite_bwd", p(rgbsigma), p(z), p(rays), p(c(d_rgb)), p(c(d_depth)), p(c(d_weights)), n, K,
                  ctx.white, p(d), _lib.stream_ptr())
        return d, None, None, None


class NeRFRendererDGS(torch.nn.Module):
    """NeRF renderer with depth-guided sampling on hand-written sm_100a kernels (see module docstring).

    :param n_samples: samples per ray evaluated by the field
    :param n_depth_candidates: stratified candidates scored against the source views' depth maps
    :param n_gaussian: samples drawn from the gaussian fitted to the occlusion-aware likelihood
    :param eval_batch_size: kept for signature compatibility (the reference chunks the MLP batch with it,
           nerf_renderer.py:325-333); chunking here is by `rows_per_chunk` rows of the per-view MLP batch
    :param white_bkgd: composite onto white (True) or black
    """

    def __init__(self, n_samples=40, n_depth_candidates=1000, n_gaussian=15, eval_batch_size=100000, white_bkgd=True,
                 rows_per_chunk=1 << 20):
        super().__init__()
        self.n_samples = n_samples
        self.n_depth_candidates = n_depth_candidates
        self.n_gaussian = n_gaussian
        self.eval_batch_size = eval_batch_size
        self.white_bkgd = white_bkgd
        self.rows_per_chunk = rows_per_chunk

    def forward(self, model, rays, want_weights=False, noise=None, return_samples=False):
        """
        :param model: PixelNeRF-like object after encode() (see diner_b200.scene)
        :param rays: [origin(3), direction(3), near, far]  (SB, B, 8), CUDA fp32
        :param want_weights: also return compositing weights (SB, B, K)
        :param noise: optional recorded jitter (parity mode), see sample_rays
        :return: result with .fine.rgb (SB,B,3), .fine.depth (SB,B) [, .fine.weights (SB,B,K)]
        """
        assert len(rays.shape) == 3
        assert self.n_samples >= self.n_gaussian
        if not rays.is_cuda:
            raise RuntimeError("diner_b200.NeRFRendererDGS needs CUDA tensors (no CPU path)")
        if rays.shape[-1] != 8:
            raise ValueError("rays must be (SB, B, 8)")
        SB, NR, _ = rays.shape
        sc = get_scene(model)
        if SB != sc.SB:
            raise ValueError(f"rays batch {SB} does not match the encoded scenes {sc.SB}")
        rays = rays.detach().to(torch.float32).contiguous()
        K = self.n_samples
        if NR == 0:                                  # nothing to render: empty results of the right shapes
            empty = RenderResult(rgb=rays.new_zeros(SB, 0, 3), depth=rays.new_zeros(SB, 0))
            if want_weights:
                empty.weights = rays.new_zeros(SB, 0, K)
            if return_samples:
                empty.z = rays.new_zeros(SB, 0, K)
            return RenderResult(fine=empty)
        if noise is not None and noise.get("z") is not None:
            # parity hook: evaluate field + compositing on externally supplied (sorted) sample depths
            z = noise["z"].detach().to(device=rays.device, dtype=torch.float32).reshape(SB * NR, K).contiguous()
        else:
            z = sample_rays(sc, rays, K, self.n_depth_candidates, self.n_gaussian, noise=noise)

        mlp = model.mlp_fine
        names = [n for n, _ in mlp.named_parameters()]
        params = [q for _, q in mlp.named_parameters()]
        latent = model.encoder.latent
        flat = rays.view(SB * NR, 8)
        n_rays = SB * NR
        per_chunk = max(1, self.rows_per_chunk // (K * sc.NV))
        track = torch.is_grad_enabled() and (latent.requires_grad or any(q.requires_grad for q in params))
        rgbs, depths, weights = [], [], []
        for r0 in range(0, n_rays, per_chunk):
            r1 = min(n_rays, r0 + per_chunk)
            rc, zc = flat[r0:r1], z[r0:r1]
            if track:
                rgbsigma = _FieldFn.apply(sc, mlp, names, rc, zc, NR, r0, latent, *params)
            else:       # inference: nothing is saved, the MLP runs as one persistent kernel
                rgbsigma, _ = F_.field_forward(sc, F_.get_packed(mlp, need_transposed=False), rc, zc, NR, r0, save=False)
            rgb, depth, w = _CompositeFn.apply(rgbsigma, zc, rc, self.white_bkgd)
            rgbs.append(rgb); depths.append(depth); weights.append(w)
        cat = (lambda xs: xs[0] if len(xs) == 1 else torch.cat(xs, 0))
        out = RenderResult(rgb=cat(rgbs).view(SB, NR, 3), depth=cat(depths).view(SB, NR))
        if want_weights:
            out.weights = cat(weights).view(SB, NR, K)
        if return_samples:
            out.z = z.view(SB, NR, K)
        return RenderResult(fine=out)
