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

import os

import torch

from . import _lib, field as F_
from .scene import DEPTH_DIFF_MAX, STD_PAD, SceneBuffers, get_scene, lin_table


# Evaluation renders go through ONE C-ABI call (dnr_render_fwd, csrc/render.cu: the same launches in the same order, issued
# by the library) unless a test or the bench's per-call instrumentation asks for the call-by-call sequence below.
USE_NATIVE_FRAME = os.environ.get("DNR_NATIVE_FRAME", "1") != "0"
_workspaces: dict = {}


def _native_descs(sc: SceneBuffers, pk, lin: torch.Tensor):
    """ctypes images of dnr_scene_desc / dnr_mlp_desc, cached on the scene buffers / packed weights they point into."""
    sd = getattr(sc, "_desc", None)
    if sd is None:
        p = _lib.ptr
        sd = _lib.SceneDesc(poses=p(sc.poses), focal=p(sc.focal), cxy=p(sc.cxy), depth=p(sc.depth),
                            depth_std_padded=p(sc.std_padded), normal=p(sc.normal), latent_nhwc_bf16=p(sc.latent_nhwc),
                            freqs=p(sc.freqs), phases=p(sc.phases), lin_tab=None, SB=sc.SB, NV=sc.NV, H=sc.H, W=sc.W,
                            Hl=sc.Hl, Wl=sc.Wl, L=sc.L, NF=sc.num_freqs, pad=STD_PAD, img_w=sc.img_w, img_h=sc.img_h,
                            std_sx=sc.std_sx, std_sy=sc.std_sy, lat_sx=sc.lat_sx, lat_sy=sc.lat_sy,
                            depth_diff_max=DEPTH_DIFF_MAX)
        sc._desc = sd
    sd.lin_tab = lin.data_ptr()
    md = getattr(pk, "_desc", None)
    if md is None:
        import ctypes as C
        cast = lambda a: C.cast(a, C.c_void_p)
        md = _lib.MlpDesc(w128=cast(pk.w128_ptrs), w256=cast(pk.w_ptrs), pair_biases=cast(pk.pair_b_ptrs),
                          w_out=pk.w_out.data_ptr(), b_out=pk.b_out.data_ptr())
        pk._desc = md
    return sd, md


class RenderResult(dict):
    """Attribute-access dict (stands in for dotmap.DotMap, nerf_renderer.py:421-430)."""

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name) from None       # hasattr / getattr-with-default / copy / pickle rely on this

    __setattr__ = dict.__setitem__


def sample_rays(sc: SceneBuffers, rays: torch.Tensor, K: int, C: int, G: int, noise: dict | None = None,
                seed: int | None = None, want_debug: bool = False, cand_offset=None):
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
                  p(sc.normal), p(lin), p(u_coarse), p(n_gauss), p(u_fill), p(goff), p(foff), p(cand_offset),
                  seed & (2 ** 64 - 1),
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
    """rgbsigma = field(rays, z) over all ray chunks of one forward call; differentiable w.r.t. encoder.latent and
    the 30 MLP parameter tensors.  The chunks only tile the work: one flat parameter-gradient buffer and one feature
    gradient map are zero-filled, accumulated into by every chunk and unpacked ONCE per backward call."""

    @staticmethod
    def forward(ctx, sc, mlp, names, rays, z, NR, chunks, precision, latent, *params):
        need_grad = any(ctx.needs_input_grad[8:])
        pk = F_.get_packed(mlp, need_transposed=need_grad, split=precision == "split")
        out = torch.empty(z.shape[0] * z.shape[1], 4, device=z.device, dtype=torch.float32)
        K = z.shape[1]
        saved = []
        for r0, r1 in chunks:
            _, sv = F_.field_forward(sc, pk, rays[r0:r1], z[r0:r1], NR, r0, save=need_grad, out=out[r0 * K:r1 * K],
                                     precision=precision)
            saved.append(sv)
        ctx.sc, ctx.mlp, ctx.names, ctx.pk, ctx.saved = sc, mlp, names, pk, saved
        ctx.rays, ctx.z, ctx.NR, ctx.chunks = rays, z, NR, chunks
        ctx.latent_shape = latent.shape
        return out

    @staticmethod
    def backward(ctx, d_out):
        sc, mlp = ctx.sc, ctx.mlp
        dev = d_out.device
        with torch.cuda.device(dev):
            params = dict(mlp.named_parameters())
            # one zero fill for all parameter gradients (views into a flat buffer, 16-byte aligned starts)
            offs, total = {}, 0
            for n in ctx.names:
                offs[n] = total
                total += (params[n].numel() + 3) // 4 * 4
            flat = torch.zeros(total, device=dev, dtype=torch.float32)
            grads = {n: flat[offs[n]:offs[n] + params[n].numel()].view(params[n].shape) for n in ctx.names}
            d_lat_nhwc = None
            if ctx.needs_input_grad[8]:
                d_lat_nhwc = torch.zeros(sc.SB, sc.NV, sc.Hl, sc.Wl, sc.L, device=dev, dtype=torch.float32)
            d_out = d_out.contiguous()
            K = ctx.z.shape[1]
            for (r0, r1), sv in zip(ctx.chunks, ctx.saved):
                F_.field_backward(sc, ctx.pk, mlp, sv, ctx.rays[r0:r1], ctx.z[r0:r1], ctx.NR, r0,
                                  d_out[r0 * K:r1 * K], grads, d_lat_nhwc)
            d_latent = None
            if d_lat_nhwc is not None:
                d_latent = torch.empty(ctx.latent_shape, device=dev, dtype=torch.float32)
                _lib.call("dnr_unpack_dlatent_nchw", _lib.ptr(d_lat_nhwc), sc.SB * sc.NV, sc.L, sc.Hl, sc.Wl,
                          _lib.ptr(d_latent), _lib.stream_ptr())
            ctx.saved = None
            pg = tuple(grads[n].to(params[n].dtype) if ctx.needs_input_grad[9 + i] else None
                       for i, n in enumerate(ctx.names))
        return (None, None, None, None, None, None, None, None, d_latent, *pg)


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


def balanced_chunks(n_rays: int, per_chunk: int):
    """[(r0, r1)] covering n_rays in ceil(n_rays / per_chunk) chunks whose sizes differ by at most one ray (no
    degenerate tail chunk: 65 536 rays at 13 107 per chunk are 6 chunks of 10 922 / 10 923, not 5 + one ray)."""
    n_chunks = max(1, -(-n_rays // max(1, per_chunk)))
    base, extra = divmod(n_rays, n_chunks)
    out, r0 = [], 0
    for i in range(n_chunks):
        r1 = r0 + base + (1 if i < extra else 0)
        out.append((r0, r1))
        r0 = r1
    return out


class NeRFRendererDGS(torch.nn.Module):
    """NeRF renderer with depth-guided sampling on hand-written sm_100a kernels (see module docstring).

    :param n_samples: samples per ray evaluated by the field
    :param n_depth_candidates: stratified candidates scored against the source views' depth maps
    :param n_gaussian: samples drawn from the gaussian fitted to the occlusion-aware likelihood
    :param eval_batch_size: kept for signature compatibility (the reference chunks the MLP batch with it,
           nerf_renderer.py:325-333); chunking here is by `rows_per_chunk` rows of the per-view MLP batch
    :param white_bkgd: composite onto white (True) or black
    :param rows_per_chunk: per-view MLP rows evaluated per kernel sequence; None = ROWS_INFERENCE without autograd
           (1 152 B of packed operands per row), ROWS_TRAINING with (every layer's operands are kept for backward)
    :param precision: "bf16" (default: bf16 x bf16 -> fp32 tensor-core MLP) | "split" (verification mode: every MLP
           operand is carried as a bf16 hi + lo pair and every product as hi*hi + hi*lo + lo*hi on the same kernels,
           layouts and epilogues -- ~2^-16 relative per product instead of 2^-8; about 3.5x slower)
    """
    ROWS_INFERENCE = 1 << 22
    ROWS_TRAINING = 1 << 20

    def __init__(self, n_samples=40, n_depth_candidates=1000, n_gaussian=15, eval_batch_size=100000, white_bkgd=True,
                 rows_per_chunk=None, precision="bf16"):
        super().__init__()
        if precision not in ("bf16", "split"):
            raise ValueError("precision must be 'bf16' or 'split'")
        self.n_samples = n_samples
        self.n_depth_candidates = n_depth_candidates
        self.n_gaussian = n_gaussian
        self.eval_batch_size = eval_batch_size
        self.white_bkgd = white_bkgd
        self.rows_per_chunk = rows_per_chunk
        self.precision = precision

    def forward(self, model, rays, want_weights=False, noise=None, return_samples=False):
        """
        :param model: PixelNeRF-like object after encode() (see diner_b200.scene)
        :param rays: [origin(3), direction(3), near, far]  (SB, B, 8), CUDA fp32
        :param want_weights: also return compositing weights (SB, B, K)
        :param noise: optional recorded jitter (parity mode), see sample_rays; dict(seed=int) fixes the in-kernel jitter
        :return: result with .fine.rgb (SB,B,3), .fine.depth (SB,B) [, .fine.weights (SB,B,K)]
        """
        assert len(rays.shape) == 3
        assert self.n_samples >= self.n_gaussian
        if not rays.is_cuda:
            raise RuntimeError("diner_b200.NeRFRendererDGS needs CUDA tensors (no CPU path)")
        if rays.shape[-1] != 8:
            raise ValueError("rays must be (SB, B, 8)")
        lat_dev = model.encoder.latent.device
        if lat_dev != rays.device:
            raise RuntimeError(f"diner_b200: rays live on {rays.device} but the encoded scene on {lat_dev}")
        with torch.cuda.device(rays.device):     # every launch goes to the stream / context of the rays' device
            return self._forward(model, rays, want_weights, noise, return_samples)

    def _forward(self, model, rays, want_weights, noise, return_samples):
        SB, NR, _ = rays.shape
        if getattr(model, "deformation_layer", None) is not None:
            raise NotImplementedError("diner_b200.NeRFRendererDGS: this is a NOVEL_PE field -- use diner_b200.novel_pe")
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
        mlp = model.mlp_fine
        names = [n for n, _ in mlp.named_parameters()]
        params = [q for _, q in mlp.named_parameters()]
        if params[0].device != rays.device:
            raise RuntimeError(f"diner_b200: rays live on {rays.device} but model.mlp_fine on {params[0].device}")
        latent = model.encoder.latent
        flat = rays.view(SB * NR, 8)
        n_rays = SB * NR
        track = torch.is_grad_enabled() and (latent.requires_grad or any(q.requires_grad for q in params))
        rows = self.rows_per_chunk or (self.ROWS_TRAINING if track or self.precision == "split" else self.ROWS_INFERENCE)
        # one library call for the whole render when nothing asks for the call-by-call sequence (recorded reference draws,
        # another MLP kernel, the verification mode, per-call timing)
        native = (USE_NATIVE_FRAME and not track and self.precision == "bf16" and sc.NV <= 4 and _lib.profile is None
                  and F_.FORWARD_KERNEL == "quad" and F_.MERGE_POOLED and F_.PAIR_SHARE is None and F_.PAIR_PROF is None
                  and (noise is None or set(noise) <= {"z", "seed"}))
        z = None
        if noise is not None and noise.get("z") is not None:
            # parity hook: evaluate field + compositing on externally supplied (sorted) sample depths
            z = noise["z"].detach().to(device=rays.device, dtype=torch.float32).reshape(SB * NR, K).contiguous()
        elif not native:
            draws = {k: v for k, v in (noise or {}).items() if k != "seed"} or None
            z = sample_rays(sc, rays, K, self.n_depth_candidates, self.n_gaussian, noise=draws,
                            seed=(noise or {}).get("seed"))
        if native:
            return self._render_native(sc, mlp, flat, SB, NR, z, rows, (noise or {}).get("seed"), want_weights,
                                       return_samples)
        chunks = balanced_chunks(n_rays, max(1, rows // (K * sc.NV)))
        if track:
            rgbsigma = _FieldFn.apply(sc, mlp, names, flat, z, NR, chunks, self.precision, latent, *params)
        else:           # inference: nothing is saved, the MLP runs as one persistent kernel per chunk
            pk = F_.get_packed(mlp, need_transposed=False, split=self.precision == "split")
            rgbsigma = torch.empty(n_rays * K, 4, device=rays.device, dtype=torch.float32)
            for r0, r1 in chunks:
                F_.field_forward(sc, pk, flat[r0:r1], z[r0:r1], NR, r0, save=False, out=rgbsigma[r0 * K:r1 * K],
                                 precision=self.precision)
        # one compositing launch per call (nerf_renderer.py:344-365), whatever the chunking of the field
        rgb, depth, w = _CompositeFn.apply(rgbsigma, z, flat, self.white_bkgd)
        out = RenderResult(rgb=rgb.view(SB, NR, 3), depth=depth.view(SB, NR))
        if want_weights:
            out.weights = w.view(SB, NR, K)
        if return_samples:
            out.z = z.view(SB, NR, K)
        return RenderResult(fine=out)

    def _render_native(self, sc, mlp, flat, SB, NR, z_in, rows, seed, want_weights, return_samples):
        """The evaluation render as one C-ABI call (dnr_render_fwd): same launches, same order, same bits."""
        import ctypes as C
        dev = flat.device
        K, Cn, G = self.n_samples, self.n_depth_candidates, self.n_gaussian
        n_rays = SB * NR
        pk = F_.get_packed(mlp, need_transposed=False)
        sd, md = _native_descs(sc, pk, lin_table(Cn, dev))
        need = _lib.load().dnr_render_workspace_bytes(sc.NV, n_rays, K, rows)
        key = (str(dev), torch.cuda.current_stream(dev).cuda_stream)          # one per stream: calls on two streams may overlap
        ws = _workspaces.get(key)
        if ws is None or ws.numel() < need:
            _workspaces.pop(key, None)
            ws = None                                                         # (release before growing)
            ws = _workspaces[key] = torch.empty(need, device=dev, dtype=torch.uint8)
        f32 = dict(device=dev, dtype=torch.float32)
        rgb, depth = torch.empty(n_rays, 3, **f32), torch.empty(n_rays, **f32)
        w = torch.empty(n_rays, K, **f32) if want_weights else None
        z_out = (z_in if z_in is not None else torch.empty(n_rays, K, **f32)) if return_samples else None
        if seed is None and z_in is None:
            seed = int(torch.empty((), dtype=torch.int64).random_().item())
        info = (C.c_int * 8)()
        p = _lib.ptr
        _lib.call("dnr_render_fwd", C.byref(sd), C.byref(md), p(flat), NR, K, Cn, G, (seed or 0) & (2 ** 64 - 1), p(z_in),
                  int(self.white_bkgd), rows, p(ws), ws.numel(), p(rgb), p(depth), p(w), p(z_out), info, _lib.stream_ptr())
        _lib.launch_count += info[0] - 1                                      # (call() counted one)
        F_.LAST_SPLIT.update(tiles_main=info[2], tiles_side=info[3], tiles=info[4], kernel="quad")
        out = RenderResult(rgb=rgb.view(SB, NR, 3), depth=depth.view(SB, NR))
        if want_weights:
            out.weights = w.view(SB, NR, K)
        if return_samples:
            out.z = z_out.view(SB, NR, K)
        return RenderResult(fine=out)
