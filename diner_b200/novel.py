"""NOVEL renderer variant -- evaluation and training step.

Drop-in for the reference's `src.models.novel.nerf_novel_renderer.NeRFRendererDGS` (selected by the NOVEL YAMLs'
`renderer.module`), same constructor and the same

    forward(model, rays, target_vertices, tgt_in_offsets, tgt_gen_offsets, want_weights=False)

What the variant changes against the DINER path (reference file:line), and where it runs here:

  * every sampler CANDIDATE is moved by the offset of its nearest target-mesh vertex before it is projected into the
    source views (nerf_novel_renderer.py:40-50,104-107)  -> `dnr_novel_deform` (candidates) + `dnr_sampler_fwd(cand_offset)`
  * every SAMPLE point is deformed twice, with tgt_in_offsets and tgt_gen_offsets (:441-442)   -> `dnr_novel_deform`
  * the field reads a second, learned feature map `model.gen_latent` through one generic camera and adds it to the
    encoder feature (novel_pixelnerf.py:46,73-80,108-141,163-196)                             -> `dnr_encode_novel_fwd`
  * MLP and compositing are the DINER kernels.

NOVEL_PE (src/models/novel_pe/): the same renderer (pe_nerf_novel_renderer.py differs in one line, :444) over another
field -- recognised by `model.deformation_layer`; see `diner_b200.novel_pe`, `field.PackedPE`, `dnr_encode_novel_pe_fwd`.

`model` is the reference's novel `PixelNeRF` after `encode()` and `encode_gen()` (duck-typed): everything
`diner_b200.scene` reads, plus `gen_poses (SB,1,4,4)`, `gen_focal / gen_c (SB,1,2)`, `gen_image_shape (2,)`,
`gen_latent (512,Hg,Wg)`.

Training: the NOVEL training step renders with offsets taken from the batch and optimises `self.nerf.parameters()`
(src/models/novel/novel.py:280-285,361-363) -- the MLP, the encoder behind `encoder.latent`, and `gen_latent`.  With
autograd enabled the call is differentiable w.r.t. exactly those: the DINER backward (csrc/gemm.cu, wgrad.cu,
mlp_misc.cu) plus `dnr_encode_novel_bwd`, which scatters the feature gradient into both maps.  The deformed points carry
no gradient (nearest-vertex selection has none, and the reference's own offsets are data): offsets that require grad
raise NotImplementedError rather than silently returning detached results.

The nearest-vertex search is exact (equal to the brute force `diner_b200.deform.nearest_vertex` bit for bit) but pruned:
the mesh is sorted along a Morton curve into groups of 32 vertices with a bounding sphere each (`build_vertex_groups`).
"""
from __future__ import annotations

import torch

from . import _lib, field as F_
from .renderer import RenderResult, _CompositeFn, balanced_chunks, sample_rays
from .scene import SceneBuffers, get_scene, lin_table

GROUP = 32
SUB_SPHERES = True       # third pruning level of the nearest-vertex search (A/B switch; the result is the same bits)


def _morton3(q: torch.Tensor) -> torch.Tensor:
    """q (.., 3) int64 in [0, 1024) -> 30-bit Morton code."""
    def spread(x):
        x = (x | (x << 16)) & 0x030000FF
        x = (x | (x << 8)) & 0x0300F00F
        x = (x | (x << 4)) & 0x030C30C3
        x = (x | (x << 2)) & 0x09249249
        return x
    return spread(q[..., 0]) | (spread(q[..., 1]) << 1) | (spread(q[..., 2]) << 2)


def build_vertex_groups(vertices: torch.Tensor):
    """vertices (SB,V,3) -> (verts_sorted (SB, G*32, 4) fp32 [x, y, z, original index as int32 bits],
    spheres (SB, G, 4) fp32 [centre, radius], supers (SB, ceil(G/8), 2, 4) fp32 [box lo | box hi of 8 consecutive groups],
    subs (SB, G, 4, 4) fp32 [centre, radius of each run of 8 vertices]).  Padding slots sit at 1e18 with index INT_MAX
    (never the minimum)."""
    SB, V, _ = vertices.shape
    dev = vertices.device
    v = vertices.detach().to(torch.float32)
    lo, hi = v.amin(1, keepdim=True), v.amax(1, keepdim=True)
    q = ((v - lo) / (hi - lo).clamp(min=1e-20) * 1023.0).long().clamp(0, 1023)
    order = torch.argsort(_morton3(q), dim=1, stable=True)                       # SB,V
    G = (V + GROUP - 1) // GROUP
    sv = torch.gather(v, 1, order.unsqueeze(-1).expand(-1, -1, 3))
    pad = G * GROUP - V
    idx_bits = order.to(torch.int32)
    if pad:
        sv = torch.cat([sv, torch.full((SB, pad, 3), 1e18, device=dev)], 1)
        idx_bits = torch.cat([idx_bits, torch.full((SB, pad), 2 ** 31 - 1, device=dev, dtype=torch.int32)], 1)
    grp = sv.view(SB, G, GROUP, 3)
    valid = (torch.arange(G * GROUP, device=dev) < V).view(1, G, GROUP, 1)
    big = torch.finfo(torch.float32).max
    gmin = torch.where(valid, grp, torch.full_like(grp, big)).amin(2)
    gmax = torch.where(valid, grp, torch.full_like(grp, -big)).amax(2)
    centre = 0.5 * (gmin + gmax)
    rad = torch.where(valid[..., 0], (grp - centre.unsqueeze(2)).norm(dim=-1), torch.zeros((), device=dev)).amax(2)
    spheres = torch.cat([centre, (rad * (1 + 1e-5) + 1e-9).unsqueeze(-1)], -1).contiguous()
    verts_sorted = torch.cat([sv, idx_bits.view(torch.float32).unsqueeze(-1)], -1).contiguous()
    # super-groups: the axis-aligned box of 8 consecutive groups (a little larger than tight, so that rounding cannot prune)
    S = (G + 7) // 8
    padg = S * 8 - G
    gmin_p = torch.cat([gmin, torch.full((SB, padg, 3), big, device=dev)], 1) if padg else gmin
    gmax_p = torch.cat([gmax, torch.full((SB, padg, 3), -big, device=dev)], 1) if padg else gmax
    smin = gmin_p.view(SB, S, 8, 3).amin(2)
    smax = gmax_p.view(SB, S, 8, 3).amax(2)
    ext = (smax - smin).abs().amax(-1, keepdim=True)
    smin, smax = smin - (1e-6 * ext + 1e-9), smax + (1e-6 * ext + 1e-9)
    zero = torch.zeros(SB, S, 1, device=dev)
    supers = torch.stack([torch.cat([smin, zero], -1), torch.cat([smax, zero], -1)], 2).contiguous()      # SB,S,2,4
    # third level: the four runs of 8 consecutive (Morton-sorted) vertices of a group, a sphere each
    if not SUB_SPHERES:
        return verts_sorted, spheres, supers, None
    run, rvalid = sv.view(SB, G, 4, 8, 3), valid.view(1, G, 4, 8, 1)
    rmin = torch.where(rvalid, run, torch.full_like(run, big)).amin(3)
    rmax = torch.where(rvalid, run, torch.full_like(run, -big)).amax(3)
    rc = 0.5 * (rmin + rmax)                                                           # (empty runs: centre 0, radius 0)
    rr = torch.where(rvalid[..., 0], (run - rc.unsqueeze(3)).norm(dim=-1), torch.zeros((), device=dev)).amax(3)
    subs = torch.cat([rc, (rr * (1 + 1e-5) + 1e-9).unsqueeze(-1)], -1).contiguous()   # SB,G,4,4
    return verts_sorted, spheres, supers, subs


def _deform(rays, z, lin, u_coarse, seed, groups, off_a, off_b, SB, NR, per_ray, V, candidates, want_idx=False):
    verts_sorted, spheres, supers, subs = groups
    dev = rays.device
    n = SB * NR * per_ray
    out_a = torch.empty(n, 3, device=dev, dtype=torch.float32)
    out_b = torch.empty(n, 3, device=dev, dtype=torch.float32) if off_b is not None else None
    idx = torch.empty(n, device=dev, dtype=torch.int32) if want_idx else None
    p = _lib.ptr
    _lib.call("dnr_novel_deform", p(rays), p(z), p(lin), p(u_coarse), seed & (2 ** 64 - 1), p(verts_sorted), p(spheres),
              p(supers), p(subs), p(off_a), p(off_b), SB, NR, per_ray, V, spheres.shape[1], 1 if candidates else 0, p(out_a), p(out_b), p(idx),
              _lib.stream_ptr())
    return out_a, out_b, idx


_gen_cache: dict = {}


def _gen_latent_nhwc(model) -> torch.Tensor:
    """model.gen_latent (512,Hg,Wg) -> (Hg,Wg,512) bf16, cached on (data_ptr, _version) (see scene.invalidate)."""
    gl = model.gen_latent
    key = (gl.data_ptr(), gl._version, tuple(gl.shape))
    hit = _gen_cache.get(id(model))
    if hit is not None and hit[0] == key:
        return hit[1]
    C, Hg, Wg = gl.shape
    if C != 512:
        raise NotImplementedError("diner_b200.novel: gen_latent must have 512 channels")
    src = gl.detach().to(torch.float32).contiguous()
    out = torch.empty(Hg, Wg, C, device=gl.device, dtype=torch.bfloat16)
    _lib.call("dnr_pack_latent_nhwc", _lib.ptr(src), 1, C, Hg, Wg, _lib.ptr(out), _lib.stream_ptr())
    _gen_cache.clear()
    _gen_cache[id(model)] = (key, out)
    return out


def _is_pe(model) -> bool:
    """NOVEL_PE field (src/models/novel_pe/pe_novel_pixelnerf.py): a `deformation_layer` in front of a 518-latent MLP."""
    return getattr(model, "deformation_layer", None) is not None


def _named_params(model) -> dict:
    """Everything the renderer's output depends on among `nerf.parameters()`, apart from the two feature maps."""
    named = dict(model.mlp_fine.named_parameters())
    if _is_pe(model):
        named.update({f"deformation_layer.{n}": q for n, q in model.deformation_layer.named_parameters()})
    return named


def _packed(model, need_transposed: bool):
    if _is_pe(model):
        return F_.get_packed_pe(model, need_transposed)
    return F_.get_packed(model.mlp_fine, need_transposed=need_transposed)


class _NovelFieldFn(torch.autograd.Function):
    """rgbsigma = field(deformed sample points) over all ray chunks of one forward call (NOVEL variant); differentiable
    w.r.t. encoder.latent, gen_latent and the MLP parameter tensors.  Same structure as renderer._FieldFn."""

    @staticmethod
    def forward(ctx, sc, model, names, rays, z, NR, chunks, gen, pts_in, pts_gen, latent, gen_latent, *params):
        pk = _packed(model, True)
        K = z.shape[1]
        out = torch.empty(z.shape[0] * K, 4, device=z.device, dtype=torch.float32)
        saved = []
        for r0, r1 in chunks:
            novel = dict(gen, pts_in=pts_in[r0 * K:r1 * K], pts_gen=pts_gen[r0 * K:r1 * K])
            _, sv = F_.field_forward(sc, pk, rays[r0:r1], z[r0:r1], NR, r0, save=True, out=out[r0 * K:r1 * K], novel=novel)
            saved.append(sv)
        ctx.sc, ctx.model, ctx.names, ctx.pk, ctx.saved = sc, model, names, pk, saved
        ctx.rays, ctx.z, ctx.NR, ctx.chunks, ctx.gen, ctx.pts = rays, z, NR, chunks, gen, (pts_in, pts_gen)
        ctx.latent_shape, ctx.gen_shape = latent.shape, gen_latent.shape
        return out

    @staticmethod
    def backward(ctx, d_out):
        sc, mlp, gen = ctx.sc, ctx.model.mlp_fine, ctx.gen
        dev = d_out.device
        p = _lib.ptr
        with torch.cuda.device(dev):
            params = _named_params(ctx.model)
            offs, total = {}, 0
            for n in ctx.names:
                offs[n] = total
                total += (params[n].numel() + 3) // 4 * 4
            flat = torch.zeros(total, device=dev, dtype=torch.float32)
            grads = {n: flat[offs[n]:offs[n] + params[n].numel()].view(params[n].shape) for n in ctx.names}
            d_lat_nhwc = d_gen_nhwc = None
            if ctx.needs_input_grad[10]:
                d_lat_nhwc = torch.zeros(sc.SB, sc.NV, sc.Hl, sc.Wl, sc.L, device=dev, dtype=torch.float32)
            if ctx.needs_input_grad[11]:
                d_gen_nhwc = torch.zeros(gen["Hg"], gen["Wg"], sc.L, device=dev, dtype=torch.float32)
            d_out = d_out.contiguous()
            K = ctx.z.shape[1]
            pts_in, pts_gen = ctx.pts
            for (r0, r1), sv in zip(ctx.chunks, ctx.saved):
                novel = dict(gen, pts_in=pts_in[r0 * K:r1 * K], pts_gen=pts_gen[r0 * K:r1 * K])
                F_.field_backward(sc, ctx.pk, mlp, sv, ctx.rays[r0:r1], ctx.z[r0:r1], ctx.NR, r0, d_out[r0 * K:r1 * K],
                                  grads, d_lat_nhwc, novel=novel, d_gen_nhwc=d_gen_nhwc)
            d_latent = d_gen = None
            if d_lat_nhwc is not None:
                d_latent = torch.empty(ctx.latent_shape, device=dev, dtype=torch.float32)
                _lib.call("dnr_unpack_dlatent_nchw", p(d_lat_nhwc), sc.SB * sc.NV, sc.L, sc.Hl, sc.Wl, p(d_latent),
                          _lib.stream_ptr())
            if d_gen_nhwc is not None:
                d_gen = torch.empty(ctx.gen_shape, device=dev, dtype=torch.float32)
                _lib.call("dnr_unpack_dlatent_nchw", p(d_gen_nhwc), 1, sc.L, gen["Hg"], gen["Wg"], p(d_gen),
                          _lib.stream_ptr())
            ctx.saved = None
            pg = tuple(grads[n].to(params[n].dtype) if ctx.needs_input_grad[12 + i] else None
                       for i, n in enumerate(ctx.names))
        return (None,) * 10 + (d_latent, d_gen, *pg)


class NeRFRendererDGS(torch.nn.Module):
    """NOVEL variant of the depth-guided renderer (reference src/models/novel/nerf_novel_renderer.py:12-556); see the
    module docstring.  Same constructor as the DINER renderer."""

    ROWS_INFERENCE = 1 << 22
    ROWS_TRAINING = 1 << 20

    def __init__(self, n_samples=40, n_depth_candidates=1000, n_gaussian=15, eval_batch_size=100000, white_bkgd=True,
                 rows_per_chunk=None):
        super().__init__()
        self.n_samples = n_samples
        self.n_depth_candidates = n_depth_candidates
        self.n_gaussian = n_gaussian
        self.eval_batch_size = eval_batch_size
        self.white_bkgd = white_bkgd
        self.rows_per_chunk = rows_per_chunk

    def deform_points(self, points, target_vertices, offsets):
        """nerf_novel_renderer.py:40-50 (brute force, differentiable w.r.t. points and offsets)."""
        from .deform import deform_points
        return deform_points(points, target_vertices, offsets)

    def forward(self, model, rays, target_vertices, tgt_in_offsets, tgt_gen_offsets, want_weights=False, noise=None,
                return_samples=False):
        assert len(rays.shape) == 3
        assert self.n_samples >= self.n_gaussian
        if not rays.is_cuda:
            raise RuntimeError("diner_b200.novel.NeRFRendererDGS needs CUDA tensors (no CPU path)")
        if torch.is_grad_enabled() and any(t.requires_grad for t in (tgt_in_offsets, tgt_gen_offsets, target_vertices)):
            raise NotImplementedError("diner_b200.novel: no gradient flows to the offsets / vertices (the reference's "
                                      "training step takes them from the batch, novel.py:280-285) -- detach them")
        with torch.cuda.device(rays.device):
            return self._forward(model, rays, target_vertices, tgt_in_offsets, tgt_gen_offsets, want_weights, noise,
                                 return_samples)

    @staticmethod
    def _pe_inputs(model, SB, NV, f):
        """NOVEL_PE: the positional maps and the target camera (pe_novel_pixelnerf.py:75,90-99,133-198)."""
        src, tgt = model.pos_encodings, model.tgt_pos_encoding
        if src.dim() != 5 or tuple(src.shape[:3]) != (SB, NV, 3):
            raise ValueError(f"model.pos_encodings must be (SB,NV,3,H,W), got {tuple(src.shape)}")
        if tgt.dim() != 5 or tuple(tgt.shape[:3]) != (SB, 1, 3):
            raise ValueError(f"model.tgt_pos_encoding must be (SB,1,3,H,W), got {tuple(tgt.shape)}")
        if tuple(model.tgt_poses.shape) != (SB, 1, 4, 4):
            raise ValueError(f"model.tgt_poses must be (SB,1,4,4), got {tuple(model.tgt_poses.shape)}")
        fp = float(model.encoder.feature_padding)

        def rescale(t):                                           # :150-152, :185-187 (W then H)
            size = torch.tensor([t.shape[-1], t.shape[-2]])
            sc_ = ((size - fp * 2) / size).float()
            return float(sc_[0]), float(sc_[1])

        (sx, sy), (tx, ty) = rescale(src), rescale(tgt)
        tshape = model.tgt_image_shape.detach().float().cpu()
        return dict(src=f(src), Hp=src.shape[-2], Wp=src.shape[-1], sx=sx, sy=sy, tgt=f(tgt), Ht=tgt.shape[-2],
                    Wt=tgt.shape[-1], tgt_sx=tx, tgt_sy=ty, tgt_poses=f(model.tgt_poses), tgt_focal=f(model.tgt_focal),
                    tgt_cxy=f(model.tgt_c), tgt_img_w=float(tshape[0]), tgt_img_h=float(tshape[1]))

    def _forward(self, model, rays, verts, off_in, off_gen, want_weights, noise, return_samples):
        SB, NR, _ = rays.shape
        dev = rays.device
        sc: SceneBuffers = get_scene(model)
        if SB != sc.SB:
            raise ValueError(f"rays batch {SB} does not match the encoded scenes {sc.SB}")
        for name, t in (("target_vertices", verts), ("tgt_in_offsets", off_in), ("tgt_gen_offsets", off_gen)):
            if t.dim() != 3 or t.shape[0] != SB or t.shape[-1] != 3 or t.shape != verts.shape or t.device != dev:
                raise ValueError(f"{name} must be (SB, V, 3) on {dev}, got {tuple(t.shape)} on {t.device}")
        V = verts.shape[1]
        K, C, G = self.n_samples, self.n_depth_candidates, self.n_gaussian
        rays = rays.detach().to(torch.float32).contiguous()
        f = lambda t: t.detach().to(torch.float32).contiguous()
        off_in, off_gen = f(off_in), f(off_gen)
        groups = build_vertex_groups(verts)
        if NR == 0:
            empty = RenderResult(rgb=rays.new_zeros(SB, 0, 3), depth=rays.new_zeros(SB, 0))
            if want_weights:
                empty.weights = rays.new_zeros(SB, 0, K)
            return RenderResult(fine=empty)
        noise = dict(noise or {})
        if noise.get("z") is not None:
            z = noise["z"].detach().to(device=dev, dtype=torch.float32).reshape(SB * NR, K).contiguous()
        else:
            seed = noise.get("seed")
            if seed is None:
                seed = int(torch.empty((), dtype=torch.int64).random_().item())
            u_coarse = noise.get("u_coarse")
            if u_coarse is not None:
                u_coarse = u_coarse.to(device=dev, dtype=torch.float32).contiguous()
            cand_off, _, _ = _deform(rays, None, lin_table(C, dev), u_coarse, seed, groups, off_in, None, SB, NR, C, V, True)
            noise.pop("seed", None)
            z = sample_rays(sc, rays, K, C, G, noise=noise or None, seed=seed, cand_offset=cand_off)
            del cand_off
        flat = rays.view(SB * NR, 8)
        n_rays = SB * NR
        pts_in, pts_gen, _ = _deform(rays, z, None, None, 0, groups, off_in, off_gen, SB, NR, K, V, False)
        gen_nhwc = _gen_latent_nhwc(model)
        Hg, Wg = gen_nhwc.shape[:2]
        gsize = torch.tensor([Wg, Hg])
        gscale = ((gsize - float(model.encoder.feature_padding) * 2) / gsize).float()      # novel_pixelnerf.py:127-128
        gshape = model.gen_image_shape.detach().float().cpu()
        gen = dict(poses=f(model.gen_poses), focal=f(model.gen_focal), cxy=f(model.gen_c), latent=gen_nhwc, Hg=Hg, Wg=Wg,
                   img_w=float(gshape[0]), img_h=float(gshape[1]), sx=float(gscale[0]), sy=float(gscale[1]))
        if gen["poses"].shape != (SB, 1, 4, 4):
            raise ValueError(f"model.gen_poses must be (SB,1,4,4), got {tuple(gen['poses'].shape)}")
        if _is_pe(model):
            gen["pe"] = self._pe_inputs(model, SB, sc.NV, f)
        named = _named_params(model)
        names, params = list(named), list(named.values())
        latent, gen_latent = model.encoder.latent, model.gen_latent
        track = torch.is_grad_enabled() and (latent.requires_grad or gen_latent.requires_grad or
                                             any(q.requires_grad for q in params))
        rows = self.rows_per_chunk or (self.ROWS_TRAINING if track else self.ROWS_INFERENCE)
        chunks = balanced_chunks(n_rays, max(1, rows // (K * sc.NV)))
        if track:
            rgbsigma = _NovelFieldFn.apply(sc, model, names, flat, z, NR, chunks, gen, pts_in, pts_gen, latent, gen_latent,
                                           *params)
        else:
            pk = _packed(model, False)
            rgbsigma = torch.empty(n_rays * K, 4, device=dev, dtype=torch.float32)
            for r0, r1 in chunks:
                novel = dict(gen, pts_in=pts_in[r0 * K:r1 * K], pts_gen=pts_gen[r0 * K:r1 * K])
                F_.field_forward(sc, pk, flat[r0:r1], z[r0:r1], NR, r0, save=False, out=rgbsigma[r0 * K:r1 * K],
                                 novel=novel)
        rgb, depth, w = _CompositeFn.apply(rgbsigma, z, flat, self.white_bkgd)
        out = RenderResult(rgb=rgb.view(SB, NR, 3), depth=depth.view(SB, NR))
        if want_weights:
            out.weights = w.view(SB, NR, K)
        if return_samples:
            out.z = z.view(SB, NR, K)
        return RenderResult(fine=out)
