"""Metric helpers shared by the round-2 GPU parity tests (tests/test_gpu_parity_r2.py) and tools/parity_probe.py
(which prints the same metrics without asserting, to choose / re-check the tolerances on a B200).

Test infrastructure: everything here compares the CUDA path (through diner_b200's public classes, i.e. through the
C ABI) with oracle/ or with the reference's golden vectors.
"""
from __future__ import annotations

import dataclasses

import numpy as np
import torch


def grad_metrics(got: torch.Tensor, want: torch.Tensor) -> dict:
    """relative L2 error, cosine, and norm ratio of a gradient against its reference."""
    a, b = got.detach().double().flatten().cpu(), want.detach().double().flatten().cpu()
    nb = max(b.norm().item(), 1e-300)
    return dict(rel_l2=(a - b).norm().item() / nb, cos=float(torch.dot(a, b) / max(a.norm().item() * nb, 1e-300)),
                ratio=a.norm().item() / nb)


def golden_loss(rgb, depth):
    """The scalar loss oracle/make_golden.py differentiates: sum(rgb * w_rgb) + sum(depth * w_dep), seed 11."""
    gen = torch.Generator().manual_seed(11)
    w_rgb = torch.randn(rgb.shape, generator=gen).to(rgb.device)
    w_dep = torch.randn(depth.shape, generator=gen).to(rgb.device)
    return (rgb * w_rgb).sum() + (depth * w_dep).sum()


def render_on_golden_z(scene, rays, g, dev, precision="bf16", train=False):
    """Field + compositing on the reference's own sample depths; returns (out, model on dev).  With train=True the
    latent and the MLP parameters require grad and the golden loss has been back-propagated."""
    from diner_b200 import NeRFRendererDGS
    wl = scene.workload
    m = scene.model.to_device(dev)
    for p in m.mlp_fine.parameters():
        p.grad = None
    m.encoder.latent = m.encoder.latent.detach().requires_grad_(train)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd, precision=precision)
    noise = dict(z=torch.from_numpy(g["z"]))
    if train:
        out = r(m, rays.to(dev), want_weights=True, noise=noise)
        golden_loss(out.fine.rgb, out.fine.depth).backward()
    else:
        with torch.no_grad():
            out = r(m, rays.to(dev), want_weights=True, noise=noise)
    return out, m


def golden_gradient_metrics(scene, rays, g, dev, precision="bf16") -> dict:
    """CUDA backward vs. the UNMODIFIED reference's gradients stored in tests/golden/*.npz (oracle/make_golden.py:85-101):
    six full parameter gradients, the norm of every parameter gradient, 4 096 entries + norm + sum of d_latent."""
    out, m = render_on_golden_z(scene, rays, g, dev, precision, train=True)
    res = {}
    try:
        for k, v in m.mlp_fine.named_parameters():
            if f"grad.{k}" in g:
                res[f"full.{k}"] = grad_metrics(v.grad, torch.from_numpy(g[f"grad.{k}"]))
            res[f"norm.{k}"] = v.grad.double().norm().item() / max(float(g[f"gradnorm.{k}"]), 1e-300)
        glat = m.encoder.latent.grad
        pick = glat.flatten()[torch.from_numpy(g["glat_idx"]).to(dev)]
        res["glat.picked"] = grad_metrics(pick, torch.from_numpy(g["glat_val"]))
        res["glat.norm"] = glat.double().norm().item() / float(g["glat_norm"])
        res["glat.sum"] = (glat.double().sum().item() - float(g["glat_sum"])) / max(abs(float(g["glat_sum"])), float(g["glat_norm"]))
        res["loss"] = abs(float(golden_loss(out.fine.rgb.detach(), out.fine.depth.detach())) - float(g["loss"])) / abs(float(g["loss"]))
    finally:
        m.encoder.latent = m.encoder.latent.detach()
        for p in m.mlp_fine.parameters():
            p.grad = None
        m.to_device("cpu")
    return res


def error_metrics(got: torch.Tensor, want: torch.Tensor) -> dict:
    d = (got.detach().float().cpu() - want.detach().float().cpu()).abs().flatten()
    scale = max(float(want.abs().max()), 1e-30)
    return dict(mean=d.mean().item(), q99=d.quantile(0.99).item() if d.numel() < 16_000_000 else float("nan"),
                q999=d.quantile(0.999).item() if d.numel() < 16_000_000 else float("nan"), max=d.max().item(),
                rel_max=d.max().item() / scale)


def strided_subset(scene, n: int = 256, offset: int = 37):
    """~n rays spread over the target frame: a regular grid over the whole image (object and background),
    or a plain stride when the ray batch is not a full frame."""
    rays, wl = scene.rays, scene.workload
    if rays.shape[1] == wl.tgt_H * wl.tgt_W and wl.tgt_H * wl.tgt_W > 4 * n:
        side = int(round(n ** 0.5))
        ys = torch.linspace(0.02 * wl.tgt_H, 0.98 * wl.tgt_H - 1, side).long()
        xs = torch.linspace(0.02 * wl.tgt_W, 0.98 * wl.tgt_W - 1, side).long()
        idx = (ys[:, None] * wl.tgt_W + xs[None, :]).reshape(-1)
        return rays[:, idx].contiguous()
    stride = max(1, rays.shape[1] // n)
    return rays[:, offset::stride][:, :n].contiguous()


def unpacked_noise(n_rays: int, wl, seed: int = 9) -> dict:
    gen = torch.Generator().manual_seed(seed)
    return dict(u_coarse=torch.rand(n_rays, wl.n_candidates, generator=gen),
                n_gauss_full=torch.randn(n_rays, wl.n_gaussian, generator=gen),
                u_fill_full=torch.rand(n_rays, wl.n_samples, generator=gen))


def oracle_sampler(model, rays, wl, noise) -> dict:
    """oracle sampler on whatever device `model` / `rays` live on (torch CPU or torch CUDA ops)."""
    from oracle import dgs_oracle as O
    with torch.no_grad():
        z0, _, dbg = O.sample_depthguided(model, rays, wl.n_samples, wl.n_candidates, wl.n_gaussian, noise=dict(noise))
        z, _ = O.fill_up_uniform(z0.clone(), rays, None, noise["u_fill_full"])
    K, M = wl.n_samples, wl.n_samples - wl.n_gaussian
    order = dbg.order.view(-1, K)[:, :M]
    p = dbg.p.view(-1, wl.n_candidates)
    idx = torch.where(torch.gather(p, 1, order) > 0, order, torch.full_like(order, -1))
    return dict(z=z.view(-1, K), p=p, idx=idx, rm=dbg.ray_mask.view(-1).int(), nm=(z0.view(-1, K) == 0).sum(-1).int())


def zoom_sources(scene, factor: float):
    """Re-render the synthetic source views with `factor` x the focal length, so that the object (and with it the
    non-zero depth std) runs over the image border: exercises the exponentially padded std ring and the border /
    zero padding of the depth / normal taps with LIVE values."""
    from tools import synthetic as S
    wl, m = scene.workload, scene.model
    g = torch.Generator().manual_seed(4242)
    Ks = scene.src_intrinsics.clone()
    Ks[..., 0, 0] *= factor
    Ks[..., 1, 1] *= factor
    depths, stds = [], []
    for sb in range(wl.SB):
        d_s, s_s = [], []
        for v in range(wl.NV):
            d = S.render_ellipsoid_depth(m.poses[sb, v].double(), Ks[sb, v].double(), wl.H, wl.W, wl.scale).float()
            conf = torch.rand(wl.H, wl.W, generator=g) * 0.8 + 0.2
            std = S.conf_to_std(conf, wl.std_law) * (d != 0)
            d_s.append(d); s_s.append(std)
        depths.append(torch.stack(d_s)); stds.append(torch.stack(s_s))
    depths = torch.stack(depths).unsqueeze(2)
    m.encoder.depths = depths
    m.encoder.depths_std = torch.stack(stds).unsqueeze(2)
    m.encoder.normals = S.depth_to_normals(depths.flatten(0, 1), Ks.flatten(0, 1)).view(wl.SB, wl.NV, 3, wl.H, wl.W)
    m.focal = torch.stack((Ks[..., 0, 0], Ks[..., 1, 1]), -1)
    scene.src_intrinsics = Ks
    return scene


def config_shape_parity(workload: str, dev, n_rays: int = 256) -> dict:
    """A strided subset of a BASELINE.json config's frame: sampler vs. the oracle on torch CPU ops (fraction of rays
    whose 40 samples agree) and on torch CUDA ops (likelihoods / indices / masks, bit for bit), rgb / depth on the
    oracle's samples vs. the fp32 CPU oracle."""
    from diner_b200 import NeRFRendererDGS, sample_rays
    from diner_b200.scene import get_scene
    from oracle import dgs_oracle as O
    from tools.synthetic import make_scene
    scene = make_scene(workload, seed=0)
    wl = scene.workload
    K = wl.n_samples
    sub = strided_subset(scene, n_rays)
    n = sub.shape[0] * sub.shape[1]
    noise = unpacked_noise(n, wl)
    oc = oracle_sampler(scene.model, sub, wl, noise)                          # torch CPU
    with torch.no_grad():
        _, rgb_ref, depth_ref, _ = O.composite(scene.model, sub, oc["z"].view(sub.shape[0], -1, K), wl.white_bkgd)
    m = scene.model.to_device(dev)
    res = {}
    try:
        nd = {k: v.to(dev) for k, v in noise.items()}
        og = oracle_sampler(m, sub.to(dev), wl, nd)                           # torch CUDA
        z, d = sample_rays(get_scene(m), sub.to(dev), K, wl.n_candidates, wl.n_gaussian, noise=nd, want_debug=True)
        res["p_equal"] = bool(torch.equal(d["p"], og["p"]))
        res["p_nonzero"] = int((og["p"] != 0).sum())
        res["idx_equal"] = bool(torch.equal(d["sel_idx"].long(), og["idx"]))
        res["mask_equal"] = bool(torch.equal(d["ray_mask"], og["rm"]) and torch.equal(d["n_missing"], og["nm"]))
        res["z_vs_cuda_oracle"] = error_metrics(z, og["z"])
        res["frac_fg"] = float(og["rm"].float().mean())
        res["rays_equal_cpu_oracle"] = float(torch.isclose(z.cpu(), oc["z"], rtol=1e-5, atol=1e-6).all(-1).float().mean())
        # how far torch itself is from bit-stability across its two backends on these rays (DESIGN.md "Parity")
        res["cuda_oracle_equal_cpu_oracle"] = float(torch.isclose(og["z"].cpu(), oc["z"], rtol=1e-5, atol=1e-6).all(-1).float().mean())
        res["sorted"] = bool((z[:, 1:] >= z[:, :-1]).all())
        r = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        with torch.no_grad():
            out = r(m, sub.to(dev), noise=dict(z=oc["z"]))
        res["rgb"] = error_metrics(out.fine.rgb, rgb_ref)
        res["depth"] = error_metrics(out.fine.depth, depth_ref)
        rs = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd, precision="split")
        with torch.no_grad():
            out = rs(m, sub.to(dev), noise=dict(z=oc["z"]))
        res["rgb_split"] = error_metrics(out.fine.rgb, rgb_ref)
        res["depth_split"] = error_metrics(out.fine.depth, depth_ref)
    finally:
        m.to_device("cpu")
    return res


def train_shape_gradient_metrics(dev, precision="bf16") -> dict:
    """BASELINE config 4 (SB = 2 scenes x 4 096 rays, NV = 2): forward + backward of the whole batch through the CUDA
    path vs. the fp32 oracle evaluated with torch CUDA ops + autograd on the SAME sample depths (so that the full
    batch fits the test budget; the oracle's CPU and CUDA evaluations agree to fp32 round-off)."""
    from diner_b200 import NeRFRendererDGS, sample_rays
    from diner_b200.scene import get_scene
    from oracle import dgs_oracle as O
    from tools.synthetic import make_scene
    scene = make_scene("facescape_train", seed=0)
    wl = scene.workload
    K = wl.n_samples
    m = scene.model.to_device(dev)
    rays = scene.rays.to(dev)
    gen = torch.Generator().manual_seed(5)
    gt = torch.rand(rays.shape[0], rays.shape[1], 3, generator=gen).to(dev)
    z = sample_rays(get_scene(m), rays, K, wl.n_candidates, wl.n_gaussian, seed=21).view(wl.SB, -1, K)
    res = {}
    try:
        # oracle (fp32, torch CUDA ops)
        lat = m.encoder.latent.detach()
        m.encoder.latent = lat.clone().requires_grad_(True)
        _, rgb_ref, _, _ = O.composite(m, rays, z, wl.white_bkgd)
        torch.nn.functional.mse_loss(rgb_ref, gt).backward()
        want = {k: v.grad.clone() for k, v in m.mlp_fine.named_parameters()}
        want_lat = m.encoder.latent.grad.clone()
        for v in m.mlp_fine.parameters():
            v.grad = None
        # CUDA path
        m.encoder.latent = lat.clone().requires_grad_(True)
        r = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd, precision=precision)
        out = r(m, rays, noise=dict(z=z))
        torch.nn.functional.mse_loss(out.fine.rgb, gt).backward()
        res["rgb"] = error_metrics(out.fine.rgb, rgb_ref)
        for k, v in m.mlp_fine.named_parameters():
            res[k] = grad_metrics(v.grad, want[k])
        res["latent"] = grad_metrics(m.encoder.latent.grad, want_lat)
        per_scene = [m.encoder.latent.grad[s].abs().sum().item() for s in range(wl.SB)]
        res["latent_per_scene_nonzero"] = all(x > 0 for x in per_scene)
    finally:
        m.encoder.latent = m.encoder.latent.detach()
        for v in m.mlp_fine.parameters():
            v.grad = None
        m.to_device("cpu")
    return res
