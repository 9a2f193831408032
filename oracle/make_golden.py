"""Generate tests/golden/*.npz by running the UNMODIFIED reference renderer (build container only).

    python -m oracle.make_golden            # writes tests/golden/{tiny,tiny3,plumbing64}.npz

For each case the synthetic scene is rebuilt deterministically from (workload, seed) by
tools.synthetic.make_scene, its tensors are injected into the reference's own PixelNeRF
(so the reference's index_* / forward / ResnetFC code runs on them), the reference
NeRFRendererDGS.forward is run under torch's CPU backend with the three random draws recorded,
and inputs-by-seed + noise + outputs (+ gradients) are stored.  Test infrastructure.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tools.synthetic import make_scene  # noqa: E402
from oracle import ref_shim  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = {
    # name: (workload, scene seed, ray subset or None, noise seed)
    "tiny": ("tiny", 0, None, 3),
    "tiny3": ("tiny3", 1, None, 4),
    "plumbing64": ("plumbing64", 0, slice(1536, 1536 + 192), 5),
}


def build_reference_model(ref, scene):
    """The reference's PixelNeRF with the synthetic scene's tensors in place of encode()'s."""
    A = ref.DotMap
    nerf = ref.PixelNeRF(
        poscode_conf=A(kwargs=dict(num_freqs=6, freq_factor=6.28, include_input=True)),
        encoder_conf=A(module="src.models.image_encoder.SpatialEncoder",
                       kwargs=dict(image_padding=64, padding_pe=4, pretrained=False)),
        mlp_fine_conf=A(module="src.models.resnetfc.ResnetFC",
                        kwargs=dict(n_blocks=5, d_hidden=512, combine_layer=3, combine_type="average")))
    m = scene.model
    nerf.mlp_fine.load_state_dict(m.mlp_fine.state_dict())
    enc = nerf.encoder
    enc.latent = m.encoder.latent.clone()
    enc.depths, enc.depths_std, enc.normals = m.encoder.depths, m.encoder.depths_std, m.encoder.normals
    enc.nviews, enc.nobjects = m.encoder.nviews, m.encoder.nobjects
    nerf.poses, nerf.focal, nerf.c = m.poses, m.focal, m.c
    nerf.image_shape = m.image_shape.clone()
    return nerf


def run_case(ref, name):
    wl_name, seed, sub, noise_seed = CASES[name]
    scene = make_scene(wl_name, seed=seed)
    wl = scene.workload
    nerf = build_reference_model(ref, scene)
    rays = scene.rays if sub is None else scene.rays[:, sub].contiguous()
    renderer = ref.NeRFRendererDGS(n_samples=wl.n_samples, n_depth_candidates=wl.n_candidates,
                                   n_gaussian=wl.n_gaussian, white_bkgd=wl.white_bkgd)
    nerf.encoder.latent.requires_grad_(True)
    torch.manual_seed(noise_seed)
    noise = {}
    with ref_shim.record_noise(noise):
        out = renderer.forward(model=nerf, rays=rays, want_weights=True)
    rgb, depth, weights = out.fine.rgb, out.fine.depth, out.fine.weights

    # stage outputs with the same noise (the reference has no hooks; re-run its own methods)
    with torch.no_grad(), ref_shim.record_noise({}):
        torch.manual_seed(noise_seed)
        z_unf = renderer.sample_depthguided(rays, nerf, n_samples=wl.n_samples, n_candidates=wl.n_candidates,
                                            n_gaussian=wl.n_gaussian)
        z = renderer.fill_up_uniform_samples(z_unf.clone(), rays)

    # gradients of a fixed scalar loss
    g = torch.Generator().manual_seed(11)
    w_rgb = torch.randn(rgb.shape, generator=g)
    w_dep = torch.randn(depth.shape, generator=g)
    loss = (rgb * w_rgb).sum() + (depth * w_dep).sum()
    loss.backward()
    glat = nerf.encoder.latent.grad
    nz = glat.flatten().nonzero().flatten()
    pick = nz[torch.linspace(0, len(nz) - 1, 4096).long()] if len(nz) > 0 else nz
    grads = {f"grad.{k}": v.grad.numpy() for k, v in nerf.mlp_fine.named_parameters()
             if k in ("lin_out.weight", "lin_out.bias", "lin_in.bias", "blocks.4.fc_0.bias", "blocks.1.fc_1.bias",
                      "lin_z.1.bias")}
    for k, v in nerf.mlp_fine.named_parameters():
        grads[f"gradnorm.{k}"] = np.float64(v.grad.double().norm().item())
    sig = out_sig = None
    data = dict(
        workload=wl_name, seed=seed, noise_seed=noise_seed,
        ray_start=-1 if sub is None else sub.start, ray_stop=-1 if sub is None else sub.stop,
        u_coarse=noise["u_coarse"].numpy(),
        n_gauss=noise["n_gauss"].numpy() if noise["n_gauss"] is not None else np.zeros((0, wl.n_gaussian), np.float32),
        u_fill=noise["u_fill"].numpy() if noise["u_fill"] is not None else np.zeros((0,), np.float32),
        z_unfilled=z_unf.numpy(), z=z.numpy(),
        rgb=rgb.detach().numpy(), depth=depth.detach().numpy(), weights=weights.detach().numpy(),
        loss=np.float64(loss.item()),
        glat_idx=pick.numpy(), glat_val=glat.flatten()[pick].numpy(),
        glat_norm=np.float64(glat.double().norm().item()), glat_sum=np.float64(glat.double().sum().item()),
        **grads)
    frac_fg = float((weights.sum(-1) > 0.5).float().mean())
    print(f"[{name}] rays={tuple(rays.shape)} n_masked={data['n_gauss'].shape[0]} n_fill={data['u_fill'].shape[0]} "
          f"sum_w in [{weights.sum(-1).min():.4f},{weights.sum(-1).max():.4f}] frac(sum_w>.5)={frac_fg:.3f} "
          f"|glat|={data['glat_norm']:.4e}")
    return data


def main():
    ref = ref_shim.load()
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    torch.set_num_threads(os.cpu_count() or 1)
    for name in (sys.argv[1:] or list(CASES)):
        data = run_case(ref, name)
        path = os.path.join(GOLDEN_DIR, f"{name}.npz")
        np.savez_compressed(path, **data)
        print("wrote", path, f"{os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
