"""Generate tests/golden/novel{,_pe}_tiny{,3}.npz (forward) and novel{,_pe}_tiny{,3}_grad.npz (gradients) by running the reference's NOVEL renderer variant (build container only).

    python -m oracle.make_golden_novel

The reference classes `src.models.novel.nerf_novel_renderer.NeRFRendererDGS` and `src.models.novel.novel_pixelnerf.PixelNeRF`
(and, for the `novel_pe_*` cases, `src.models.novel_pe.pe_nerf_novel_renderer.NeRFRendererDGS` and
`src.models.novel_pe.pe_novel_pixelnerf.PixelNeRF`) run UNMODIFIED on a seeded synthetic scene with the three random draws recorded; the only substitution is
`pytorch3d.ops.knn.knn_points` -> oracle.novel_oracle.nearest_vertex (pytorch3d is not in the build image; published K = 1
semantics).  Test infrastructure.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_shim  # noqa: E402
from tools.synthetic import add_novel_inputs, add_novel_pe_inputs, make_scene  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
CASES = {"novel_tiny": ("tiny", 0, 6, 400), "novel_tiny3": ("tiny3", 1, 7, 300),
         # NOVEL_PE (src/models/novel_pe/): same scenes + positional maps, target camera, deformation layer, 518-latent MLP
         "novel_pe_tiny": ("tiny", 0, 8, 400), "novel_pe_tiny3": ("tiny3", 1, 9, 300)}


def build_reference_model(ref, scene):
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
    nerf.gen_poses, nerf.gen_focal, nerf.gen_c = m.gen_poses, m.gen_focal, m.gen_c
    nerf.gen_image_shape = m.gen_image_shape.clone()
    nerf.gen_latent = torch.nn.Parameter(m.gen_latent.clone())
    if getattr(m, "deformation_layer", None) is not None:            # NOVEL_PE
        nerf.deformation_layer.load_state_dict(m.deformation_layer.state_dict())
        nerf.pos_encodings, nerf.tgt_pos_encoding = m.pos_encodings, m.tgt_pos_encoding
        nerf.tgt_poses, nerf.tgt_focal, nerf.tgt_c = m.tgt_poses, m.tgt_focal, m.tgt_c
        nerf.tgt_image_shape = m.tgt_image_shape.clone()
    return nerf


def run_case(ref, name):
    wl_name, seed, noise_seed, n_vert = CASES[name]
    scene = add_novel_inputs(make_scene(wl_name, seed=seed), n_vertices=n_vert, seed=seed)
    if name.startswith("novel_pe"):
        scene = add_novel_pe_inputs(scene, seed=seed)
    wl = scene.workload
    nerf = build_reference_model(ref, scene)
    rays = scene.rays
    renderer = ref.NeRFRendererDGS(n_samples=wl.n_samples, n_depth_candidates=wl.n_candidates,
                                   n_gaussian=wl.n_gaussian, white_bkgd=wl.white_bkgd)
    V, oi, og = scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets
    torch.manual_seed(noise_seed)
    noise = {}
    with torch.no_grad(), ref_shim.record_noise(noise):
        out = renderer.forward(nerf, rays, V, oi, og, want_weights=True)
    with torch.no_grad(), ref_shim.record_noise({}):
        torch.manual_seed(noise_seed)
        z_unf = renderer.sample_depthguided(rays, nerf, n_samples=wl.n_samples, n_candidates=wl.n_candidates,
                                            target_vertices=V, tgt_in_offsets=oi, n_gaussian=wl.n_gaussian)
        z = renderer.fill_up_uniform_samples(z_unf.clone(), rays)
        # the same scene through the reference's DINER sampler (no deformation): the deformation must matter
        dref = ref_shim.load()
        torch.manual_seed(noise_seed)
        z_plain = dref.NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian).sample_depthguided(
            rays, nerf, n_samples=wl.n_samples, n_candidates=wl.n_candidates, n_gaussian=wl.n_gaussian)
    grad_data = reference_gradients(ref, nerf, renderer, rays, V, oi, og, noise_seed)
    assert np.array_equal(grad_data.pop("rgb"), out.fine.rgb.numpy())          # same draws, same frame
    moved = float((z_plain != z_unf).any(-1).float().mean())
    data = dict(workload=wl_name, seed=seed, noise_seed=noise_seed, n_vertices=n_vert,
                u_coarse=noise["u_coarse"].numpy(),
                n_gauss=noise["n_gauss"].numpy() if noise["n_gauss"] is not None else np.zeros((0, wl.n_gaussian), np.float32),
                u_fill=noise["u_fill"].numpy() if noise["u_fill"] is not None else np.zeros((0,), np.float32),
                z_unfilled=z_unf.numpy(), z=z.numpy(), rgb=out.fine.rgb.numpy(), depth=out.fine.depth.numpy(),
                weights=out.fine.weights.numpy())
    print(f"[{name}] rays={tuple(rays.shape)} vertices={n_vert} rays whose samples the deformation moved: {moved:.3f} "
          f"sum_w in [{out.fine.weights.sum(-1).min():.4f},{out.fine.weights.sum(-1).max():.4f}]")
    return data, grad_data


def reference_gradients(ref, nerf, renderer, rays, V, oi, og, noise_seed):
    """Gradients of a fixed scalar loss through the UNMODIFIED reference NOVEL renderer w.r.t. what its training step
    optimises (novel.py:361-363, `self.nerf.parameters()`): the MLP, gen_latent, and encoder.latent (the encoder's
    output: the encoder trunk itself is outside the path)."""
    nerf.encoder.latent.requires_grad_(True)
    torch.manual_seed(noise_seed)
    with ref_shim.record_noise({}):
        out = renderer.forward(nerf, rays, V, oi, og, want_weights=True)
    rgb, depth = out.fine.rgb, out.fine.depth
    g = torch.Generator().manual_seed(11)
    w_rgb = torch.randn(rgb.shape, generator=g)
    w_dep = torch.randn(depth.shape, generator=g)
    loss = (rgb * w_rgb).sum() + (depth * w_dep).sum()
    loss.backward()

    def picked(t, prefix):
        nz = t.flatten().nonzero().flatten()
        pick = nz[torch.linspace(0, len(nz) - 1, 4096).long()]
        return {f"{prefix}_idx": pick.numpy(), f"{prefix}_val": t.flatten()[pick].numpy(),
                f"{prefix}_norm": np.float64(t.double().norm().item()), f"{prefix}_sum": np.float64(t.double().sum().item()),
                f"{prefix}_nnz": np.int64(len(nz))}

    data = dict(loss=np.float64(loss.item()), rgb=rgb.detach().numpy())
    data.update(picked(nerf.encoder.latent.grad, "glat"))
    data.update(picked(nerf.gen_latent.grad, "ggen"))
    for k, v in nerf.mlp_fine.named_parameters():
        if k in ("lin_out.weight", "lin_out.bias", "lin_in.bias", "blocks.4.fc_0.bias", "blocks.1.fc_1.bias", "lin_z.1.bias"):
            data[f"grad.{k}"] = v.grad.numpy()
        data[f"gradnorm.{k}"] = np.float64(v.grad.double().norm().item())
    if getattr(nerf, "deformation_layer", None) is not None:         # NOVEL_PE: the layers with positional columns, in full
        # (the 518-column weights: the six positional columns in full + every 61st entry of the rest -- small fixtures)
        for k, v in (("deformation_layer.weight", nerf.deformation_layer.weight),
                     ("lin_z.0.weight", nerf.mlp_fine.lin_z[0].weight), ("lin_z.1.weight", nerf.mlp_fine.lin_z[1].weight),
                     ("lin_z.2.weight", nerf.mlp_fine.lin_z[2].weight)):
            data[f"gradpe.{k}"] = v.grad[:, 512:].numpy().copy()
            data[f"gradsub.{k}"] = v.grad[:, :512].flatten()[::61].numpy().copy()
            data[f"gradnorm.{k}"] = np.float64(v.grad.double().norm().item())
        for k, v in (("deformation_layer.bias", nerf.deformation_layer.bias), ("lin_in.weight", nerf.mlp_fine.lin_in.weight)):
            data[f"grad.{k}"] = v.grad.numpy()
            data[f"gradnorm.{k}"] = np.float64(v.grad.double().norm().item())
    nerf.encoder.latent.requires_grad_(False)
    print(f"  reference gradients: loss={loss.item():.5f} |d latent|={data['glat_norm']:.4e} ({data['glat_nnz']} entries) "
          f"|d gen_latent|={data['ggen_norm']:.4e} ({data['ggen_nnz']} entries)")
    return data


def main():
    ref, ref_pe = ref_shim.load_novel(), ref_shim.load_novel_pe()
    torch.set_num_threads(os.cpu_count() or 1)
    for name in (sys.argv[1:] or list(CASES)):
        data, grad_data = run_case(ref_pe if name.startswith("novel_pe") else ref, name)
        for suffix, d in (("", data), ("_grad", grad_data)):
            path = os.path.join(GOLDEN, f"{name}{suffix}.npz")
            np.savez_compressed(path, **d)
            print("wrote", path, f"{os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
