"""GPU tests of the NOVEL renderer variant, run with `-m gpu` on a B200, through the C ABI.

Checkers: oracle/novel_oracle.py (bit-pinned on the CPU to the reference's NOVEL renderer run unmodified except for
pytorch3d's knn_points, tests/test_oracle.py) and the golden vectors of that run (tests/golden/novel_*.npz).
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()
    return torch.device("cuda", 0)


ALL = ["novel_tiny", "novel_tiny3", "novel_pe_tiny", "novel_pe_tiny3"]      # novel_pe_*: the NOVEL_PE field


def _load(name):
    from tools.synthetic import add_novel_inputs, add_novel_pe_inputs, make_scene
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", f"{name}.npz"), allow_pickle=False))
    scene = add_novel_inputs(make_scene(str(g["workload"]), seed=int(g["seed"])), n_vertices=int(g["n_vertices"]),
                             seed=int(g["seed"]))
    if name.startswith("novel_pe"):
        scene = add_novel_pe_inputs(scene, seed=int(g["seed"]))
    noise = dict(u_coarse=torch.from_numpy(g["u_coarse"]), n_gauss=torch.from_numpy(g["n_gauss"]),
                 u_fill=torch.from_numpy(g["u_fill"]))
    return g, scene, noise


@pytest.mark.parametrize("V,kind", [(5023, "surface"), (777, "cloud"), (33, "surface"), (1, "cloud"), (640, "ties")])
def test_pruned_nearest_vertex_equals_brute_force(V, kind):
    """dnr_novel_deform (Morton-sorted groups of 32 + bounding spheres) against dnr_deform_points (brute force) on the
    same points: identical vertex indices -- including duplicated vertices (ties go to the lowest original index) -- and
    identical deformed points, for points near the mesh and far from it."""
    dev = _cuda()
    from diner_b200.deform import nearest_vertex
    from diner_b200.novel import _deform, build_vertex_groups
    gen = torch.Generator().manual_seed(V)
    SB, NR, K = 2, 300, 40
    if kind == "surface":
        u = torch.randn(SB, V, 3, generator=gen)
        verts = u / u.norm(dim=-1, keepdim=True) * torch.tensor([0.11, 0.15, 0.12])
    else:
        verts = torch.randn(SB, V, 3, generator=gen) * 0.1
    if kind == "ties":
        verts[:, V // 2:] = verts[:, : V - V // 2]                       # every vertex twice
    off_a, off_b = torch.randn(SB, V, 3, generator=gen) * 0.01, torch.randn(SB, V, 3, generator=gen) * 0.01
    rays = torch.zeros(SB, NR, 8)
    rays[..., :3] = torch.randn(SB, NR, 3, generator=gen) * 0.05 + torch.tensor([0.0, 0.0, -1.7])
    d = torch.randn(SB, NR, 3, generator=gen) * 0.08 + torch.tensor([0.0, 0.0, 1.0])
    rays[..., 3:6] = d / d.norm(dim=-1, keepdim=True)
    rays[..., 6], rays[..., 7] = 1.0, 2.5
    z = (1.0 + 1.5 * torch.rand(SB, NR, K, generator=gen)).sort(-1).values
    rays, z, verts, off_a, off_b = (t.to(dev).contiguous() for t in (rays, z, verts, off_a, off_b))
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
    want_a, want_idx, _ = nearest_vertex(pts, verts, off_a)
    want_b, _, _ = nearest_vertex(pts, verts, off_b)
    groups = build_vertex_groups(verts)
    got_a, got_b, idx = _deform(rays.view(-1, 8), z.view(-1, K), None, None, 0, groups, off_a, off_b, SB, NR, K, V, False,
                                want_idx=True)
    torch.cuda.synchronize()
    assert torch.equal(idx.view(SB, -1), want_idx)
    assert torch.equal(got_a.view(SB, -1, 3), want_a) and torch.equal(got_b.view(SB, -1, 3), want_b)


@pytest.mark.parametrize("name", ["novel_tiny", "novel_tiny3"])
def test_novel_sampler_bit_exact_vs_cuda_oracle(name):
    """Candidates deformed by dnr_novel_deform + dnr_sampler_fwd(cand_offset) against the NOVEL oracle's sampler executed
    with torch CUDA ops: per-candidate likelihoods, shortlist indices in rank order, masks -- bit for bit."""
    dev = _cuda()
    import parity_metrics as PM
    from diner_b200 import sample_rays
    from diner_b200.novel import _deform, build_vertex_groups
    from diner_b200.scene import get_scene, lin_table
    from oracle import dgs_oracle as O, novel_oracle as N
    from tools.synthetic import novel_to_device
    g, scene, _ = _load(name)
    wl = scene.workload
    rays = scene.rays
    SB, NR = rays.shape[:2]
    noise = PM.unpacked_noise(SB * NR, wl, seed=31)
    novel_to_device(scene, dev)
    m = scene.model
    try:
        nd = {k: v.to(dev) for k, v in noise.items()}
        rd = rays.to(dev)
        V = scene.target_vertices.shape[1]
        deform = lambda x: N.deform_points(x, scene.target_vertices, scene.tgt_in_offsets)
        with torch.no_grad():
            z0, _, dbg = O.sample_depthguided(m, rd, wl.n_samples, wl.n_candidates, wl.n_gaussian, noise=dict(nd), deform=deform)
            z_ref, _ = O.fill_up_uniform(z0.clone(), rd, None, nd["u_fill_full"])
        K, M = wl.n_samples, wl.n_samples - wl.n_gaussian
        order = dbg.order.view(-1, K)[:, :M]
        p_ref = dbg.p.view(-1, wl.n_candidates)
        idx_ref = torch.where(torch.gather(p_ref, 1, order) > 0, order, torch.full_like(order, -1))
        groups = build_vertex_groups(scene.target_vertices)
        cand_off, _, _ = _deform(rd.view(-1, 8), None, lin_table(wl.n_candidates, dev), nd["u_coarse"], 0, groups,
                                 scene.tgt_in_offsets.contiguous(), None, SB, NR, wl.n_candidates, V, True)
        z, d = sample_rays(get_scene(m), rd, K, wl.n_candidates, wl.n_gaussian, noise=nd, want_debug=True, cand_offset=cand_off)
        assert int((p_ref != 0).sum()) > 1000
        assert torch.equal(d["p"], p_ref)
        assert torch.equal(d["sel_idx"].long(), idx_ref)
        assert torch.equal(d["ray_mask"], dbg.ray_mask.view(-1).int())
        torch.testing.assert_close(z, z_ref.view(-1, K), rtol=1e-5, atol=1e-6)
        # and the deformation matters: the DINER sampler on the same rays picks other samples
        z_plain = sample_rays(get_scene(m), rd, K, wl.n_candidates, wl.n_gaussian, noise=nd)
        assert float((z_plain != z).any(-1).float().mean()) > 0.3
    finally:
        novel_to_device(scene, "cpu")


@pytest.mark.parametrize("name", ALL)
def test_novel_render_against_reference_golden(name):
    """diner_b200.novel.NeRFRendererDGS.forward (the reference's signature) against the golden outputs of the reference's
    NOVEL renderer: sampler with the reference's recorded draws (most rays agree; the rest is torch's CPU-vs-CUDA
    rounding, as for the DINER sampler), field with the second latent + compositing on the reference's own samples."""
    dev = _cuda()
    import parity_metrics as PM
    from diner_b200.novel import NeRFRendererDGS
    from tools.synthetic import novel_to_device
    from oracle import dgs_oracle as O, novel_oracle as N
    g, scene, noise = _load(name)
    wl = scene.workload
    # the reference packs its gaussian / fill draws by its OWN ray masks: locate every ray's slice with the CPU oracle's
    # masks (as tests/test_gpu_stages.py does for the DINER sampler), so that one ray on which torch-CPU and CUDA
    # disagree does not shift the draws of all rays behind it
    with torch.no_grad():
        z0, _, dbg = O.sample_depthguided(scene.model, scene.rays, wl.n_samples, wl.n_candidates, wl.n_gaussian,
                                          noise=dict(noise), deform=lambda x: N.deform_points(
                                              x, scene.target_vertices, scene.tgt_in_offsets))
    rm = dbg.ray_mask.view(-1).int()
    nm = (z0.view(-1, wl.n_samples) == 0).sum(-1).int()
    noise = dict(noise, gauss_off=torch.cumsum(rm, 0) - rm, fill_off=torch.cumsum(nm, 0) - nm)
    novel_to_device(scene, dev)
    try:
        r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        args = (scene.model, scene.rays.to(dev), scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets)
        with torch.no_grad():
            out = r(*args, want_weights=True, noise={k: v.to(dev) for k, v in noise.items()}, return_samples=True)
            out_z = r(*args, want_weights=True, noise=dict(z=torch.from_numpy(g["z"])))
        K = wl.n_samples
        same = torch.isclose(out.fine.z.cpu().view(-1, K), torch.from_numpy(g["z"]).view(-1, K), rtol=1e-5, atol=1e-6).all(-1)
        assert same.float().mean().item() >= 0.85
        for key in ("rgb", "depth", "weights"):
            e = PM.error_metrics(out_z.fine[key], torch.from_numpy(g[key]))
            # bf16 MLP tolerance: the mean is the meaningful figure; the second feature map roughly doubles
            # the magnitude of the latent operand, so the tail is a little wider than on the DINER goldens
            assert e["mean"] < 4e-3 and e["q99"] < 6e-2, (key, e)
        assert out.fine.rgb.shape == g["rgb"].shape and torch.isfinite(out.fine.rgb).all()
        # no gradient flows to the offsets (data in the reference's training step): asking for one fails loudly
        scene.tgt_in_offsets.requires_grad_(True)
        with pytest.raises(NotImplementedError):
            r(*args)
        scene.tgt_in_offsets.requires_grad_(False)
        # in-kernel jitter (no recorded draws): deterministic for a fixed seed, sorted samples
        with torch.no_grad():
            a = r(*args, noise=dict(seed=5), return_samples=True)
            b = r(*args, noise=dict(seed=5), return_samples=True)
        assert torch.equal(a.fine.rgb, b.fine.rgb) and bool((a.fine.z[..., 1:] >= a.fine.z[..., :-1]).all())
    finally:
        novel_to_device(scene, "cpu")


def test_novel_encode_rows_vs_oracle():
    """dnr_encode_novel_fwd: the gathered feature row equals gen_latent[gen_uv] + latent[uv] and the geometric code is
    built from the `in`-deformed point (oracle/novel_oracle.field_inputs), to bf16 operand accuracy."""
    dev = _cuda()
    from diner_b200 import _lib, field as F_
    from diner_b200.novel import NeRFRendererDGS, _deform, _gen_latent_nhwc, build_vertex_groups
    from diner_b200.scene import get_scene
    from oracle import novel_oracle as N
    from tools.synthetic import novel_to_device
    g, scene, _ = _load("novel_tiny3")
    wl = scene.workload
    K = wl.n_samples
    rays = scene.rays[:, :50].contiguous()
    SB, NR = rays.shape[:2]
    z = torch.from_numpy(g["z"])[:, :50].contiguous()
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
    dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
    with torch.no_grad():
        zx = N.field_inputs(scene.model, N.deform_points(pts, scene.target_vertices, scene.tgt_in_offsets),
                            N.deform_points(pts, scene.target_vertices, scene.tgt_gen_offsets), dirs)      # SB,NV,B,567
    novel_to_device(scene, dev)
    m = scene.model
    try:
        sc = get_scene(m)
        NV = sc.NV
        rd, zd = rays.to(dev).view(-1, 8), z.to(dev).view(-1, K)
        V = scene.target_vertices.shape[1]
        groups = build_vertex_groups(scene.target_vertices)
        pin, pgen, _ = _deform(rd, zd, None, None, 0, groups, scene.tgt_in_offsets.contiguous(),
                               scene.tgt_gen_offsets.contiguous(), SB, NR, K, V, False)
        gen_nhwc = _gen_latent_nhwc(m)
        Hg, Wg = gen_nhwc.shape[:2]
        gsize = torch.tensor([Wg, Hg])
        gscale = ((gsize - float(m.encoder.feature_padding) * 2) / gsize).float()
        f = lambda t: t.detach().float().contiguous()
        p, st = _lib.ptr, _lib.stream_ptr()
        n = SB * NR
        mv = _lib.load().dnr_view_row_tiles(n * K, NV)
        lat = torch.empty(mv * 128 * 512, device=dev, dtype=torch.bfloat16)
        code = torch.empty(mv * 128 * 64, device=dev, dtype=torch.bfloat16)
        gshape = m.gen_image_shape.cpu()
        _lib.call("dnr_encode_novel_fwd", p(rd), p(zd), p(pin), p(pgen), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth),
                  p(sc.latent_nhwc), p(f(m.gen_poses)), p(f(m.gen_focal)), p(f(m.gen_c)), p(gen_nhwc), Hg, Wg,
                  float(gshape[0]), float(gshape[1]), float(gscale[0]), float(gscale[1]), p(sc.freqs), p(sc.phases),
                  *F_._encode_args(sc, NR, K, 0, n), p(lat), p(code), st)
        lat_rows = torch.empty(mv * 128, 512, device=dev)
        code_rows = torch.empty(mv * 128, 64, device=dev)
        _lib.call("dnr_unpack_rows", p(lat), mv * 128, 512, p(lat_rows), st)
        _lib.call("dnr_unpack_rows", p(code), mv * 128, 64, p(code_rows), st)
        torch.cuda.synchronize()
        spw = 32 // NV
        s_ = torch.arange(n * K)
        for v in range(NV):
            row = ((s_ // spw) * 32 + (s_ % spw) * NV + v).to(dev)
            want = torch.cat([zx[sb, v] for sb in range(SB)], 0)            # rows follow the flattened (scene, ray, sample) order
            got_l, got_c = lat_rows[row].cpu(), code_rows[row].cpu()
            assert (got_l - want[:, :512]).abs().max() <= 2.0 ** -6 * want[:, :512].abs().max() + 1e-6
            assert torch.allclose(got_c[:, :55], want[:, 512:], rtol=2.0 ** -8, atol=2e-3)
    finally:
        novel_to_device(scene, "cpu")


@pytest.mark.parametrize("n_rays,name", [(50, "novel_tiny3"), (97, "novel_tiny")])
def test_novel_encode_bwd_vs_oracle_autograd(n_rays, name):
    """dnr_encode_novel_bwd: a random upstream gradient of the feature rows, scattered into encoder.latent (footprint of
    the `in`-deformed point per view) and into gen_latent (footprint of the `gen`-deformed point, summed over views and
    scenes) -- against torch autograd through oracle/novel_oracle.field_inputs (fp32, CPU).  Also each output alone
    (the other pointer null)."""
    dev = _cuda()
    from diner_b200 import _lib, field as F_
    from diner_b200.novel import _deform, build_vertex_groups
    from diner_b200.scene import get_scene
    from oracle import novel_oracle as N
    from tools.synthetic import novel_to_device
    g, scene, _ = _load(name)
    wl = scene.workload
    K = wl.n_samples
    rays = scene.rays[:, :n_rays].contiguous()
    SB, NR = rays.shape[:2]
    z = torch.from_numpy(g["z"])[:, :n_rays].contiguous()
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
    dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
    m = scene.model
    lat_leaf = m.encoder.latent.detach().clone().requires_grad_(True)
    gen_leaf = m.gen_latent.detach().clone().requires_grad_(True)
    m.encoder.latent, m.gen_latent = lat_leaf, gen_leaf
    zx = N.field_inputs(m, N.deform_points(pts, scene.target_vertices, scene.tgt_in_offsets),
                        N.deform_points(pts, scene.target_vertices, scene.tgt_gen_offsets), dirs)      # SB,NV,B,567
    NV = zx.shape[1]
    gen = torch.Generator().manual_seed(n_rays)
    G = torch.randn(SB, NV, NR * K, 512, generator=gen)
    (zx[..., :512] * G).sum().backward()
    want_dlat, want_dgen = lat_leaf.grad.clone(), gen_leaf.grad.clone()
    m.encoder.latent, m.gen_latent = lat_leaf.detach(), gen_leaf.detach()
    novel_to_device(scene, dev)
    try:
        sc = get_scene(m)
        rd, zd = rays.to(dev).view(-1, 8), z.to(dev).view(-1, K)
        V = scene.target_vertices.shape[1]
        pin, pgen, _ = _deform(rd, zd, None, None, 0, build_vertex_groups(scene.target_vertices),
                               scene.tgt_in_offsets.contiguous(), scene.tgt_gen_offsets.contiguous(), SB, NR, K, V, False)
        Hg, Wg = m.gen_latent.shape[1:]
        gsize = torch.tensor([Wg, Hg])
        gscale = ((gsize - float(m.encoder.feature_padding) * 2) / gsize).float()
        gshape = m.gen_image_shape.cpu()
        f = lambda t: t.detach().float().contiguous()
        p, st = _lib.ptr, _lib.stream_ptr()
        n = SB * NR
        mv = _lib.load().dnr_view_row_tiles(n * K, NV)
        spw = 32 // NV
        s_ = torch.arange(n * K)
        g_rows = torch.randn(mv * 128, 512, generator=gen)               # padding rows hold junk on purpose
        for v in range(NV):
            row = (s_ // spw) * 32 + (s_ % spw) * NV + v
            g_rows[row] = torch.cat([G[sb, v] for sb in range(SB)], 0)
        g_xr = torch.empty(mv * 128 * 512, device=dev)
        _lib.call("dnr_rows_to_xr", p(g_rows.to(dev).contiguous()), mv * 128, 512, mv * 128, p(g_xr), st)
        gp, gf, gc = f(m.gen_poses), f(m.gen_focal), f(m.gen_c)

        def run(want_lat, want_gen):
            d_lat = torch.zeros(sc.SB, sc.NV, sc.Hl, sc.Wl, 512, device=dev) if want_lat else None
            d_gen = torch.zeros(Hg, Wg, 512, device=dev) if want_gen else None
            _lib.call("dnr_encode_novel_bwd", p(rd), p(zd), p(pin), p(pgen), p(sc.poses), p(sc.focal), p(sc.cxy),
                      p(sc.depth), p(gp), p(gf), p(gc), Hg, Wg, float(gshape[0]), float(gshape[1]), float(gscale[0]),
                      float(gscale[1]), p(sc.freqs), p(sc.phases), *F_._encode_args(sc, NR, K, 0, n), p(g_xr), None, p(d_lat),
                      p(d_gen), st)
            torch.cuda.synchronize()
            return d_lat, d_gen

        d_lat, d_gen = run(True, True)
        got_lat, got_gen = d_lat.permute(0, 1, 4, 2, 3).cpu(), d_gen.permute(2, 0, 1).cpu()
        assert got_lat.shape == want_dlat.shape and got_gen.shape == want_dgen.shape
        for got, want in ((got_lat, want_dlat), (got_gen, want_dgen)):
            err = (got - want).abs().max().item()
            assert err <= 1e-4 * max(want.abs().max().item(), 1.0), err      # fp32 sums in a different order
            assert (want != 0).sum() > 10000 and ((got != 0) == (want != 0)).float().mean() > 0.999
        only_lat, none_gen = run(True, False)
        none_lat, only_gen = run(False, True)
        assert none_gen is None and none_lat is None
        # same rows, same reductions: equal up to the order of the atomic additions
        assert torch.allclose(only_lat, d_lat, rtol=1e-5, atol=1e-5) and torch.allclose(only_gen, d_gen, rtol=1e-5, atol=1e-5)
    finally:
        novel_to_device(scene, "cpu")


@pytest.mark.parametrize("name", ALL)
def test_novel_training_step_against_reference_golden_gradients(name):
    """The NOVEL training step (novel.py:280-296,361-363: render with offsets from the batch, optimise nerf.parameters())
    on the reference's own samples: gradients of loss = sum(rgb * w_rgb) + sum(depth * w_dep) w.r.t. the MLP, gen_latent
    and encoder.latent against those of the UNMODIFIED reference (tests/golden/novel_*_grad.npz, oracle/make_golden_novel.py).

    The forward state is bf16 here (the variant has no split mode), and the synthetic density head is calibrated to logits
    of +-150, so -- exactly as for the DINER goldens (test_gpu_parity_r2.py::test_backward_against_reference_golden_gradients,
    bf16 case) -- the deviation is the loss's sensitivity to the forward value; the scatter itself is held to 1e-4 by
    test_novel_encode_bwd_vs_oracle_autograd and the MLP backward kernels by the DINER split-mode test.  Asserted here:
    direction and scale of every gradient, and that the gradients land on the right ENTRIES of both maps."""
    dev = _cuda()
    import parity_metrics as PM
    from diner_b200.novel import NeRFRendererDGS
    from tools.synthetic import novel_to_device
    g, scene, _ = _load(name)
    gg = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", f"{name}_grad.npz"), allow_pickle=False))
    wl = scene.workload
    novel_to_device(scene, dev)
    m = scene.model
    try:
        m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
        m.gen_latent = m.gen_latent.detach().requires_grad_(True)
        named = dict(m.mlp_fine.named_parameters())
        if name.startswith("novel_pe"):
            named.update({f"deformation_layer.{k}": v for k, v in m.deformation_layer.named_parameters()})
        for q in named.values():
            q.requires_grad_(True)
            q.grad = None
        r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        out = r(m, scene.rays.to(dev), scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets,
                noise=dict(z=torch.from_numpy(g["z"])))
        loss = PM.golden_loss(out.fine.rgb, out.fine.depth)
        loss.backward()
        torch.cuda.synchronize()
        # the training-mode forward (per-layer kernels, operands saved) to the tolerance of the evaluation path's test
        for key in ("rgb", "depth"):
            e = PM.error_metrics(out.fine[key], torch.from_numpy(g[key]))
            assert e["mean"] < 4e-3 and e["q99"] < 6e-2, (key, e)
        for k, v in named.items():
            ratio = v.grad.double().norm().item() / max(float(gg[f"gradnorm.{k}"]), 1e-300)
            assert 0.4 < ratio < 1.8, (k, ratio)
            if f"grad.{k}" in gg:
                e = PM.grad_metrics(v.grad, torch.from_numpy(gg[f"grad.{k}"]))
                assert e["cos"] > 0.6, (k, e)
            if f"gradpe.{k}" in gg:       # NOVEL_PE, 518-column weights: the six positional columns and a sub-sample of the rest
                e = PM.grad_metrics(v.grad[:, 512:], torch.from_numpy(gg[f"gradpe.{k}"]))
                assert e["cos"] > 0.6 and 0.4 < e["ratio"] < 1.8, (k, "positional columns", e)
                e = PM.grad_metrics(v.grad[:, :512].flatten()[::61], torch.from_numpy(gg[f"gradsub.{k}"]))
                assert e["cos"] > 0.6 and 0.4 < e["ratio"] < 1.8, (k, "feature columns", e)
        for key, t in (("glat", m.encoder.latent.grad), ("ggen", m.gen_latent.grad)):
            assert t is not None and torch.isfinite(t).all()
            pick = t.flatten()[torch.from_numpy(gg[f"{key}_idx"]).to(dev)]
            e = PM.grad_metrics(pick, torch.from_numpy(gg[f"{key}_val"]))
            assert e["cos"] > 0.8, (key, e)
            assert 0.4 < t.double().norm().item() / float(gg[f"{key}_norm"]) < 1.8
            # footprints: the entries the reference's gradient touches (ours may underflow to 0 on a few, not more)
            nnz = int((t != 0).sum())
            assert 0.9 * int(gg[f"{key}_nnz"]) <= nnz <= 1.001 * int(gg[f"{key}_nnz"]), (key, nnz, int(gg[f"{key}_nnz"]))
            want = torch.from_numpy(gg[f"{key}_val"]).to(dev)
            big = want.abs() > 1e-9 * want.abs().max()                # (entries of 1e-30 underflow in the bf16 stream)
            assert int(big.sum()) > 1000 and float((pick[big] != 0).float().mean()) > 0.99
        # chunking of the training step is invisible (two balanced chunks instead of one)
        first = {k: v.grad.clone() for k, v in named.items()}
        ggen_first = m.gen_latent.grad.clone()
        for q in (*named.values(), m.gen_latent, m.encoder.latent):
            q.grad = None
        r2 = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd,
                             rows_per_chunk=(scene.rays.shape[0] * scene.rays.shape[1] * wl.n_samples * get_nv(m)) // 2 + 1)
        out2 = r2(m, scene.rays.to(dev), scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets,
                  noise=dict(z=torch.from_numpy(g["z"])))
        assert torch.equal(out2.fine.rgb, out.fine.rgb)
        PM.golden_loss(out2.fine.rgb, out2.fine.depth).backward()
        for k, v in named.items():
            assert PM.grad_metrics(v.grad, first[k])["rel_l2"] < 1e-3, k
        assert PM.grad_metrics(m.gen_latent.grad, ggen_first)["rel_l2"] < 1e-4
    finally:
        m.encoder.latent = m.encoder.latent.detach()
        m.gen_latent = m.gen_latent.detach()
        for q in named.values():
            q.grad = None
        novel_to_device(scene, "cpu")


def get_nv(model):
    return int(model.poses.shape[1])


def test_novel_pe_operands_vs_oracle():
    """dnr_encode_novel_pe_fwd (NOVEL_PE, pe_novel_pixelnerf.py:200-275): the three operand images against the oracle's
    pieces -- encoder feature alone in lat_pk, gen_latent taps in gen_pk, and [code 55 | src pe 3 | tgt pe 3 | 0 3] in
    in_pk -- then the deformation GEMM: final_latent = gen + deformation_layer([feature | pe 6]) to bf16 operand accuracy."""
    dev = _cuda()
    from diner_b200 import _lib, field as F_
    from diner_b200.novel import NeRFRendererDGS, _deform, _gen_latent_nhwc, build_vertex_groups
    from diner_b200.scene import get_scene
    from oracle import dgs_oracle as O, novel_oracle as N
    from tools.synthetic import novel_to_device
    g, scene, _ = _load("novel_pe_tiny3")
    wl = scene.workload
    K = wl.n_samples
    rays = scene.rays[:, :50].contiguous()
    SB, NR = rays.shape[:2]
    z = torch.from_numpy(g["z"])[:, :50].contiguous()
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
    dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
    m = scene.model
    with torch.no_grad():
        p_in = N.deform_points(pts, scene.target_vertices, scene.tgt_in_offsets)
        p_gen = N.deform_points(pts, scene.target_vertices, scene.tgt_gen_offsets)
        zx = N.field_inputs_pe(m, p_in, p_gen, pts, dirs)                                   # SB,NV,B,573
        _, uv = O.project_to_views(m, p_in)
        NV = uv.shape[1]
        want_lat = O.tap_latent(m, uv).transpose(-1, -2)
        want_gen = N.tap_gen_latent(m, N.project_gen(m, p_gen, NV)).transpose(-1, -2)
    novel_to_device(scene, dev)
    try:
        sc = get_scene(m)
        rd, zd = rays.to(dev).view(-1, 8), z.to(dev).view(-1, K)
        V = scene.target_vertices.shape[1]
        pin, pgen, _ = _deform(rd, zd, None, None, 0, build_vertex_groups(scene.target_vertices),
                               scene.tgt_in_offsets.contiguous(), scene.tgt_gen_offsets.contiguous(), SB, NR, K, V, False)
        gen_nhwc = _gen_latent_nhwc(m)
        Hg, Wg = gen_nhwc.shape[:2]
        gsize = torch.tensor([Wg, Hg])
        gscale = ((gsize - float(m.encoder.feature_padding) * 2) / gsize).float()
        f = lambda t: t.detach().float().contiguous()
        gshape = m.gen_image_shape.cpu()
        gen = dict(poses=f(m.gen_poses), focal=f(m.gen_focal), cxy=f(m.gen_c), latent=gen_nhwc, Hg=Hg, Wg=Wg,
                   img_w=float(gshape[0]), img_h=float(gshape[1]), sx=float(gscale[0]), sy=float(gscale[1]),
                   pts_in=pin, pts_gen=pgen, pe=NeRFRendererDGS._pe_inputs(m, SB, NV, f))
        e = gen["pe"]
        p, st = _lib.ptr, _lib.stream_ptr()
        n = SB * NR
        mv = _lib.load().dnr_view_row_tiles(n * K, NV)
        bf = dict(device=dev, dtype=torch.bfloat16)
        lat, gpk, code = torch.empty(mv * 128 * 512, **bf), torch.empty(mv * 128 * 512, **bf), torch.empty(mv * 128 * 64, **bf)
        _lib.call("dnr_encode_novel_pe_fwd", p(rd), p(zd), p(pin), p(pgen), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth),
                  p(sc.latent_nhwc), p(gen["poses"]), p(gen["focal"]), p(gen["cxy"]), p(gen_nhwc), Hg, Wg, gen["img_w"],
                  gen["img_h"], gen["sx"], gen["sy"], p(e["src"]), e["Hp"], e["Wp"], e["sx"], e["sy"], p(e["tgt_poses"]),
                  p(e["tgt_focal"]), p(e["tgt_cxy"]), e["tgt_img_w"], e["tgt_img_h"], p(e["tgt"]), e["Ht"], e["Wt"],
                  e["tgt_sx"], e["tgt_sy"], p(sc.freqs), p(sc.phases), *F_._encode_args(sc, NR, K, 0, n), p(lat), p(gpk),
                  p(code), st)
        pk = F_.get_packed_pe(m, need_transposed=False)
        final = torch.empty(mv * 128 * 512, **bf)
        F_._gemm(lat, pk.wd, 512, mv, A1=code, B1=pk.wdc, K1=64, bias=pk.b_d, res_in_pk=gpk, act_out=final)

        def rows_of(buf, width):
            out = torch.empty(mv * 128, width, device=dev)
            _lib.call("dnr_unpack_rows", p(buf), mv * 128, width, p(out), st)
            return out

        lat_rows, gen_rows, code_rows, fin_rows = rows_of(lat, 512), rows_of(gpk, 512), rows_of(code, 64), rows_of(final, 512)
        torch.cuda.synchronize()
        spw = 32 // NV
        s_ = torch.arange(n * K)
        cat = lambda t, v: torch.cat([t[sb, v] for sb in range(SB)], 0)
        for v in range(NV):
            row = ((s_ // spw) * 32 + (s_ % spw) * NV + v).to(dev)
            w = cat(zx, v)                                       # [final 512 | src pe 3 | tgt pe 3 | code 55]
            wl_, wg_ = cat(want_lat, v), cat(want_gen, v)
            assert (lat_rows[row].cpu() - wl_).abs().max() <= 2.0 ** -7 * wl_.abs().max() + 1e-6
            assert (gen_rows[row].cpu() - wg_).abs().max() <= 2.0 ** -7 * wg_.abs().max() + 1e-6
            got_c = code_rows[row].cpu()
            assert torch.allclose(got_c[:, :55], w[:, 518:], rtol=2.0 ** -8, atol=2e-3)
            assert torch.allclose(got_c[:, 55:61], w[:, 512:518], rtol=2.0 ** -7, atol=2e-3)      # bilinear taps, bf16
            assert float(w[:, 512:518].abs().mean()) > 0.05 and (got_c[:, 61:] == 0).all()
            # bf16 operands into a 518-term fp32 sum, bf16 result
            got_f, want_f = fin_rows[row].cpu(), w[:, :512]
            assert (got_f - want_f).abs().max() <= 2.0 ** -5 * want_f.abs().max(), float((got_f - want_f).abs().max())
            assert (got_f - want_f).abs().mean() <= 2.0 ** -8 * want_f.abs().mean() * 4
    finally:
        novel_to_device(scene, "cpu")


@pytest.mark.parametrize("NR,neighbours", [(37, True), (37, False), (256, True)])
def test_candidate_tiling_does_not_change_the_result(NR, neighbours):
    """dnr_novel_deform, candidates mode: warps take 8 candidates x 4 neighbouring rays where that is more compact than 32
    candidates of one ray (decided per group of four rays).  Same offsets and vertex indices as the plain order
    (DNR_NOVEL_TILE=0), for ray counts that are not a multiple of four, neighbouring and scattered rays."""
    dev = _cuda()
    from diner_b200.novel import _deform, build_vertex_groups
    from diner_b200.scene import lin_table
    gen = torch.Generator().manual_seed(NR)
    SB, C, V = 2, 1000, 1500
    u = torch.randn(SB, V, 3, generator=gen)
    verts = u / u.norm(dim=-1, keepdim=True) * torch.tensor([0.11, 0.15, 0.12])
    off = torch.randn(SB, V, 3, generator=gen) * 0.01
    rays = torch.zeros(SB, NR, 8)
    rays[..., :3] = torch.tensor([0.0, 0.0, -1.7])
    if neighbours:          # adjacent pixels of one camera
        px = torch.arange(NR).float().view(1, NR, 1) * 0.0007
        d = torch.cat([px - 0.01, torch.zeros(1, NR, 1) + 0.003, torch.ones(1, NR, 1)], -1).expand(SB, -1, -1)
    else:
        d = torch.randn(SB, NR, 3, generator=gen) * 0.08 + torch.tensor([0.0, 0.0, 1.0])
    rays[..., 3:6] = d / d.norm(dim=-1, keepdim=True)
    rays[..., 6], rays[..., 7] = 1.0, 2.5
    u_coarse = torch.rand(SB * NR, C, generator=gen)
    rays, verts, off, u_coarse = (t.to(dev).contiguous() for t in (rays, verts, off, u_coarse))
    groups = build_vertex_groups(verts)
    run = lambda: _deform(rays.view(-1, 8), None, lin_table(C, dev), u_coarse, 0, groups, off, None, SB, NR, C, V, True,
                          want_idx=True)
    old = os.environ.get("DNR_NOVEL_TILE")
    try:
        os.environ["DNR_NOVEL_TILE"] = "0"
        want_off, _, want_idx = run()
        os.environ["DNR_NOVEL_TILE"] = "1"
        got_off, _, got_idx = run()
        torch.cuda.synchronize()
    finally:
        if old is None:
            os.environ.pop("DNR_NOVEL_TILE", None)
        else:
            os.environ["DNR_NOVEL_TILE"] = old
    assert torch.equal(got_idx, want_idx) and torch.equal(got_off, want_off)
    assert int(got_idx.unique().numel()) > 50 and int(got_idx.max()) < V


@pytest.mark.parametrize("name", ["novel_pe_tiny", "novel_pe_tiny3"])
def test_novel_pe_cluster_kernels_agree_with_per_layer_path(name):
    """NOVEL_PE evaluation: the cluster kernels with the second code K-step in fc_1 of blocks 0 / 1 (dnr_mlp_quad_fwd_pe,
    dnr_mlp_pair_fwd_pe) against the per-layer GEMMs (three K segments, field.PackedPE) on the same operands: pair == quad
    bit for bit, both within the DINER kernels' mutual tolerance of the per-layer result
    (test_gpu_field.py::test_inference_kernels_agree), deterministic, and the plain DINER path is not disturbed."""
    dev = _cuda()
    from diner_b200 import field as F_
    from diner_b200.novel import NeRFRendererDGS, _deform, _gen_latent_nhwc, build_vertex_groups
    from diner_b200.scene import get_scene
    from tools.synthetic import novel_to_device
    g, scene, _ = _load(name)
    wl = scene.workload
    K = wl.n_samples
    novel_to_device(scene, dev)
    m = scene.model
    keep = (F_.FORWARD_KERNEL, F_.PE_FORWARD_LAYERS)
    try:
        sc = get_scene(m)
        SB, NR = scene.rays.shape[:2]
        NV = sc.NV
        rd = scene.rays.to(dev).view(-1, 8)
        zd = torch.from_numpy(g["z"]).view(-1, K).to(dev)
        V = scene.target_vertices.shape[1]
        pin, pgen, _ = _deform(rd, zd, None, None, 0, build_vertex_groups(scene.target_vertices),
                               scene.tgt_in_offsets.contiguous(), scene.tgt_gen_offsets.contiguous(), SB, NR, K, V, False)
        gen_nhwc = _gen_latent_nhwc(m)
        Hg, Wg = gen_nhwc.shape[:2]
        gsize = torch.tensor([Wg, Hg])
        gscale = ((gsize - float(m.encoder.feature_padding) * 2) / gsize).float()
        f = lambda t: t.detach().float().contiguous()
        gshape = m.gen_image_shape.cpu()
        novel = dict(poses=f(m.gen_poses), focal=f(m.gen_focal), cxy=f(m.gen_c), latent=gen_nhwc, Hg=Hg, Wg=Wg,
                     img_w=float(gshape[0]), img_h=float(gshape[1]), sx=float(gscale[0]), sy=float(gscale[1]),
                     pts_in=pin, pts_gen=pgen, pe=NeRFRendererDGS._pe_inputs(m, SB, NV, f))
        pk = F_.get_packed_pe(m, need_transposed=False)
        outs = {}
        for kern in ("layers", "pair", "quad"):
            F_.FORWARD_KERNEL, F_.PE_FORWARD_LAYERS = ("quad", True) if kern == "layers" else (kern, False)
            outs[kern], _ = F_.field_forward(sc, pk, rd, zd, NR, 0, save=False, novel=novel)
            again, _ = F_.field_forward(sc, pk, rd, zd, NR, 0, save=False, novel=novel)
            torch.cuda.synchronize()
            assert torch.isfinite(outs[kern]).all() and torch.equal(outs[kern], again), kern
        assert torch.equal(outs["quad"], outs["pair"])
        a, b = outs["pair"], outs["layers"]
        smax = float(b[:, 3].max())
        assert smax > 1.0
        assert float((a[:, :3] - b[:, :3]).abs().max()) < 1.5e-2 and float((a[:, :3] - b[:, :3]).abs().mean()) < 5e-4
        assert float((a[:, 3] - b[:, 3]).abs().max()) < 1e-2 * smax and float((a[:, 3] - b[:, 3]).abs().mean()) < 5e-4 * smax
        # the positional columns matter: zeroing the second code K-step's weights moves the result
        import ctypes as C
        zero = torch.zeros_like(pk.w_zc128[0])
        pk_zero = (C.c_void_p * 2)(zero.data_ptr(), zero.data_ptr())
        real, pk.zc128_ptrs = pk.zc128_ptrs, pk_zero
        try:
            F_.FORWARD_KERNEL, F_.PE_FORWARD_LAYERS = "quad", False
            moved, _ = F_.field_forward(sc, pk, rd, zd, NR, 0, save=False, novel=novel)
        finally:
            pk.zc128_ptrs = real
        assert float((moved - outs["quad"]).abs().max()) > 1e-3
    finally:
        F_.FORWARD_KERNEL, F_.PE_FORWARD_LAYERS = keep
        novel_to_device(scene, "cpu")
