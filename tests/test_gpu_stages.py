"""GPU parity tests, stage by stage, through the C ABI (ctypes) -- run with `-m gpu` on a B200.

Checkers: the CPU oracle (oracle/dgs_oracle.py, bit-pinned to the reference's outputs by tests/test_oracle.py),
the same oracle executed with torch CUDA ops (the reference's own deployment backend), and the
precision-contract oracle (oracle/bf16_field.py) for the bf16 tensor-core MLP.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CASES = ["tiny", "tiny3", "plumbing64"]


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()           # fails loudly if the extension is missing: there is no fallback
    return torch.device("cuda", 0)


def _full_noise(g, wl, n_rays, seed=5):
    gen = torch.Generator().manual_seed(seed)
    return dict(u_coarse=torch.from_numpy(g["u_coarse"]),
                n_gauss_full=torch.randn(n_rays, wl.n_gaussian, generator=gen),
                u_fill_full=torch.rand(n_rays, wl.n_samples, generator=gen))


def _oracle_sampler(model, rays, wl, noise):
    from oracle import dgs_oracle as O
    with torch.no_grad():
        z0, _, dbg = O.sample_depthguided(model, rays, wl.n_samples, wl.n_candidates, wl.n_gaussian, noise=dict(noise))
        z, _ = O.fill_up_uniform(z0.clone(), rays, None, noise["u_fill_full"])
    K, M = wl.n_samples, wl.n_samples - wl.n_gaussian
    order = dbg.order.view(-1, K)[:, :M]
    p = dbg.p.view(-1, wl.n_candidates)
    idx = torch.where(torch.gather(p, 1, order) > 0, order, torch.full_like(order, -1))
    return dict(z=z.view(-1, K), p=p, idx=idx, rm=dbg.ray_mask.view(-1).int(), nm=(z0.view(-1, K) == 0).sum(-1).int())


# ---------------------------------------------------------------------------------------------------------
# compositing (fp32, 1e-4 relative -- BASELINE.json north_star)
# ---------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("n,K,white", [(1000, 40, True), (1000, 40, False), (37, 1, True), (129, 33, False),
                                       (64, 100, True), (1001, 40, True), (1, 40, True), (77, 48, True),
                                       (77, 47, False), (50, 2, True), (33, 50, True), (33, 64, False)])
def test_composite_fwd_bwd_matches_oracle(n, K, white):
    dev = _cuda()
    from diner_b200.renderer import _CompositeFn
    from oracle import dgs_oracle as O
    torch.manual_seed(n + K)
    z = torch.sort(torch.rand(n, K) * 1.5 + 1.0, -1).values
    rays = torch.zeros(n, 8); rays[:, 6] = 1.0; rays[:, 7] = 2.5
    rs = torch.rand(n, K, 4); rs[..., 3] = torch.relu(torch.randn(n, K) * 30)
    a = rs.clone().requires_grad_(True)
    w, rgb, dep = O.alpha_composite(a.view(1, n, K, 4), z.view(1, n, K), rays.view(1, n, 8), white)
    gr, gd, gw = torch.randn(1, n, 3), torch.randn(1, n), torch.randn(1, n, K)
    ((rgb * gr).sum() + (dep * gd).sum() + (w * gw).sum()).backward()
    b = rs.clone().to(dev).requires_grad_(True)
    rgb2, dep2, w2 = _CompositeFn.apply(b.view(n * K, 4), z.to(dev), rays.to(dev), white)
    ((rgb2 * gr[0].to(dev)).sum() + (dep2 * gd[0].to(dev)).sum() + (w2 * gw[0].to(dev)).sum()).backward()
    tol = dict(rtol=1e-4, atol=1e-5)            # north_star: 1e-4 relative in fp32
    torch.testing.assert_close(rgb2.cpu(), rgb[0], **tol)
    torch.testing.assert_close(dep2.cpu(), dep[0], **tol)
    torch.testing.assert_close(w2.cpu(), w[0], **tol)
    torch.testing.assert_close(b.grad.cpu(), a.grad, rtol=1e-4, atol=1e-5 * float(a.grad.abs().max()))


def test_composite_known_answers_and_empty():
    dev = _cuda()
    from diner_b200.renderer import _CompositeFn
    K = 40
    z = torch.linspace(1.0, 2.0, K).view(1, K).to(dev)
    rays = torch.tensor([[0, 0, 0, 0, 0, 1, 1.0, 2.5]], device=dev)
    sig = 3.0
    rs = torch.cat([torch.full((K, 3), 0.25), torch.full((K, 1), sig)], -1).to(dev)
    rgb, depth, w = _CompositeFn.apply(rs, z, rays, False)
    want = 1 - np.exp(-sig * 1.5)               # constant density: sum w = 1 - exp(-sigma (far - z0))
    assert abs(w.sum().item() - want) < 1e-5 and abs(rgb[0, 0].item() - 0.25 * want) < 1e-5
    rs = torch.zeros(K, 4, device=dev)
    rs[7] = torch.tensor([0.1, 0.5, 0.9, 1e6])  # one opaque sample
    rgb, depth, w = _CompositeFn.apply(rs, z, rays, True)
    assert torch.allclose(rgb[0].cpu(), torch.tensor([0.1, 0.5, 0.9]), atol=1e-6)
    assert abs(depth.item() - z[0, 7].item()) < 1e-6
    rgb, depth, w = _CompositeFn.apply(torch.zeros(0, 4, device=dev), torch.zeros(0, K, device=dev),
                                       torch.zeros(0, 8, device=dev), True)
    assert rgb.shape == (0, 3) and depth.shape == (0,)


# ---------------------------------------------------------------------------------------------------------
# sampler: indices / sort / merge order bit-exact against the reference algorithm on torch CUDA ops
# ---------------------------------------------------------------------------------------------------------

@pytest.mark.parametrize("name", CASES)
def test_sampler_bit_exact_vs_cuda_oracle(golden_cache, name):
    dev = _cuda()
    from diner_b200 import sample_rays
    from diner_b200.scene import get_scene
    g, scene, rays, _ = golden_cache(name)
    wl = scene.workload
    noise = _full_noise(g, wl, rays.shape[0] * rays.shape[1])
    oc = _oracle_sampler(scene.model, rays, wl, noise)                                    # torch CPU
    scene.model.to_device(dev)
    try:
        noise_d = {k: v.to(dev) for k, v in noise.items()}
        og = _oracle_sampler(scene.model, rays.to(dev), wl, noise_d)                      # torch CUDA
        z, d = sample_rays(get_scene(scene.model), rays.to(dev), wl.n_samples, wl.n_candidates, wl.n_gaussian,
                           noise=noise_d, want_debug=True)
        # per-candidate likelihoods, shortlist indices in rank order, masks and fill counts: bit-exact
        assert torch.equal(d["p"], og["p"])
        assert torch.equal(d["sel_idx"].long(), og["idx"])
        assert torch.equal(d["ray_mask"], og["rm"]) and torch.equal(d["n_missing"], og["nm"])
        # values: shortlisted / uniform samples are bit-exact, gaussian draws differ by the reduction order of
        # the two moments (torch's parallel sum vs. a warp tree) -> fp32 round-off only
        torch.testing.assert_close(z, og["z"], rtol=1e-5, atol=1e-6)
        assert bool((z[:, 1:] >= z[:, :-1]).all())
        # against the CPU oracle torch itself is not bit-stable (CPU vs CUDA erf / matmul rounding flips a few
        # texels); most rays must still agree
        close = torch.isclose(z.cpu(), oc["z"], rtol=1e-5, atol=1e-6).all(-1).float().mean().item()
        cpu_vs_cuda = torch.isclose(og["z"].cpu(), oc["z"], rtol=1e-5, atol=1e-6).all(-1).float().mean().item()
        assert close >= 0.9 and abs(close - cpu_vs_cuda) < 0.02
    finally:
        scene.model.to_device("cpu")


@pytest.mark.parametrize("name", CASES)
def test_sampler_packed_reference_noise_vs_golden(golden_cache, name):
    """The reference's own recorded draws, packed in the reference's order.  (a) With each ray's slice located by
    the CPU reference's own masks, every ray on which torch-CPU and CUDA agree reproduces the golden z; (b) the
    counting pass (dnr_sampler_fwd count_only=1 + prefix sums) locates the same slices wherever the masks agree."""
    dev = _cuda()
    from diner_b200 import sample_rays
    from diner_b200.scene import get_scene
    from oracle import dgs_oracle as O
    g, scene, rays, noise = golden_cache(name)
    wl = scene.workload
    K = wl.n_samples
    with torch.no_grad():
        z0, _, dbg = O.sample_depthguided(scene.model, rays, K, wl.n_candidates, wl.n_gaussian, noise=dict(noise))
    rm = dbg.ray_mask.view(-1).int()
    nm = (z0.view(-1, K) == 0).sum(-1).int()
    scene.model.to_device(dev)
    try:
        sc = get_scene(scene.model)
        nz = {k: v.to(dev) for k, v in noise.items()}
        gz = torch.from_numpy(g["z"]).view(-1, K)
        za = sample_rays(sc, rays.to(dev), K, wl.n_candidates, wl.n_gaussian,
                         noise=dict(nz, gauss_off=torch.cumsum(rm, 0) - rm, fill_off=torch.cumsum(nm, 0) - nm))
        ok = torch.isclose(za.cpu(), gz, rtol=1e-5, atol=1e-6).all(-1)
        assert ok.float().mean().item() >= 0.9
        zb, d = sample_rays(sc, rays.to(dev), K, wl.n_candidates, wl.n_gaussian, noise=nz, want_debug=True)
        assert bool((zb[:, 1:] >= zb[:, :-1]).all()) and torch.isfinite(zb).all()
        agree = (d["ray_mask"].cpu() == rm) & (d["n_missing"].cpu() == nm)
        prefix = int((~agree).nonzero()[0]) if (~agree).any() else len(agree)
        assert torch.equal(za[:prefix], zb[:prefix])
    finally:
        scene.model.to_device("cpu")


def test_sampler_philox_jitter_equals_explicit_uniforms():
    """Performance mode (in-kernel Philox, four candidates per call) scores exactly the candidates the parity mode
    scores when it is handed the same uniforms: likelihoods and shortlist indices bit-equal."""
    dev = _cuda()
    from diner_b200 import sample_rays
    from diner_b200.scene import get_scene
    from tools.synthetic import make_scene
    from oracle.philox import philox4x32_10 as _philox4x32_10
    scene = make_scene("tiny", seed=0, calibrate=False)
    wl = scene.workload
    scene.model.to_device(dev)
    rays = scene.rays.to(dev)
    sc = get_scene(scene.model)
    K, C, G = wl.n_samples, wl.n_candidates, wl.n_gaussian
    n = rays.shape[0] * rays.shape[1]
    for C_ in (C, 998, 130):                         # C not a multiple of 4 / of 128: ragged last pass
        seed = 0x1234_5678_9ABC_DEF0 + C_
        _, d = sample_rays(sc, rays, K, C_, G, seed=seed, want_debug=True)
        ray = np.arange(n, dtype=np.uint64)[:, None]
        blk = np.arange((C_ + 3) // 4, dtype=np.uint64)[None, :]
        r = _philox4x32_10(ray & np.uint64(0xFFFFFFFF), ray >> np.uint64(32), blk + 0 * ray, 0 * blk + 0 * ray,
                           seed & 0xFFFFFFFF, seed >> 32)
        u = np.stack(r, -1).reshape(n, -1)[:, :C_]
        u = ((u & np.uint64(0xFFFFFF)).astype(np.float32) * np.float32(1.0 / 16777216.0)).astype(np.float32)
        noise = dict(u_coarse=torch.from_numpy(u), n_gauss_full=torch.zeros(n, G), u_fill_full=torch.zeros(n, K))
        _, d2 = sample_rays(sc, rays, K, C_, G, noise=noise, want_debug=True)
        assert int((d["p"] != 0).sum()) > 0
        assert torch.equal(d["p"], d2["p"]) and torch.equal(d["sel_idx"], d2["sel_idx"])
        assert torch.equal(d["ray_mask"], d2["ray_mask"])


def test_sampler_background_rays_are_stratified_uniform():
    """A ray that sees no depth (std == 0 everywhere) gets K stratified samples in [near, far)."""
    dev = _cuda()
    from diner_b200 import sample_rays
    from diner_b200.scene import get_scene
    from tools.synthetic import make_scene
    scene = make_scene("tiny", seed=0, calibrate=False)
    scene.model.encoder.depths_std = torch.zeros_like(scene.model.encoder.depths_std)
    wl = scene.workload
    scene.model.to_device(dev)
    rays = scene.rays.to(dev)
    z, d = sample_rays(get_scene(scene.model), rays, wl.n_samples, wl.n_candidates, wl.n_gaussian, seed=7,
                       want_debug=True)
    K = wl.n_samples
    step = (wl.far - wl.near) / K
    lo = wl.near + step * torch.arange(K, device=dev)
    assert bool((z >= lo - 1e-6).all()) and bool((z <= lo + step + 1e-6).all())
    assert int(d["ray_mask"].sum()) == 0 and bool((d["n_missing"] == K).all()) and bool((d["sel_idx"] == -1).all())
    z2 = sample_rays(get_scene(scene.model), rays, wl.n_samples, wl.n_candidates, wl.n_gaussian, seed=7)
    assert torch.equal(z, z2), "Philox jitter must be deterministic for a fixed seed"
    z3 = sample_rays(get_scene(scene.model), rays, wl.n_samples, wl.n_candidates, wl.n_gaussian, seed=8)
    assert not torch.equal(z, z3)


# ------------------------------------------------------------------------------------------------ gather stage
@pytest.mark.parametrize("NV,n_rays", [(1, 37), (2, 100), (3, 45), (4, 64), (2, 1)])
def test_encode_stage_fwd_bwd_all_view_counts(NV, n_rays):
    """Projection + gather + geometric code (dnr_encode_fwd) and its backward scatter (dnr_encode_bwd) for every view
    count, on ragged sample counts (padding rows in the last slab / tile): forward against the fp32 oracle's
    field_inputs (bf16 rounding of the operands is the only difference), backward against autograd through the oracle's
    grid_sample on the same upstream gradient."""
    dev = _cuda()
    import dataclasses
    from diner_b200 import _lib
    from diner_b200.field import _encode_args
    from diner_b200.scene import get_scene
    from tools.synthetic import WORKLOADS, make_scene
    from oracle import dgs_oracle as O
    wl = dataclasses.replace(WORKLOADS["tiny"], name=f"enc_nv{NV}", NV=NV, tgt_H=16, tgt_W=16)
    scene = make_scene(wl, seed=10 + NV)
    K = wl.n_samples
    rays = scene.rays[:, :n_rays].contiguous()
    gen = torch.Generator().manual_seed(NV)
    # depth-guided-like samples: most of a ray's samples cluster (shared bilinear footprints), the rest spread out
    u = torch.cat([0.45 + 0.02 * torch.rand(1, n_rays, K - 10, generator=gen), torch.rand(1, n_rays, 10, generator=gen)], -1)
    z = rays[..., 6:7] + (rays[..., 7:8] - rays[..., 6:7]) * u.sort(-1).values
    SB, NR = 1, n_rays
    pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
    dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
    m = scene.model
    lat_leaf = m.encoder.latent.detach().clone().requires_grad_(True)
    m.encoder.latent = lat_leaf
    zx = O.field_inputs(m, pts, dirs)                                  # SB,NV,B,567 (fp32, torch CPU)
    G = torch.randn(SB, NV, NR * K, 512, generator=gen)
    (zx[..., :512] * G).sum().backward()
    want_dlat = lat_leaf.grad.clone()                                  # SB,NV,512,Hl,Wl
    m.encoder.latent = lat_leaf.detach()
    m.to_device(dev)
    try:
        sc = get_scene(m)
        p, st = _lib.ptr, _lib.stream_ptr()
        ns = NR * K
        mv = _lib.load().dnr_view_row_tiles(ns, NV)
        lat = torch.empty(mv * 128 * 512, device=dev, dtype=torch.bfloat16)
        code = torch.empty(mv * 128 * 64, device=dev, dtype=torch.bfloat16)
        r_d, z_d = rays.view(-1, 8).to(dev), z.view(-1, K).to(dev)
        _lib.call("dnr_encode_fwd", p(r_d), p(z_d), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth), p(sc.latent_nhwc),
                  p(sc.freqs), p(sc.phases), *_encode_args(sc, NR, K, 0, NR), p(lat), p(code), st)
        lat_rows = torch.empty(mv * 128, 512, device=dev)
        code_rows = torch.empty(mv * 128, 64, device=dev)
        _lib.call("dnr_unpack_rows", p(lat), mv * 128, 512, p(lat_rows), st)
        _lib.call("dnr_unpack_rows", p(code), mv * 128, 64, p(code_rows), st)
        spw = 32 // NV
        s_ = torch.arange(ns)
        g_rows = torch.zeros(mv * 128, 512)
        valid = torch.zeros(mv * 128, dtype=torch.bool)
        for v in range(NV):
            row = (s_ // spw) * 32 + (s_ % spw) * NV + v
            valid[row] = True
            w = zx[0, v].detach()
            got_l, got_c = lat_rows[row.to(dev)].cpu(), code_rows[row.to(dev)].cpu()
            # operands are bf16: relative 2^-9 of the value, plus the latent itself stored in bf16 before blending
            assert (got_l - w[:, :512]).abs().max() <= 2.0 ** -7 * w[:, :512].abs().max() + 1e-6
            assert torch.allclose(got_c[:, :55], w[:, 512:], rtol=2.0 ** -8, atol=2e-3)      # sin of large arguments
            assert (got_c[:, 55:] == 0).all()
            g_rows[row] = G[0, v]
        assert (lat_rows[(~valid).to(dev)] == 0).all() and (code_rows[(~valid).to(dev)] == 0).all()    # padding rows
        # backward: upstream gradient in XR layout -> NHWC fp32 latent gradient
        g_xr = torch.empty(mv * 128 * 512, device=dev)
        _lib.call("dnr_rows_to_xr", p(g_rows.to(dev).contiguous()), mv * 128, 512, mv * 128, p(g_xr), st)
        d_nhwc = torch.zeros(sc.SB, sc.NV, sc.Hl, sc.Wl, sc.L, device=dev)
        _lib.call("dnr_encode_bwd", p(r_d), p(z_d), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth), p(sc.freqs),
                  p(sc.phases), *_encode_args(sc, NR, K, 0, NR), p(g_xr), p(d_nhwc), st)
        torch.cuda.synchronize()
        got = d_nhwc.permute(0, 1, 4, 2, 3).cpu()
        assert got.shape == want_dlat.shape
        err = (got - want_dlat).abs().max().item()
        assert err <= 1e-4 * max(want_dlat.abs().max().item(), 1.0), err      # fp32 sums in a different order
        assert (got != 0).sum() > 0 and ((got != 0) == (want_dlat != 0)).float().mean() > 0.999
    finally:
        m.to_device("cpu")
