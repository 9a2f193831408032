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
                                       (64, 100, True)])
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
    from diner_b200.synthetic import make_scene
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
    from diner_b200.synthetic import make_scene
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
