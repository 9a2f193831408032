"""Full-size, size-independent property tests of the whole render path at BASELINE.json shapes (run with -m gpu).

The oracle cannot render a full frame in seconds, so these check what the domain guarantees: sorted samples in
range, weights that sum to <= 1, white/black background identity, determinism for a fixed jitter seed,
invariance to how the frame is chunked or sharded, a frame-sized training step with finite gradients, and a
strided sample of rays against the CPU oracle.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()
    return torch.device("cuda", 0)


@pytest.fixture(scope="module")
def facescape():
    from tools.synthetic import make_scene
    dev = _cuda()
    scene = make_scene("facescape", seed=0)
    scene.model.to_device(dev)
    return scene


def test_full_frame_properties(facescape):
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS, sample_rays
    from diner_b200.scene import get_scene
    scene = facescape
    wl = scene.workload
    rays = scene.rays.to(dev)                                  # (1, 65536, 8)
    K = wl.n_samples
    sc = get_scene(scene.model)
    z, d = sample_rays(sc, rays, K, wl.n_candidates, wl.n_gaussian, seed=11, want_debug=True)
    assert bool((z[:, 1:] >= z[:, :-1]).all())
    fg = d["ray_mask"].bool()
    frac_fg = fg.float().mean().item()
    assert 0.3 < frac_fg < 0.9                                  # non-degenerate scene
    bgz = z[~fg]
    assert bool((bgz >= wl.near).all()) and bool((bgz < wl.far + 1e-5).all())
    # foreground rays concentrate their samples: the shortlisted + gaussian ones sit within a few cm
    spread = (z[fg].quantile(0.75, dim=1) - z[fg].quantile(0.25, dim=1)).median().item()
    assert spread < 0.05 * (wl.far - wl.near)

    r = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=True)
    with torch.no_grad():
        o1 = r(scene.model, rays, want_weights=True, noise=dict(z=z))
        o2 = r(scene.model, rays, want_weights=True, noise=dict(z=z))
    w = o1.fine.weights
    assert torch.isfinite(o1.fine.rgb).all() and torch.isfinite(o1.fine.depth).all()
    assert torch.equal(o1.fine.rgb, o2.fine.rgb) and torch.equal(w, o2.fine.weights)      # deterministic
    sw = w.sum(-1)
    assert float(w.min()) >= 0 and float(sw.max()) <= 1 + 1e-4
    assert float(sw.min()) < 0.1 and float(sw.max()) > 0.9                                  # live density head
    # white background adds exactly (1 - sum w) to every channel
    rb = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=False)
    with torch.no_grad():
        ob = rb(scene.model, rays, noise=dict(z=z))
    torch.testing.assert_close(o1.fine.rgb, ob.fine.rgb + (1 - sw).unsqueeze(-1), rtol=1e-5, atol=1e-5)
    # depth is a convex-ish combination of the samples
    assert float(o1.fine.depth.max()) <= wl.far + 1e-3
    # chunking / sharding invariance: any split of the frame gives the same pixels
    from diner_b200.sharding import shard_range
    small = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=True, rows_per_chunk=1 << 17)
    lo, hi = shard_range(rays.shape[1], 1, 3)
    with torch.no_grad():
        part = small(scene.model, rays[:, lo:hi].contiguous(), noise=dict(z=z.view(1, -1, K)[:, lo:hi]))
    assert torch.equal(part.fine.rgb, o1.fine.rgb[:, lo:hi])


def test_strided_rays_against_cpu_oracle(facescape):
    """256 rays spread over the full frame: sampler (unpacked jitter) and rendered colour vs. the CPU oracle."""
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS
    from oracle import dgs_oracle as O
    scene = facescape
    wl = scene.workload
    K = wl.n_samples
    sub = scene.rays[:, 37::256].contiguous()                   # 256 rays
    n = sub.shape[1]
    gen = torch.Generator().manual_seed(9)
    noise = dict(u_coarse=torch.rand(n, wl.n_candidates, generator=gen),
                 n_gauss_full=torch.randn(n, wl.n_gaussian, generator=gen), u_fill_full=torch.rand(n, K, generator=gen))
    scene.model.to_device("cpu")
    try:
        with torch.no_grad():
            z0, _, _ = O.sample_depthguided(scene.model, sub, K, wl.n_candidates, wl.n_gaussian, noise=dict(noise))
            z_ref, _ = O.fill_up_uniform(z0, sub, None, noise["u_fill_full"])
            _, rgb_ref, depth_ref, _ = O.composite(scene.model, sub, z_ref, wl.white_bkgd)
    finally:
        scene.model.to_device(dev)
    r = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    with torch.no_grad():
        out = r(scene.model, sub.to(dev), noise={k: v.to(dev) for k, v in noise.items()}, return_samples=True)
        out_z = r(scene.model, sub.to(dev), noise=dict(z=z_ref))
    same = torch.isclose(out.fine.z.cpu().view(-1, K), z_ref.view(-1, K), rtol=1e-5, atol=1e-6).all(-1)
    assert same.float().mean().item() >= 0.9                     # rest: torch CPU-vs-CUDA rounding
    assert float((out_z.fine.rgb.cpu() - rgb_ref).abs().mean()) < 3e-3
    assert float((out_z.fine.depth.cpu() - depth_ref).abs().mean()) < 3e-3
    agree = same.view(1, -1)
    assert float((out.fine.rgb.cpu() - rgb_ref)[agree].abs().mean()) < 3e-3


def test_training_step_shape():
    """BASELINE config 4: SB=2 scenes, one 64x64 patch each, forward + backward through all four stages."""
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS
    from tools.synthetic import make_scene
    scene = make_scene("facescape_train", seed=0)
    wl = scene.workload
    m = scene.model.to_device(dev)
    m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    rays = scene.rays.to(dev)
    assert rays.shape == (2, 4096, 8)
    out = r(m, rays)
    gt = torch.rand_like(out.fine.rgb)
    loss = torch.nn.functional.mse_loss(out.fine.rgb, gt)
    loss.backward()
    g = m.encoder.latent.grad
    assert g.shape == m.encoder.latent.shape and torch.isfinite(g).all() and float(g.abs().sum()) > 0
    # both scenes receive feature gradients, and only inside their own maps
    assert float(g[0].abs().sum()) > 0 and float(g[1].abs().sum()) > 0
    for name, p in m.mlp_fine.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all(), name
        assert float(p.grad.abs().sum()) > 0, name
    # a second identical step accumulates exactly like autograd does
    g1 = m.mlp_fine.lin_out.weight.grad.clone()
    out = r(m, rays, noise=None)
    torch.nn.functional.mse_loss(out.fine.rgb, gt).backward()
    assert float((m.mlp_fine.lin_out.weight.grad - g1).abs().sum()) > 0


def test_empty_and_tiny_ray_batches():
    """Zero rays give empty outputs of the right shapes; 1, 2 and 129 rays (less than one tile, ragged) render and
    back-propagate; a renderer whose n_samples / n_gaussian were mutated after construction (as
    python_scripts/create_prediction_folder.py:51-52 does) uses the new values."""
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS
    from tools.synthetic import make_scene
    scene = make_scene("tiny", seed=0)
    wl = scene.workload
    m = scene.model.to_device(dev)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=True)
    for n in (0, 1, 2, 129):
        rays = scene.rays[:, :n].contiguous().to(dev)
        with torch.no_grad():
            out = r(m, rays, want_weights=True)
        assert out.fine.rgb.shape == (1, n, 3) and out.fine.depth.shape == (1, n)
        assert out.fine.weights.shape == (1, n, wl.n_samples) and torch.isfinite(out.fine.rgb).all()
        m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
        out2 = r(m, rays, want_weights=True)
        if n:
            (out2.fine.rgb.sum() + out2.fine.depth.sum()).backward()
            assert torch.isfinite(m.encoder.latent.grad).all()
        m.encoder.latent = m.encoder.latent.detach()
    r.n_samples, r.n_gaussian = 24, 8
    with torch.no_grad():
        out = r(m, scene.rays[:, :50].contiguous().to(dev), want_weights=True, return_samples=True)
    assert out.fine.weights.shape == (1, 50, 24) and out.fine.z.shape == (1, 50, 24)
    assert bool((out.fine.z[..., 1:] >= out.fine.z[..., :-1]).all()) and torch.isfinite(out.fine.rgb).all()


@pytest.mark.parametrize("workload,max_rays", [("dtu", None), ("multiface", 131072)])
def test_full_frame_properties_other_shapes(workload, max_rays):
    """BASELINE.json configs 2 and 5: DTU shape (3 views of 300 x 400, black background, near / far 0.32 / 1.20) at its
    full 120 000 rays, Multiface shape (4 views of 1024 x 672, std 1e-3) on 131 072 of its 1 M rays -- the same
    size-independent properties as the FaceScape frame, plus agreement of the three inference MLP kernels."""
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS, field as F_, sample_rays
    from diner_b200.scene import get_scene
    from tools.synthetic import make_scene
    scene = make_scene(workload, seed=0)
    m = scene.model.to_device(dev)
    try:
        wl = scene.workload
        rays = scene.rays.to(dev)
        if max_rays is not None:                                   # a centred band of the frame
            mid = rays.shape[1] // 2
            rays = rays[:, mid - max_rays // 2: mid + max_rays // 2].contiguous()
        K = wl.n_samples
        sc = get_scene(m)
        z, d = sample_rays(sc, rays, K, wl.n_candidates, wl.n_gaussian, seed=3, want_debug=True)
        assert bool((z[:, 1:] >= z[:, :-1]).all())
        fg = d["ray_mask"].bool()
        assert 0.05 < fg.float().mean().item() < 0.95
        bgz = z[~fg]
        assert bool((bgz >= wl.near).all()) and bool((bgz < wl.far + 1e-5).all())
        r = NeRFRendererDGS(K, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        outs = {}
        keep = F_.FORWARD_KERNEL
        try:
            for kern in ("quad", "pair"):
                F_.FORWARD_KERNEL = kern
                with torch.no_grad():
                    outs[kern] = r(m, rays, want_weights=True, noise=dict(z=z))
        finally:
            F_.FORWARD_KERNEL = keep
        o = outs["quad"]
        assert torch.equal(o.fine.rgb, outs["pair"].fine.rgb) and torch.equal(o.fine.weights, outs["pair"].fine.weights)
        w = o.fine.weights
        sw = w.sum(-1)
        assert torch.isfinite(o.fine.rgb).all() and torch.isfinite(o.fine.depth).all()
        assert float(w.min()) >= 0 and float(sw.max()) <= 1 + 1e-4
        assert float(sw.max()) > 0.9                               # the calibrated density head is live
        if not wl.white_bkgd:                                      # black background: colour bounded by the weight sum
            assert bool((o.fine.rgb.amax(-1) <= sw + 1e-5).all())
        assert float(o.fine.depth.max()) <= wl.far + 1e-3 and float(o.fine.depth.min()) >= 0
    finally:
        m.to_device("cpu")


@pytest.mark.parametrize("workload,n_rays,rows", [("facescape", 65536, None), ("facescape", 4096, None),
                                                  ("facescape", 20000, 1 << 19), ("tiny3", 288, None), ("tiny", 129, 1 << 12),
                                                  ("multiface", 6000, None), ("tiny", 1, None)])
def test_single_call_render_equals_call_by_call_sequence(workload, n_rays, rows):
    """dnr_render_fwd (csrc/render.cu: the evaluation render as ONE C-ABI call, sequenced by the library) against the
    same launches issued call by call from Python (renderer.py / field.py): rgb, depth, weights and samples are the same
    bits, for the in-kernel jitter (same seed) and for caller-supplied samples; view counts 2 / 3 / 4, one and several
    chunks (with and without a 2-CTA-cluster share), ragged last tiles, two scenes per call."""
    dev = _cuda()
    import diner_b200.renderer as R
    from diner_b200 import NeRFRendererDGS, _lib
    from tools.synthetic import make_scene
    scene = make_scene(workload, seed=3)
    wl = scene.workload
    m = scene.model.to_device(dev)
    per_scene = max(1, n_rays // wl.SB)
    rays = scene.rays[:, :per_scene].contiguous().to(dev)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd, rows_per_chunk=rows)
    old = R.USE_NATIVE_FRAME
    try:
        outs = {}
        with torch.no_grad():
            r(m, rays)                                           # (per-scene / per-weights repacks happen here)
        for native in (True, False):
            R.USE_NATIVE_FRAME = native
            n0 = _lib.launch_count
            with torch.no_grad():
                a = r(m, rays, want_weights=True, return_samples=True, noise=dict(seed=77))
                b = r(m, rays, noise=dict(z=a.fine.z))
                torch.manual_seed(5)
                c = r(m, rays)                                   # seed drawn from torch's CPU generator
            outs[native] = (a, b, c, _lib.launch_count - n0)
        torch.cuda.synchronize()
    finally:
        R.USE_NATIVE_FRAME = old
        m.to_device("cpu")
    (a1, b1, c1, l1), (a0, b0, c0, l0) = outs[True], outs[False]
    for k in ("rgb", "depth", "weights", "z"):
        assert torch.equal(a1.fine[k], a0.fine[k]), k
    for k in ("rgb", "depth"):
        assert torch.equal(b1.fine[k], b0.fine[k]) and torch.equal(c1.fine[k], c0.fine[k]), k
        assert torch.equal(b1.fine[k], a1.fine[k])               # the renderer on its own samples
    assert torch.isfinite(a1.fine.rgb).all() and l1 == l0        # and the same number of kernel launches is claimed
