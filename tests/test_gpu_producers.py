"""GPU parity tests of the input producers (gen_rays, depth2normal) through the C ABI -- run with `-m gpu`.

Checkers: outputs of the unmodified reference (tests/golden/producers.npz, oracle/make_golden_producers.py), the numpy
restatement (oracle/producers_oracle.py), and size-independent properties at the BASELINE.json frame sizes.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "producers.npz")


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()
    return torch.device("cuda", 0)


@pytest.mark.parametrize("case,W,H", [("A", 80, 64), ("B", 37, 23), ("C", 37, 23)])
def test_gen_rays_matches_reference(case, W, H):
    dev = _cuda()
    from diner_b200 import gen_rays
    g = np.load(GOLD)
    t = lambda k: torch.from_numpy(g[case + k]).to(dev)
    rays = gen_rays(t("_E"), t("_K"), W, H, t("_near"), t("_far")).cpu().numpy()
    ref = g[case + "_rays"]
    assert rays.shape == ref.shape
    assert np.array_equal(rays[..., 6:], ref[..., 6:])
    np.testing.assert_allclose(rays[..., :6], ref[..., :6], rtol=0, atol=1e-6)       # 3-term sums: order is the BLAS's
    # every ray of a camera starts at the same centre, and the centre projects to itself: E [c; 1] = 0
    c = torch.from_numpy(rays[:, 0, 0, :3])
    E = torch.from_numpy(g[case + "_E"])
    assert np.array_equal(rays[..., :3], np.broadcast_to(rays[:, :1, :1, :3], rays[..., :3].shape))
    assert float((torch.einsum("bij,bj->bi", E[:, :3, :3], c) + E[:, :3, 3]).abs().max()) < 1e-5


@pytest.mark.parametrize("case", ["A", "B", "C"])
def test_depth2normal_matches_reference(case):
    dev = _cuda()
    from diner_b200 import depth2normal
    from oracle import producers_oracle as P
    g = np.load(GOLD)
    d, K = torch.from_numpy(g[case + "_depth"]).to(dev), torch.from_numpy(g[case + "_K"]).to(dev)
    n = depth2normal(d, K).cpu().numpy()
    ref = g[case + "_normals"]
    assert n.shape == ref.shape
    assert np.array_equal(np.isnan(n), np.isnan(ref))                       # 0/0 exactly where the reference has it
    assert np.array_equal(n == 0, ref == 0)
    np.testing.assert_allclose(np.nan_to_num(n), np.nan_to_num(ref), rtol=0, atol=1e-5)
    # against the restatement every operation is one rounding on both sides: bit-exact
    assert np.array_equal(n, P.depth2normal(g[case + "_depth"], g[case + "_K"]), equal_nan=True)


def test_producers_full_frame_properties_and_renderer_round_trip():
    """FaceScape frame sizes: unit directions, pixel-centre round trip through the projection, normals of the
    synthetic scene's depth maps face the camera, and a frame rendered from device-generated rays equals the one
    rendered from the host-generated rays of the scene generator."""
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS, depth2normal, gen_rays
    from tools.synthetic import make_scene
    scene = make_scene("facescape", seed=0)
    wl = scene.workload
    m = scene.model.to_device(dev)
    E, K = scene.target_extrinsics.to(dev), scene.target_intrinsics.to(dev)
    near = torch.full((E.shape[0],), wl.near, device=dev)
    far = torch.full((E.shape[0],), wl.far, device=dev)
    rays = gen_rays(E, K, wl.tgt_W, wl.tgt_H, near, far)
    assert rays.shape == (E.shape[0], wl.tgt_H, wl.tgt_W, 8)
    d = rays[..., 3:6]
    assert float((d.norm(dim=-1) - 1).abs().max()) < 1e-6
    # a point one unit along each ray projects back onto its pixel centre
    p = rays[..., :3] + d
    pc = torch.einsum("bij,bhwj->bhwi", E[:, :3, :3], p) + E[:, None, None, :3, 3]
    u = pc[..., 0] / pc[..., 2] * K[:, 0, 0].view(-1, 1, 1) + K[:, 0, 2].view(-1, 1, 1)
    v = pc[..., 1] / pc[..., 2] * K[:, 1, 1].view(-1, 1, 1) + K[:, 1, 2].view(-1, 1, 1)
    xs = torch.arange(wl.tgt_W, device=dev) + 0.5
    ys = torch.arange(wl.tgt_H, device=dev) + 0.5
    assert float((u - xs.view(1, 1, -1)).abs().max()) < 1e-3 and float((v - ys.view(1, -1, 1)).abs().max()) < 1e-3
    host = scene.rays.to(dev).view(rays.shape)
    torch.testing.assert_close(rays, host, rtol=0, atol=1e-6)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    sub = slice(0, 8192)
    with torch.no_grad():
        a = r(m, rays.view(1, -1, 8)[:, sub].contiguous(), noise=None)
    assert torch.isfinite(a.fine.rgb).all()
    # normals of the scene's own depth maps
    depth = m.encoder.depths.flatten(0, 1)                                    # (SB*NV,1,H,W)
    Ks = torch.zeros(depth.shape[0], 3, 3, device=dev)
    Ks[:, 0, 0], Ks[:, 1, 1] = m.focal.flatten(0, 1)[:, 0], m.focal.flatten(0, 1)[:, 1]
    Ks[:, 0, 2], Ks[:, 1, 2] = m.c.flatten(0, 1)[:, 0], m.c.flatten(0, 1)[:, 1]
    Ks[:, 2, 2] = 1
    n = depth2normal(depth, Ks)
    fg = depth[:, 0] != 0
    assert bool((n.permute(0, 2, 3, 1)[~fg] == 0).all())
    nf = n.permute(0, 2, 3, 1)[fg]
    ok = torch.isfinite(nf).all(-1)
    assert ok.float().mean().item() > 0.999
    assert float((nf[ok].norm(dim=-1) - 1).abs().max()) < 1e-5
    assert (nf[ok][:, 2] < 0).float().mean().item() > 0.99                    # facing the camera (z towards -view)
    # idempotent on repeated calls / deterministic
    assert torch.equal(torch.nan_to_num(n), torch.nan_to_num(depth2normal(depth, Ks)))


# ---------------------------------------------------------------------------------------------------------
# caller-side ray handling: selected-pixel rays, patches, full-frame / sweep rendering
# ---------------------------------------------------------------------------------------------------------

def test_gen_rays_at_equals_gather_of_full_frame_bit_exact():
    dev = _cuda()
    from diner_b200 import gen_rays, gen_rays_at
    from oracle import producers_oracle as P
    g = np.load(GOLD)
    W, H = 80, 64
    t = lambda k: torch.from_numpy(g["A" + k]).to(dev)
    E, K, near, far = t("_E"), t("_K"), t("_near"), t("_far")
    B = E.shape[0]
    full = gen_rays(E, K, W, H, near, far).view(B, H * W, 8)
    gen = torch.Generator().manual_seed(5)
    for n in (1, 4096, 7):
        pix = torch.randint(-H * W, H * W, (B, n), generator=gen)            # negative indices count from the end
        got = gen_rays_at(E, K, W, H, near, far, pix.to(dev))
        want = full[torch.arange(B, device=dev)[:, None], pix.to(dev)]
        assert torch.equal(got, want)
        ref = P.rays_at(g["A_E"], g["A_K"], W, H, g["A_near"], g["A_far"], pix.numpy())
        assert np.array_equal(got.cpu().numpy()[..., 6:], ref[..., 6:])
        np.testing.assert_allclose(got.cpu().numpy(), ref, rtol=0, atol=1e-6)
    assert gen_rays_at(E, K, W, H, near, far, torch.zeros(B, 0, dtype=torch.int64, device=dev)).shape == (B, 0, 8)
    with pytest.raises(IndexError):
        gen_rays_at(E, K, W, H, near, far, torch.full((B, 3), H * W, dtype=torch.int64, device=dev))
    with pytest.raises(IndexError):
        gen_rays_at(E, K, W, H, near, far, torch.zeros(B, 3, device=dev))     # float indices
    with pytest.raises(ValueError):
        gen_rays_at(E, K, W, H, near, far, torch.zeros(B + 1, 3, dtype=torch.int64, device=dev))


@pytest.mark.parametrize("patch", [64, 5, 1])
def test_patch_pixels_match_restatement(patch):
    dev = _cuda()
    from diner_b200.callers import patch_pixels, patch_pixels_at, training_rays
    from oracle import producers_oracle as P
    SB, H, W = 3, 96, 112
    torch.manual_seed(patch)
    alpha = (torch.rand(SB, 1, H, W) > 0.4).float()
    gen = torch.Generator(device=dev).manual_seed(11)
    idx, centres = patch_pixels(alpha.to(dev), patch, generator=gen)
    assert idx.shape == (SB, patch * patch) and centres.shape == (SB, 2)
    flat = (centres[:, 0] + centres[:, 1] * W).cpu().numpy()
    assert np.array_equal(idx.cpu().numpy(), P.patch_pixel_indices(flat, W, patch))
    assert np.array_equal(patch_pixels_at(torch.from_numpy(flat), W, patch).numpy(), idx.cpu().numpy())
    elig = P.eligible_patch_centres(alpha.numpy(), patch)
    for s in range(SB):                                                     # drawn from the eligible set only
        assert elig[s, int(centres[s, 1]), int(centres[s, 0])] > 0
    assert int(idx.min()) >= 0 and int(idx.max()) < H * W
    x, y = idx % W, idx // W                                                # a full window inside the image
    assert ((x.max(1).values - x.min(1).values) == patch - 1).all() and ((y.max(1).values - y.min(1).values) == patch - 1).all()
    # colours and rays of those pixels
    g = np.load(GOLD)
    E = torch.from_numpy(g["A_E"]).to(dev)[:1].expand(SB, 4, 4).contiguous()
    K = torch.from_numpy(g["A_K"]).to(dev)[:1].expand(SB, 3, 3).contiguous()
    rgb = torch.rand(SB, 3, H, W, device=dev)
    rays, gt = training_rays(E, K, W, H, 1.0, 2.5, idx, target_rgb=rgb, check_bounds=True)
    assert rays.shape == (SB, patch * patch, 8) and gt.shape == (SB, patch * patch, 3)
    s, i = SB - 1, patch * patch - 1
    assert torch.equal(gt[s, i], rgb[s, :, int(y[s, i]), int(x[s, i])])
    assert float(rays[..., 6].min()) == 1.0 and float(rays[..., 7].max()) == 2.5


def test_predict_imgs_and_sweep_match_direct_renderer_calls():
    dev = _cuda()
    from diner_b200 import NeRFRendererDGS, gen_rays
    from diner_b200.callers import predict_imgs, render_sweep
    from tools.synthetic import make_scene
    scene = make_scene("tiny", seed=0)
    wl = scene.workload
    model = scene.model.to_device(dev)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    H, W = wl.tgt_H, wl.tgt_W
    E, K = scene.target_extrinsics.to(dev), scene.target_intrinsics.to(dev)
    # noise-free comparison: the renderer draws from the torch CUDA generator, so seed before each call
    def run(fn):
        torch.manual_seed(3); torch.cuda.manual_seed(3)
        return fn()
    rgb, depth = run(lambda: predict_imgs(r, model, E, K, W, H, wl.near, wl.far, return_depth=True))
    assert rgb.shape == (1, 3, H, W) and depth.shape == (1, 1, H, W)
    rays = gen_rays(E, K, W, H, torch.tensor([wl.near], device=dev), torch.tensor([wl.far], device=dev)).view(1, H * W, 8)
    with torch.no_grad():
        out = run(lambda: r.forward(model=model, rays=rays))
    assert torch.equal(rgb, out.fine.rgb.view(1, H, W, 3).permute(0, 3, 1, 2))
    assert torch.equal(depth, out.fine.depth.view(1, H, W, 1).permute(0, 3, 1, 2))
    assert torch.isfinite(rgb).all() and float(rgb.min()) >= 0 and float(rgb.max()) <= 1 + 1e-5
    # split like the reference loop: same shapes, finite (noise is consumed in a different order)
    rgb_s = predict_imgs(r, model, E, K, W, H, wl.near, wl.far, ray_batch_size=100)
    assert rgb_s.shape == rgb.shape and torch.isfinite(rgb_s).all()
    # sweep: frame 0 of a two-frame sweep is the single-frame render when the frames go one per call
    E2 = torch.cat([E, E])
    sw_rgb, sw_depth = run(lambda: render_sweep(r, model, E2, K, W, H, wl.near, wl.far, frames_per_call=1))
    assert sw_rgb.shape == (2, 3, H, W) and sw_depth.shape == (2, 1, H, W)
    assert torch.equal(sw_rgb[0], rgb[0])
    both = render_sweep(r, model, E2, K, W, H, wl.near, wl.far, frames_per_call=2)[0]
    assert both.shape == (2, 3, H, W) and torch.isfinite(both).all()
