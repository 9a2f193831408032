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
    from diner_b200.synthetic import make_scene
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
