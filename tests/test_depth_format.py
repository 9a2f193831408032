"""On-disk depth format: oracle vs outputs of the unmodified reference dataset code (CPU), and the
CUDA decoder through the C ABI vs both (`-m gpu`), bit for bit -- the decoding is integer -> fp32 arithmetic with one
rounding per reference operation."""
import os

import numpy as np
import pytest
import torch

from oracle import depth_format_oracle as D

GOLD = os.path.join(os.path.dirname(__file__), "golden", "depth_format.npz")
FS_LAW = (-1.582e-2, 1.649e-2)
DTU_LAW = (-2.5679e-2, 3.2818e-2)


@pytest.mark.parametrize("depth_type", ["original", "mesh", "merge"])
def test_oracle_facescape_matches_reference(depth_type):
    g = np.load(GOLD)
    for n in range(g["fs_tri"].shape[0]):
        p, c = D.read_depth_facescape(g["fs_tri"][n], g["fs_mesh"][n], depth_type)
        assert p.dtype == np.float32 and c.dtype == np.float32
        assert np.array_equal(p, g[f"fs_{depth_type}_depth"][n])
        assert np.array_equal(c, g[f"fs_{depth_type}_conf"][n])
        assert np.array_equal(D.conf2std(c, *FS_LAW), g[f"fs_{depth_type}_std"][n])
    with pytest.raises(ValueError):
        D.read_depth_facescape(g["fs_tri"][0], g["fs_mesh"][0], "median")


def test_oracle_dtu_and_multiface_match_reference():
    g = np.load(GOLD)
    sf = float(g["dtu_scale_factor"])
    assert np.array_equal(D.read_depth_dtu_png(g["dtu_depth_u16"], sf)[:, None], g["dtu_depth"])
    conf = D.read_depth_dtu_png(g["dtu_conf_u16"], sf)
    assert np.array_equal(D.conf2std(conf, *DTU_LAW)[:, None], g["dtu_std"])
    md = g["mf_depth_u16"].astype(np.float32) * np.float32(1e-4)
    mc = g["mf_conf_u16"].astype(np.float32) * np.float32(1e-4)
    assert np.array_equal(md[:, None], g["mf_depth"])
    assert np.array_equal(D.multiface_std(md, None)[:, None], g["mf_std_const"])
    assert np.array_equal(D.multiface_std(md, mc)[:, None], g["mf_std_conf"])
    assert (g["mf_std_conf"] == 0).any() and (g["mf_std_conf"] > 0).any()          # the clip and the hole rule both fire


def test_host_layer_rejects_cpu_and_bad_inputs():
    from diner_b200 import depth_format as F
    tri = torch.zeros(1, 4, 12, dtype=torch.uint16)
    with pytest.raises(ValueError, match="CUDA"):
        F.decode_facescape(tri)
    with pytest.raises(ValueError, match="original, mesh or merge"):
        F.decode_facescape(tri, None, "median")


# ------------------------------------------------------------------------------------------------ GPU
def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()
    return torch.device("cuda", 0)


@pytest.mark.gpu
@pytest.mark.parametrize("depth_type", ["original", "mesh", "merge"])
def test_gpu_facescape_bit_exact(depth_type):
    dev = _cuda()
    from diner_b200 import depth_format as F
    g = np.load(GOLD)
    tri, mesh = torch.from_numpy(g["fs_tri"]).to(dev), torch.from_numpy(g["fs_mesh"]).to(dev)
    depth, std = F.decode_facescape(tri, mesh, depth_type)
    assert depth.shape == std.shape == (2, 1, 48, 40)
    assert np.array_equal(depth.cpu().numpy(), g[f"fs_{depth_type}_depth"])
    assert np.array_equal(std.cpu().numpy(), g[f"fs_{depth_type}_std"])
    p, c = F.read_depth(tri[1], mesh[1], depth_type)                               # the reference's own return values
    assert np.array_equal(p.cpu().numpy(), g[f"fs_{depth_type}_depth"][1])
    assert np.array_equal(c.cpu().numpy(), g[f"fs_{depth_type}_conf"][1])
    if depth_type != "original":
        with pytest.raises(ValueError):
            F.decode_facescape(tri, None, depth_type)
    # int16 storage with the same bits (older torch builds have no uint16 transfer)
    d2, s2 = F.decode_facescape(tri.view(torch.int16), mesh.view(torch.int16), depth_type)
    assert torch.equal(d2, depth) and torch.equal(s2, std)


@pytest.mark.gpu
def test_gpu_dtu_and_multiface_bit_exact():
    dev = _cuda()
    from diner_b200 import depth_format as F
    g = np.load(GOLD)
    t = lambda k: torch.from_numpy(g[k]).to(dev)
    depth, std = F.decode_dtu(t("dtu_depth_u16"), t("dtu_conf_u16"), float(g["dtu_scale_factor"]))
    assert np.array_equal(depth.cpu().numpy(), g["dtu_depth"])
    assert np.array_equal(std.cpu().numpy(), g["dtu_std"])
    depth, std = F.decode_multiface(t("mf_depth_u16"))
    assert np.array_equal(depth.cpu().numpy(), g["mf_depth"]) and np.array_equal(std.cpu().numpy(), g["mf_std_const"])
    depth, std = F.decode_multiface(t("mf_depth_u16"), t("mf_conf_u16"))
    assert np.array_equal(depth.cpu().numpy(), g["mf_depth"]) and np.array_equal(std.cpu().numpy(), g["mf_std_conf"])


@pytest.mark.gpu
def test_gpu_full_size_against_oracle_and_feeds_the_renderer_inputs():
    """FaceScape frame size (2 views, 256 x 256), random planes: kernel == numpy oracle bit for bit; empty batch."""
    dev = _cuda()
    from diner_b200 import depth_format as F
    rng = np.random.default_rng(0)
    N, H, W = 2, 256, 256
    tri = rng.integers(0, 65536, size=(N, H, 3 * W)).astype(np.uint16)
    mesh = rng.integers(0, 65536, size=(N, H, W)).astype(np.uint16)
    tri[rng.random(tri.shape) < 0.3] = 0
    mesh[rng.random(mesh.shape) < 0.5] = 0
    for dt in ("original", "mesh", "merge"):
        depth, std = F.decode_facescape(torch.from_numpy(tri).to(dev), torch.from_numpy(mesh).to(dev), dt)
        for n in range(N):
            p, c = D.read_depth_facescape(tri[n], mesh[n], dt)
            assert np.array_equal(depth[n].cpu().numpy(), p)
            assert np.array_equal(std[n].cpu().numpy(), D.conf2std(c, *FS_LAW))
    d0, s0 = F.decode_facescape(torch.zeros(0, H, 3 * W, dtype=torch.uint16, device=dev))
    assert d0.shape == (0, 1, H, W) and s0.shape == (0, 1, H, W)
