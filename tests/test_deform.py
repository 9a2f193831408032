"""Nearest-vertex deformation (first piece): numpy restatement known answers (CPU) and the CUDA kernel
through the C ABI against it (`-m gpu`).  Parity with pytorch3d itself is unpinned (not installed in the build image)."""
import numpy as np
import pytest
import torch

from oracle import deform_oracle as D


def test_oracle_known_answers():
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [1, 0, 0]], np.float32)            # vertex 3 duplicates vertex 1
    p = np.array([[0.1, 0, 0], [0.9, 0.1, 0], [0, 1.6, 0.2], [0.5, 0, 0]], np.float32)
    idx, d2 = D.nearest_vertex(p, v)
    assert idx.tolist() == [0, 1, 2, 0]                                                # ties: the lowest index (0.5: v0 vs v1; v1 vs v3)
    np.testing.assert_allclose(d2, [0.01, 0.02, 0.2, 0.25], rtol=1e-6)
    off = np.arange(12, dtype=np.float32).reshape(1, 4, 3)
    out = D.deform_points(p[None], v[None], off)
    np.testing.assert_array_equal(out[0], p + off[0][idx])


def test_host_layer_rejects_cpu_and_bad_shapes():
    from diner_b200.deform import deform_points
    p, v = torch.zeros(1, 5, 3), torch.zeros(1, 4, 3)
    with pytest.raises(ValueError, match="CUDA"):
        deform_points(p, v, v)


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diner_b200 import _lib
    _lib.load()
    return torch.device("cuda", 0)


@pytest.mark.gpu
@pytest.mark.parametrize("SB,B,V", [(1, 1000, 37), (2, 4097, 1024), (1, 300, 2500), (3, 1, 1), (1, 0, 5)])
def test_gpu_deform_matches_oracle(SB, B, V):
    dev = _cuda()
    from diner_b200.deform import deform_points, nearest_vertex
    rng = np.random.default_rng(SB * 1000 + V)
    p = rng.normal(size=(SB, B, 3)).astype(np.float32)
    v = rng.normal(size=(SB, V, 3)).astype(np.float32)
    off = (0.1 * rng.normal(size=(SB, V, 3))).astype(np.float32)
    t = lambda a: torch.from_numpy(a).to(dev)
    out, idx, d2 = nearest_vertex(t(p), t(v), t(off))
    assert out.shape == (SB, B, 3) and idx.shape == (SB, B)
    if B == 0:
        return
    idx_c, out_c, d2_c = idx.cpu().numpy(), out.cpu().numpy(), d2.cpu().numpy()
    for sb in range(SB):
        want_idx, want_d2 = D.nearest_vertex(p[sb], v[sb])
        same = idx_c[sb] == want_idx
        assert same.mean() > 0.999
        # a different vertex only where the two candidates are tied to fp32 round-off
        np.testing.assert_allclose(d2_c[sb][~same], want_d2[~same], rtol=2e-6)
        np.testing.assert_allclose(d2_c[sb], want_d2, rtol=2e-6)
        np.testing.assert_array_equal(out_c[sb], p[sb] + off[sb][idx_c[sb]])
    # the autograd wrapper: d points = g, d offsets = scatter-add of g by nearest vertex
    pt, ot = t(p).requires_grad_(True), t(off).requires_grad_(True)
    o2 = deform_points(pt, t(v), ot)
    g = torch.randn_like(o2)
    o2.backward(g)
    assert torch.equal(o2.detach(), out) and torch.equal(pt.grad, g)
    want = torch.zeros_like(ot)
    want.scatter_add_(1, idx.long().unsqueeze(-1).expand(-1, -1, 3), g)
    torch.testing.assert_close(ot.grad, want)


@pytest.mark.gpu
def test_gpu_deform_sample_points_of_a_frame():
    """FaceScape-frame size: 65 536 rays x 40 samples against 5 023 vertices (FLAME-sized mesh): every point's vertex is a
    true nearest neighbour (checked with torch.cdist on a strided subset), offsets of zero leave the points untouched."""
    dev = _cuda()
    from diner_b200.deform import nearest_vertex
    g = torch.Generator(device="cpu").manual_seed(0)
    pts = torch.randn(1, 65536 * 40, 3, generator=g).to(dev) * 0.15
    vts = torch.randn(1, 5023, 3, generator=g).to(dev) * 0.12
    out, idx, d2 = nearest_vertex(pts, vts)
    assert torch.equal(out, pts)
    sub = slice(0, None, 997)
    ref = torch.cdist(pts[0, sub].double(), vts[0].double()).min(dim=1).values ** 2
    torch.testing.assert_close(d2[0, sub].double(), ref, rtol=1e-4, atol=1e-9)
    assert int(idx.min()) >= 0 and int(idx.max()) < 5023
