"""Generate tests/golden/producers.npz by running the UNMODIFIED reference input producers (build container only).

    python -m oracle.make_golden_producers

gen_rays (src/util/cam_geometry.py:36-79) and depth2normal (src/util/depth2normal.py:6-87) are imported from
/root/reference through oracle/ref_shim.py and run under torch's CPU backend on seeded inputs; inputs and outputs
are stored.  Test infrastructure.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from tools.synthetic import make_scene  # noqa: E402
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "producers.npz")


def main():
    ref = ref_shim.load()
    from src.util.depth2normal import depth2normal as ref_d2n          # noqa: E402  (reference, unmodified)
    data = {}
    # case A: the synthetic "tiny" scene's own cameras and depth maps (ellipsoid with background holes)
    scene = make_scene("tiny", seed=0)
    m = scene.model
    E = m.poses[0].clone()                                                 # (NV,4,4)
    NV = E.shape[0]
    K = torch.zeros(NV, 3, 3)
    K[:, 0, 0], K[:, 1, 1] = m.focal[0, :, 0], m.focal[0, :, 1]
    K[:, 0, 2], K[:, 1, 2] = m.c[0, :, 0], m.c[0, :, 1]
    K[:, 2, 2] = 1
    depth = m.encoder.depths[0].clone()                                    # (NV,1,H,W)
    H, W = depth.shape[-2:]
    near, far = torch.tensor([1.0, 0.9][:NV]), torch.tensor([2.5, 2.75][:NV])
    data.update(A_E=E.numpy(), A_K=K.numpy(), A_depth=depth.numpy(), A_near=near.numpy(), A_far=far.numpy(),
                A_rays=ref.gen_rays(E, K, W, H, near, far).numpy(), A_normals=ref_d2n(depth, K).numpy())
    # case B: random rigid cameras, a smooth random surface with random holes, ragged size, and a principal point
    # that makes (x + 0.5 - cx) == 0 on one pixel column (the reference's `x == 0` hole test fires there too)
    g = torch.Generator().manual_seed(7)
    Nb, Hb, Wb = 3, 23, 37
    A = torch.randn(Nb, 3, 3, generator=g)
    Q, _ = torch.linalg.qr(A)
    Q = Q * torch.sign(torch.linalg.det(Q)).view(-1, 1, 1)
    Eb = torch.eye(4).repeat(Nb, 1, 1)
    Eb[:, :3, :3] = Q
    Eb[:, :3, 3] = torch.randn(Nb, 3, generator=g)
    Kb = torch.zeros(Nb, 3, 3)
    Kb[:, 0, 0] = 40 + 10 * torch.rand(Nb, generator=g)
    Kb[:, 1, 1] = 45 + 10 * torch.rand(Nb, generator=g)
    Kb[:, 0, 2] = torch.tensor([10.5, 18.25, 20.0])
    Kb[:, 1, 2] = torch.tensor([11.0, 12.5, 9.75])
    Kb[:, 2, 2] = 1
    yy, xx = torch.meshgrid(torch.arange(Hb, dtype=torch.float32), torch.arange(Wb, dtype=torch.float32), indexing="ij")
    db = 1.5 + 0.2 * torch.sin(xx / 5)[None] * torch.cos(yy / 4)[None] + 0.05 * torch.rand(Nb, Hb, Wb, generator=g)
    db = db * (torch.rand(Nb, Hb, Wb, generator=g) > 0.15)                 # 15 % holes
    db[:, 5:9, 12:20] = 0                                                  # a block hole
    db = db.unsqueeze(1).contiguous()
    nb, fb = torch.rand(Nb, generator=g) + 0.5, torch.rand(Nb, generator=g) + 2.0
    data.update(B_E=Eb.numpy(), B_K=Kb.numpy(), B_depth=db.numpy(), B_near=nb.numpy(), B_far=fb.numpy(),
                B_rays=ref.gen_rays(Eb, Kb, Wb, Hb, nb, fb).numpy(), B_normals=ref_d2n(db, Kb).numpy())
    # case C: the first two views of B (N = 2: torch.cross takes the last axis, as intended)
    data.update(C_E=Eb[:2].numpy(), C_K=Kb[:2].numpy(), C_depth=db[:2].numpy(), C_near=nb[:2].numpy(), C_far=fb[:2].numpy(),
                C_rays=ref.gen_rays(Eb[:2], Kb[:2], Wb, Hb, nb[:2], fb[:2]).numpy(),
                C_normals=ref_d2n(db[:2].contiguous(), Kb[:2]).numpy())
    np.savez_compressed(OUT, **data)
    for k in ("A_normals", "B_normals", "C_normals"):
        n = data[k]
        print(k, n.shape, "nan:", int(np.isnan(n).sum()), "nonzero:", int((n != 0).sum()))
    print("wrote", OUT, f"{os.path.getsize(OUT) / 1e3:.1f} KB")


if __name__ == "__main__":
    main()
