"""Stress the cluster kernels (development tool): random ray counts and view counts, the 4-CTA kernel (with the
pooled-row merge and the two-stream split) must reproduce the 2-CTA kernel bit for bit, run after run."""
import dataclasses, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from diner_b200 import field as F_
from diner_b200.scene import get_scene
from tools.synthetic import WORKLOADS, make_scene

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
dev = torch.device("cuda", 0)
gen = torch.Generator().manual_seed(0)
t0 = time.time()
runs = 0
for NV in (2, 3, 4, 1):
    wl = dataclasses.replace(WORKLOADS["tiny"], name=f"s{NV}", NV=NV, tgt_H=64, tgt_W=64)
    scene = make_scene(wl, seed=NV, calibrate=False)
    m = scene.model.to_device(dev)
    sc = get_scene(m)
    pk = F_.get_packed(m.mlp_fine, need_transposed=False)
    K = wl.n_samples
    t_nv = time.time()
    while time.time() - t_nv < budget / 4:
        n = int(torch.randint(1, 4096, (1,), generator=gen))
        rays = scene.rays[:, :n].contiguous().to(dev)
        u = torch.rand(1, n, K, generator=gen).sort(-1).values.to(dev)
        z = rays[..., 6:7] + (rays[..., 7:8] - rays[..., 6:7]) * u
        F_.FORWARD_KERNEL, F_.PAIR_SHARE = "pair", None
        a, _ = F_.field_forward(sc, pk, rays.view(-1, 8), z.view(-1, K), n, 0, save=False)
        F_.FORWARD_KERNEL = "quad"
        F_.PAIR_SHARE = [None, 0.0, 0.3, 0.5][runs % 4]
        for rep in range(3):
            b, _ = F_.field_forward(sc, pk, rays.view(-1, 8), z.view(-1, K), n, 0, save=False)
            torch.cuda.synchronize()
            assert torch.equal(a, b), (NV, n, F_.PAIR_SHARE, rep, float((a - b).abs().max()))
        runs += 1
    m.to_device("cpu")
print(f"stress ok: {runs} random launches x 3 repeats in {time.time() - t0:.0f} s")
