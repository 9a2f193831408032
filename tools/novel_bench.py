"""Time the NOVEL renderer variant (diner_b200.novel.NeRFRendererDGS) on a full FaceScape-shape frame against a
FLAME-sized target mesh (5 023 vertices): frame time, per-C-ABI-call breakdown, and the DINER renderer on the same frame.

    python tools/novel_bench.py [n_vertices] [steps]
    python tools/novel_bench.py [n_vertices] [steps] train     # the training step (SB = 2 x 4 096 rays, fwd + loss + bwd
                                                               # w.r.t. MLP, gen_latent, encoder.latent) beside DINER's
    python tools/novel_bench.py [n_vertices] [steps] pe | pe_train      # the same for the NOVEL_PE field
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from diner_b200 import NeRFRendererDGS as DinerRenderer, _lib
from diner_b200.novel import NeRFRendererDGS
from tools.synthetic import add_novel_inputs, add_novel_pe_inputs, make_scene, novel_to_device

V = int(sys.argv[1]) if len(sys.argv) > 1 else 5023
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
MODE = sys.argv[3] if len(sys.argv) > 3 else "eval"
TRAIN, PE = MODE in ("train", "pe_train"), MODE in ("pe", "pe_train")
TAG = "NOVEL_PE" if PE else "NOVEL"


def train_main():
    scene = add_novel_inputs(make_scene("facescape_train", seed=0), n_vertices=V, seed=0, gen_hw=(192, 192))
    diner_scene = make_scene("facescape_train", seed=0)
    diner_scene.model.to_device(dev)
    if PE:
        scene = add_novel_pe_inputs(scene, seed=0)
    novel_to_device(scene, dev)
    wl, m = scene.workload, scene.model
    md = diner_scene.model
    md.encoder.latent = md.encoder.latent.detach().requires_grad_(True)
    rays = scene.rays.to(dev)
    m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
    m.gen_latent = m.gen_latent.detach().requires_grad_(True)
    r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    rd = DinerRenderer(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def clear():
        for q in (*m.parameters(), *md.parameters(), m.gen_latent, m.encoder.latent, md.encoder.latent):
            q.grad = None

    def novel_step():
        clear()
        out = r(m, rays, scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets)
        out.fine.rgb.square().mean().backward()

    def diner_step():
        clear()
        rd(md, rays).fine.rgb.square().mean().backward()

    def timed(fn):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return min(ts), sum(ts) / len(ts)

    t_novel, t_diner = timed(novel_step), timed(diner_step)
    _lib.profile = []
    novel_step()
    torch.cuda.synchronize()
    per = {}
    for name, e0, e1, tag in _lib.profile:
        d = per.setdefault(name, [0.0, 0])
        d[0] += e0.elapsed_time(e1); d[1] += 1
    _lib.profile = None
    n = rays.shape[0] * rays.shape[1]
    rows = n * wl.n_samples * wl.NV
    bwd = per.get("dnr_encode_novel_bwd", [0.0, 0])
    print(json.dumps(dict(workload=f"facescape_train {TAG} variant (fwd + loss + bwd)", rays=n, n_vertices=V,
                          ms_best=t_novel[0], ms_mean=t_novel[1], rays_per_s=n / (t_novel[1] * 1e-3),
                          diner_ms_mean=t_diner[1],
                          breakdown_ms={k: dict(ms=v[0], launches=v[1]) for k, v in per.items()},
                          # fp32 gradient rows read once (2 KB per view row); the reductions land in L2
                          encode_novel_bwd_GBps=rows * 2048 / (bwd[0] * 1e-3) / 1e9 if bwd[0] else None,
                          d_gen_latent_nonzero=int((m.gen_latent.grad != 0).sum()),
                          finite=bool(torch.isfinite(m.gen_latent.grad).all() and torch.isfinite(m.encoder.latent.grad).all()))))


if TRAIN:
    train_main()
    sys.exit(0)
scene = add_novel_inputs(make_scene("facescape", seed=0), n_vertices=V, seed=0, gen_hw=(192, 192))
diner_model = make_scene("facescape", seed=0).model.to_device(dev)
if PE:
    scene = add_novel_pe_inputs(scene, seed=0)
novel_to_device(scene, dev)
wl = scene.workload
rays = scene.rays.to(dev)
args = (scene.model, rays, scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets)
r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
rd = DinerRenderer(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sum(ts) / len(ts)


with torch.no_grad():
    t_novel = timed(lambda: r(*args))
    t_diner = timed(lambda: rd(diner_model, rays))
    _lib.profile = []
    r(*args)
    torch.cuda.synchronize()
    per = {}
    for name, e0, e1, tag in _lib.profile:
        d = per.setdefault(name + (" (overlapped)" if tag == "overlapped" else ""), [0.0, 0])
        d[0] += e0.elapsed_time(e1); d[1] += 1
    _lib.profile = None
    out = r(*args)
n = rays.shape[0] * rays.shape[1]
deform_ms = per.get("dnr_novel_deform", [0, 0])[0]
queries = n * (wl.n_candidates + wl.n_samples)
print(json.dumps(dict(workload=f"facescape_eval {TAG} variant", rays=n, n_vertices=V, ms_best=t_novel[0], ms_mean=t_novel[1],
                      rays_per_s=n / (t_novel[1] * 1e-3), diner_ms_mean=t_diner[1],
                      breakdown_ms={k: dict(ms=v[0], launches=v[1]) for k, v in per.items()},
                      deform_queries=queries, deform_queries_per_s=queries / (deform_ms * 1e-3) if deform_ms else None,
                      brute_force_pairs_equivalent_per_s=queries * V / (deform_ms * 1e-3) if deform_ms else None,
                      finite=bool(torch.isfinite(out.fine.rgb).all()))))
