"""Stage-by-stage GPU diagnostics against the CPU oracle (development tool, not product).

    python tools/gpu_diag.py [stage ...]      stages: composite gemm wgrad sampler encode render grads

Each stage runs in its own subprocess with a timeout so that a hung kernel cannot block the rest.
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

STAGES = ["composite", "gemm", "wgrad", "sampler", "encode", "render", "fieldgrad", "grads"]


def _stats(name, got, want, tol=None):
    import torch
    got, want = got.detach().float().cpu(), want.detach().float().cpu()
    err = (got - want).abs()
    den = want.abs().max().item() + 1e-12
    print(f"  {name}: max_abs_err={err.max().item():.3e} mean_abs_err={err.mean().item():.3e} "
          f"max|want|={den:.3e} rel={err.max().item() / den:.3e} nan={int(torch.isnan(got).sum())}", flush=True)


def stage_composite():
    import torch
    from diner_b200.renderer import _CompositeFn
    from oracle import dgs_oracle as O
    torch.manual_seed(0)
    n, K = 1000, 40
    z = torch.sort(torch.rand(n, K) * 1.5 + 1.0, -1).values
    rays = torch.zeros(n, 8); rays[:, 6] = 1.0; rays[:, 7] = 2.5
    rs = torch.rand(n, K, 4); rs[..., 3] = torch.relu(torch.randn(n, K) * 30)
    for white in (True, False):
        a = rs.clone().requires_grad_(True)
        w, rgb, dep = O.alpha_composite(a.view(1, n, K, 4), z.view(1, n, K), rays.view(1, n, 8), white)
        gr, gd, gw = torch.randn(1, n, 3), torch.randn(1, n), torch.randn(1, n, K)
        (rgb * gr).sum().add((dep * gd).sum()).add((w * gw).sum()).backward()
        b = rs.clone().cuda().requires_grad_(True)
        rgb2, dep2, w2 = _CompositeFn.apply(b.view(n * K, 4), z.cuda(), rays.cuda(), white)
        ((rgb2 * gr[0].cuda()).sum() + (dep2 * gd[0].cuda()).sum() + (w2 * gw[0].cuda()).sum()).backward()
        print(f" white={white}")
        _stats("rgb", rgb2, rgb[0]); _stats("depth", dep2, dep[0]); _stats("weights", w2, w[0])
        _stats("d_rgbsigma", b.grad, a.grad)


def _pack_rows(x, rows_pad):
    import torch
    from diner_b200 import _lib
    out = torch.empty(rows_pad * x.shape[1], device="cuda", dtype=torch.bfloat16)
    _lib.call("dnr_pack_rows", _lib.ptr(x.contiguous()), x.shape[0], x.shape[1], rows_pad, _lib.ptr(out), _lib.stream_ptr())
    return out


def _unpack_rows(pk, rows, F):
    import torch
    from diner_b200 import _lib
    out = torch.empty(rows, F, device="cuda")
    _lib.call("dnr_unpack_rows", _lib.ptr(pk), rows, F, _lib.ptr(out), _lib.stream_ptr())
    return out


def _xr_to_rows(xr, rows, N):
    import torch
    from diner_b200 import _lib
    out = torch.empty(rows, N, device="cuda")
    _lib.call("dnr_xr_to_rows", _lib.ptr(xr), rows, N, _lib.ptr(out), _lib.stream_ptr())
    return out


def _rows_to_xr(x, rows_pad):
    import torch
    from diner_b200 import _lib
    out = torch.empty(rows_pad * x.shape[1], device="cuda")
    _lib.call("dnr_rows_to_xr", _lib.ptr(x.contiguous()), x.shape[0], x.shape[1], rows_pad, _lib.ptr(out), _lib.stream_ptr())
    return out


def _pack_w(w, transpose=False):
    import torch
    from diner_b200 import _lib
    rows, cols = w.shape
    n, k = (cols, rows) if transpose else (rows, cols)
    n_pad, k_pad = (n + 255) // 256 * 256, (k + 63) // 64 * 64
    out = torch.empty(n_pad * k_pad, device="cuda", dtype=torch.bfloat16)
    _lib.call("dnr_pack_weight", _lib.ptr(w.contiguous()), rows, cols, cols, int(transpose), n_pad, k_pad, _lib.ptr(out),
              _lib.stream_ptr())
    return out


def stage_gemm():
    import torch
    from diner_b200.field import _gemm
    torch.manual_seed(0)
    bfr = lambda t: t.bfloat16().float()
    for rows, K in ((128, 64), (300, 512), (5000, 512)):
        rp = (rows + 127) // 128 * 128
        A = torch.randn(rows, K, device="cuda"); W = torch.randn(512, K, device="cuda") / K ** 0.5
        bias = torch.randn(512, device="cuda")
        ref = bfr(A) @ bfr(W).T + bias
        xr = torch.empty(rp * 512, device="cuda"); act = torch.empty(rp * 512, device="cuda", dtype=torch.bfloat16)
        _gemm(_pack_rows(A, rp), _pack_w(W), K, rp // 128, bias=bias, res_out=xr, act_out=act, relu=1)
        torch.cuda.synchronize()
        print(f" rows={rows} K={K}")
        _stats("res_out", _xr_to_rows(xr, rows, 512), ref)
        _stats("act_out(relu,bf16)", _unpack_rows(act, rows, 512), torch.relu(ref))
    # two segments + residual + mask
    rows, rp = 1000, 1024
    A0 = torch.randn(rows, 64, device="cuda"); A1 = torch.randn(rows, 512, device="cuda")
    W0 = torch.randn(512, 64, device="cuda") / 8; W1 = torch.randn(512, 512, device="cuda") / 22
    R = torch.randn(rows, 512, device="cuda"); M = torch.randn(rows, 512, device="cuda")
    ref = (bfr(A0) @ bfr(W0).T + bfr(A1) @ bfr(W1).T) * (bfr(M) > 0) + R
    xr = torch.empty(rp * 512, device="cuda"); act = torch.empty(rp * 512, device="cuda", dtype=torch.bfloat16)
    _gemm(_pack_rows(A0, rp), _pack_w(W0), 64, rp // 128, A1=_pack_rows(A1, rp), B1=_pack_w(W1), K1=512,
          mask=_pack_rows(M, rp), res_in=_rows_to_xr(R, rp), res_out=xr, act_out=act)
    torch.cuda.synchronize()
    print(" two segments + mask + residual")
    _stats("res_out", _xr_to_rows(xr, rows, 512), ref)
    _stats("act_out", _unpack_rows(act, rows, 512), ref)
    # transposed weights (dgrad)
    G = torch.randn(rows, 512, device="cuda")
    ref = bfr(G) @ bfr(W1)
    _gemm(_pack_rows(G, rp), _pack_w(W1, True), 512, rp // 128, res_out=xr)
    torch.cuda.synchronize()
    print(" dgrad (W^T)")
    _stats("res_out", _xr_to_rows(xr, rows, 512), ref)
    # view mean
    from diner_b200 import _lib
    for NV in (2, 3, 4):
        ns = 777
        mv = _lib.load().dnr_view_row_tiles(ns, NV)
        spw = 32 // NV
        rowsv = mv * 128
        Av = torch.zeros(rowsv, 512, device="cuda")
        idx = torch.arange(ns * NV, device="cuda")
        s_, v_ = idx // NV, idx % NV
        row = (s_ // spw) * 32 + (s_ % spw) * NV + v_
        vals = torch.randn(ns * NV, 512, device="cuda")
        Av[row] = vals
        ref = (bfr(vals) @ bfr(W1).T + bias).view(ns, NV, 512).mean(1)
        ms = (ns + 127) // 128
        xm = torch.zeros(ms * 128 * 512, device="cuda"); am = torch.zeros(ms * 128 * 512, device="cuda", dtype=torch.bfloat16)
        _gemm(_pack_rows(Av, rowsv), _pack_w(W1), 512, mv, bias=bias, res_out=xm, act_out=am, relu=1, mean_nv=NV, n_samples=ns)
        torch.cuda.synchronize()
        print(f" view mean NV={NV}")
        _stats("res_out", _xr_to_rows(xm, ns, 512), ref)
        _stats("act_out", _unpack_rows(am, ns, 512), torch.relu(ref))
    # speed
    rows = 128 * 2048
    A = _pack_rows(torch.randn(rows, 512, device="cuda"), rows); Wp = _pack_w(W1)
    act = torch.empty(rows * 512, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        _gemm(A, Wp, 512, rows // 128, bias=bias, act_out=act, relu=1)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(10):
        _gemm(A, Wp, 512, rows // 128, bias=bias, act_out=act, relu=1)
    e1.record(); torch.cuda.synchronize()
    ms_ = e0.elapsed_time(e1) / 10
    print(f" speed: {rows}x512x512 in {ms_:.3f} ms -> {2 * rows * 512 * 512 / ms_ / 1e9:.1f} TFLOP/s")


def stage_wgrad():
    import torch
    from diner_b200 import _lib
    torch.manual_seed(0)
    bfr = lambda t: t.bfloat16().float()
    for rows, Fx, kv in ((1000, 512, 512), (4000, 64, 55), (128 * 300, 512, 512)):
        rp = (rows + 127) // 128 * 128
        dY = torch.randn(rows, 512, device="cuda"); X = torch.randn(rows, Fx, device="cuda")
        dW = torch.zeros(512, kv, device="cuda"); db = torch.zeros(512, device="cuda")
        dYp, Xp = _pack_rows(dY, rp), _pack_rows(X, rp)
        _lib.call("dnr_mlp_wgrad", _lib.ptr(dYp), 512, _lib.ptr(Xp), Fx, rp // 128,
                  _lib.ptr(dW), kv, 512, kv, _lib.ptr(db), _lib.stream_ptr())
        torch.cuda.synchronize()
        print(f" rows={rows} Fx={Fx}")
        _stats("dW", dW, (bfr(dY).T @ bfr(X))[:, :kv]); _stats("db", db, bfr(dY).sum(0))


def _golden(name):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import load_golden
    return load_golden(name)


def _full_noise(g, wl, n_rays, seed=5):
    """Unpacked per-ray draws: u_coarse from the golden file, fresh gaussian / fill draws for every ray and slot."""
    import torch
    gen = torch.Generator().manual_seed(seed)
    return dict(u_coarse=torch.from_numpy(g["u_coarse"]),
                n_gauss_full=torch.randn(n_rays, wl.n_gaussian, generator=gen),
                u_fill_full=torch.rand(n_rays, wl.n_samples, generator=gen))


def _oracle_sampler(model, rays, wl, noise):
    import torch
    from oracle import dgs_oracle as O
    with torch.no_grad():
        z0, _, dbg = O.sample_depthguided(model, rays, wl.n_samples, wl.n_candidates, wl.n_gaussian, noise=dict(noise))
        z, _ = O.fill_up_uniform(z0.clone(), rays, None, noise["u_fill_full"])
    K, G = wl.n_samples, wl.n_gaussian
    M = K - G
    order = dbg.order.view(-1, K)[:, :M]
    p = dbg.p.view(-1, wl.n_candidates)
    psel = torch.gather(p, 1, order)
    idx = torch.where(psel > 0, order, torch.full_like(order, -1))
    rm = dbg.ray_mask.view(-1).int()
    nm = (z0.view(-1, K) == 0).sum(-1).int()
    return dict(z=z.view(-1, K), p=p, idx=idx, rm=rm, nm=nm)


def stage_sampler():
    import torch
    from diner_b200 import sample_rays
    from diner_b200.scene import get_scene
    for name in ("tiny", "tiny3", "plumbing64"):
        g, scene, rays, _ = _golden(name)
        wl = scene.workload
        noise = _full_noise(g, wl, rays.shape[0] * rays.shape[1])
        oc = _oracle_sampler(scene.model, rays, wl, noise)                       # torch CPU ops
        scene.model.to_device("cuda")
        noise_d = {k: v.cuda() for k, v in noise.items()}
        og = _oracle_sampler(scene.model, rays.cuda(), wl, noise_d)             # torch CUDA ops
        sc = get_scene(scene.model)
        z, d = sample_rays(sc, rays.cuda(), wl.n_samples, wl.n_candidates, wl.n_gaussian, noise=noise_d, want_debug=True)
        torch.cuda.synchronize()
        print(f" [{name}] rays={oc['z'].shape[0]}")
        for tag, o in (("cpu-oracle", oc), ("cuda-oracle", og)):
            zc, zo = z.cpu(), o["z"].cpu()
            po, pg = o["p"].cpu(), d["p"].cpu()
            exact = (zc == zo).all(-1)
            close = torch.isclose(zc, zo, rtol=1e-5, atol=1e-6).all(-1)
            same_idx = (d["sel_idx"].cpu().long() == o["idx"].cpu()).all(-1)
            same_set = (torch.sort(d["sel_idx"].cpu().long(), -1).values == torch.sort(o["idx"].cpu(), -1).values).all(-1)
            pdiff = (po != pg)
            flips = ((po == 0) != (pg == 0))
            rel = ((po - pg).abs() / po.abs().clamp(min=1e-30))[pdiff & ~flips]
            print(f"   vs {tag}: z bit-exact rays={int(exact.sum())} close1e-5={int(close.sum())} idx exact(order)={int(same_idx.sum())} "
                  f"idx same set={int(same_set.sum())} ray_mask agree={int((d['ray_mask'].cpu() == o['rm'].cpu()).sum())} "
                  f"n_missing agree={int((d['n_missing'].cpu() == o['nm'].cpu()).sum())}")
            print(f"      p: differing candidates={int(pdiff.sum())} of {int((po != 0).sum())} nonzero; zero/nonzero flips={int(flips.sum())} "
                  f"max rel diff={float(rel.max()) if rel.numel() else 0:.3e} rays with p diff={int(pdiff.any(-1).sum())}", flush=True)
            bad = (~same_idx).nonzero().flatten()[:2]
            for r in bad.tolist():
                print("      idx mismatch ray", r, "want", o["idx"][r].tolist()[:8], "got", d["sel_idx"][r].tolist()[:8])
                c = pdiff[r].nonzero().flatten()[:4].tolist()
                print("         p diffs at", c, [float(po[r, i]) for i in c], [float(pg[r, i]) for i in c])
            badz = (~close).nonzero().flatten()[:2]
            for r in badz.tolist():
                dz = (zc[r] - zo[r]).abs()
                print("      z mismatch ray", r, "n bad slots", int((dz > 1e-5).sum()), "max", float(dz.max()))
        pdiff = (oc["p"] != og["p"].cpu())
        print(f"   cpu-oracle vs cuda-oracle: p differing={int(pdiff.sum())} flips={int(((oc['p'] == 0) != (og['p'].cpu() == 0)).sum())} "
              f"idx equal rays={int((oc['idx'] == og['idx'].cpu()).all(-1).sum())} "
              f"z exact rays={int((oc['z'] == og['z'].cpu()).all(-1).sum())} z close rays={int(torch.isclose(oc['z'], og['z'].cpu(), rtol=1e-5, atol=1e-6).all(-1).sum())}")
        # packed (reference-order) noise through the counting pass must reproduce the golden file where the masks agree
        _, _, _, pnoise = _golden(name)
        zg = sample_rays(sc, rays.cuda(), wl.n_samples, wl.n_candidates, wl.n_gaussian, noise={k: v.cuda() for k, v in pnoise.items()})
        gz = torch.from_numpy(g["z"]).view(-1, wl.n_samples)
        print(f"   packed noise vs golden: z bit-exact rays={int((zg.cpu() == gz).all(-1).sum())} close={int(torch.isclose(zg.cpu(), gz, rtol=1e-5, atol=1e-6).all(-1).sum())}")
        zp = sample_rays(sc, rays.cuda(), wl.n_samples, wl.n_candidates, wl.n_gaussian, seed=1)
        ok = bool((zp[:, 1:] >= zp[:, :-1]).all())
        print(f"   philox mode: sorted={ok} min={float(zp.min()):.3f} max={float(zp.max()):.3f}")


def stage_encode():
    import torch
    from diner_b200 import _lib
    from diner_b200.field import field_forward, get_packed
    from diner_b200.scene import get_scene
    from oracle import dgs_oracle as O
    for name in ("tiny", "tiny3"):
        g, scene, rays, noise = _golden(name)
        wl = scene.workload
        K, NV, SB = wl.n_samples, wl.NV, wl.SB
        z = torch.from_numpy(g["z"])
        NR = rays.shape[1]
        pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
        dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
        with torch.no_grad():
            zx = O.field_inputs(scene.model, pts, dirs)            # SB,NV,B,567
            want_out = O.field(scene.model, pts, dirs).reshape(-1, 4)
        scene.model.to_device("cuda")
        sc = get_scene(scene.model)
        pk = get_packed(scene.model.mlp_fine, True)
        out, sv = field_forward(sc, pk, rays.view(-1, 8).cuda(), z.view(-1, K).cuda(), NR, 0, save=True)
        torch.cuda.synchronize()
        ns = SB * NR * K
        lat = _unpack_rows(sv.lat, sv.mv * 128, 512).cpu()
        code = _unpack_rows(sv.code, sv.mv * 128, 64).cpu()
        spw = 32 // NV
        s_ = torch.arange(ns)
        print(f" [{name}]")
        for v in range(NV):
            row = (s_ // spw) * 32 + (s_ % spw) * NV + v
            sb = s_ // (NR * K)
            w = zx[sb, v, s_ % (NR * K)]
            _stats(f"latent v{v}", lat[row], w[:, :512]); _stats(f"code v{v}", code[row][:, :55], w[:, 512:])
        _stats("rgbsigma", out, want_out)
        _stats("rgb only", out[:, :3], want_out[:, :3])


def stage_render():
    import torch
    from diner_b200 import NeRFRendererDGS
    for name in ("tiny", "tiny3", "plumbing64"):
        g, scene, rays, noise = _golden(name)
        wl = scene.workload
        scene.model.to_device("cuda")
        r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        with torch.no_grad():
            out = r(scene.model, rays.cuda(), want_weights=True, noise=dict(z=torch.from_numpy(g["z"])))
        torch.cuda.synchronize()
        print(f" [{name}]")
        _stats("rgb", out.fine.rgb, torch.from_numpy(g["rgb"])); _stats("depth", out.fine.depth, torch.from_numpy(g["depth"]))
        _stats("weights", out.fine.weights, torch.from_numpy(g["weights"]))


def stage_fieldgrad():
    """MLP + gather backward in isolation: unit-scale random d(rgbsigma), compared with autograd through the oracle."""
    import torch
    from diner_b200.renderer import _FieldFn
    from diner_b200.scene import get_scene
    from oracle import dgs_oracle as O
    for name in ("tiny", "tiny3"):
        g, scene, rays, _ = _golden(name)
        wl = scene.workload
        K, SB = wl.n_samples, wl.SB
        NR = rays.shape[1]
        z = torch.from_numpy(g["z"])
        m = scene.model
        gen = torch.Generator().manual_seed(3)
        d_out = torch.randn(SB * NR * K, 4, generator=gen)
        if os.environ.get("DIAG_RGB_ONLY") == "1":
            d_out[:, 3] = 0      # no gradient through the (ill-conditioned) relu(sigma) channel
        lat = m.encoder.latent.detach().clone().requires_grad_(True)
        m.encoder.latent = lat
        pts = (rays[..., None, :3] + z.unsqueeze(-1) * rays[..., None, 3:6]).reshape(SB, NR * K, 3)
        dirs = rays[..., None, 3:6].expand(-1, -1, K, -1).reshape(SB, NR * K, 3)
        want = O.field(m, pts, dirs).reshape(-1, 4)
        (want * d_out).sum().backward()
        ref = {k: v.grad.clone() for k, v in m.mlp_fine.named_parameters()}
        ref_lat = lat.grad.clone()
        for v in m.mlp_fine.parameters():
            v.grad = None
        m.encoder.latent = lat.detach()
        m.to_device("cuda")
        m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
        sc = get_scene(m)
        mlp = m.mlp_fine
        names = [n for n, _ in mlp.named_parameters()]
        params = [q for _, q in mlp.named_parameters()]
        out = _FieldFn.apply(sc, mlp, names, rays.view(-1, 8).cuda(), z.view(-1, K).cuda(), NR, [(0, rays.shape[0] * rays.shape[1])], "bf16",
                             m.encoder.latent, *params)
        (out * d_out.cuda()).sum().backward()
        torch.cuda.synchronize()
        print(f" [{name}]")
        _stats("out", out, want)
        for k, v in mlp.named_parameters():
            e = (v.grad.cpu() - ref[k]).double().norm().item() / max(ref[k].double().norm().item(), 1e-30)
            print(f"  d {k}: rel_l2_err={e:.3e} |want|={ref[k].double().norm().item():.3e}")
        gl = m.encoder.latent.grad.cpu()
        print(f"  d latent: rel_l2_err={(gl - ref_lat).double().norm().item() / ref_lat.double().norm().item():.3e} |want|={ref_lat.double().norm().item():.3e}")


def stage_grads():
    import torch
    from diner_b200 import NeRFRendererDGS
    for name in ("tiny", "tiny3"):
        g, scene, rays, noise = _golden(name)
        wl = scene.workload
        m = scene.model
        m.to_device("cuda")
        m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
        r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        out = r(m, rays.cuda(), noise=dict(z=torch.from_numpy(g["z"])))
        gen = torch.Generator().manual_seed(11)
        w_rgb = torch.randn(out.fine.rgb.shape, generator=gen).cuda()
        w_dep = torch.randn(out.fine.depth.shape, generator=gen).cuda()
        loss = (out.fine.rgb * w_rgb).sum() + (out.fine.depth * w_dep).sum()
        loss.backward()
        torch.cuda.synchronize()
        print(f" [{name}] loss={loss.item():.6f} want={float(g['loss']):.6f}")
        glat = m.encoder.latent.grad
        _stats("glat(sample)", glat.flatten()[torch.from_numpy(g["glat_idx"]).cuda()], torch.from_numpy(g["glat_val"]))
        print(f"  |glat|={glat.double().norm().item():.5e} want={float(g['glat_norm']):.5e}")
        for k, v in m.mlp_fine.named_parameters():
            gn = v.grad.double().norm().item()
            print(f"  |d {k}|={gn:.4e} want={float(g['gradnorm.' + k]):.4e} ratio={gn / max(float(g['gradnorm.' + k]), 1e-30):.4f}")
            if f"grad.{k}" in g:
                _stats(f"d {k}", v.grad, torch.from_numpy(g[f"grad.{k}"]))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--run":
        t0 = time.time()
        globals()["stage_" + sys.argv[2]]()
        print(f"  ({time.time() - t0:.1f}s)")
        sys.exit(0)
    rc = 0
    for st in (sys.argv[1:] or STAGES):
        print(f"=== {st} ===", flush=True)
        try:
            r = subprocess.run([sys.executable, __file__, "--run", st], timeout=240)
            if r.returncode != 0:
                print(f"!!! stage {st} exited {r.returncode}", flush=True); rc = 1
        except subprocess.TimeoutExpired:
            print(f"!!! stage {st} TIMED OUT", flush=True); rc = 1
    sys.exit(rc)
