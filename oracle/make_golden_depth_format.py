"""Generates tests/golden/depth_format.npz by running the UNMODIFIED reference dataset code on synthetic 16-bit PNGs
(build container only: needs /root/reference).

    python -m oracle.make_golden_depth_format

FacescapeDataSet.read_depth is a staticmethod and is called as is.  DTUDataSet.read_depth is called unbound with a
stand-in `self` carrying the two attributes it reads (downsample = 1, scale_factor).  The Multiface std law is inline
code of MultiFaceDataset.__getitem__ (multiface.py:305-311); its three lines are executed here as torch statements on
the output of the reference's own MultiFaceDataset.read_depth.
"""
import os
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402


def _planes(rng, H, W, hole_frac):
    d = rng.integers(9000, 16000, size=(H, W)).astype(np.uint16)
    d[rng.random((H, W)) < hole_frac] = 0
    return d


def main():
    from PIL import Image
    ref_shim._install_stubs()
    sys.path.insert(0, ref_shim.REF_ROOT)
    from src.data.facescape import FacescapeDataSet
    from src.data.dtu import DTUDataSet
    from src.data.multiface import MultiFaceDataset
    rng = np.random.default_rng(7)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        # ---- FaceScape: 2 views, 48 x 40, triptych + mesh ----
        H, W, N = 48, 40, 2
        tri = np.zeros((N, H, 3 * W), np.uint16)
        mesh = np.zeros((N, H, W), np.uint16)
        for n in range(N):
            tri[n, :, :W] = _planes(rng, H, W, 0.2)
            tri[n, :, W:2 * W] = _planes(rng, H, W, 0.3)
            conf = rng.integers(0, 10001, size=(H, W)).astype(np.uint16)       # confidence 0 .. 1 in 1e-4 units
            conf[rng.random((H, W)) < 0.1] = 0
            tri[n, :, 2 * W:] = conf
            mesh[n] = _planes(rng, H, W, 0.4)
        tri[0, 0, W] = 65535; tri[0, 0, 2 * W] = 65535; mesh[0, 0, 0] = 65535      # extreme values
        out["fs_tri"], out["fs_mesh"] = tri, mesh
        ds = types.SimpleNamespace(conf2std=FacescapeDataSet._getconf2std(None))
        for dt in ("original", "mesh", "merge"):
            preds, confs = [], []
            for n in range(N):
                p, mp = os.path.join(tmp, f"t{n}.png"), os.path.join(tmp, f"m{n}.png")
                Image.fromarray(tri[n]).save(p); Image.fromarray(mesh[n]).save(mp)
                a, b = FacescapeDataSet.read_depth(p, mp, depth_type=dt)
                preds.append(a); confs.append(b)
            out[f"fs_{dt}_depth"] = torch.stack(preds).numpy()
            out[f"fs_{dt}_conf"] = torch.stack(confs).numpy()
            out[f"fs_{dt}_std"] = ds.conf2std(torch.stack(confs)).numpy()              # facescape.py:266
        # ---- DTU: 512 x 640 (asserted by the reference), PNG branch ----
        scale_factor = 0.7 / 872.0
        dself = types.SimpleNamespace(downsample=1, scale_factor=scale_factor)
        dd = _planes(rng, 512, 640, 0.25); dc = rng.integers(0, 10001, size=(512, 640)).astype(np.uint16)
        p, pc = os.path.join(tmp, "d.png"), os.path.join(tmp, "d_conf.png")
        Image.fromarray(dd).save(p); Image.fromarray(dc).save(pc)
        depth, _ = DTUDataSet.read_depth(dself, p)
        conf, _ = DTUDataSet.read_depth(dself, pc)
        std = DTUDataSet._getconf2std(None)(torch.stack([conf]))
        # the decoding is per pixel: a 40 x 64 crop of inputs and outputs pins it and keeps the fixture small
        out["dtu_depth_u16"], out["dtu_conf_u16"] = dd[None, :40, :64].copy(), dc[None, :40, :64].copy()
        out["dtu_scale_factor"] = np.float64(scale_factor)
        out["dtu_depth"], out["dtu_std"] = depth[None, :, :40, :64].numpy().copy(), std[:, :, :40, :64].numpy().copy()
        # ---- Multiface ----
        md = _planes(rng, 32, 24, 0.3); mc = rng.integers(0, 20001, size=(32, 24)).astype(np.uint16)   # conf > 1.04 clips
        p, pc = os.path.join(tmp, "mf.png"), os.path.join(tmp, "mf_c.png")
        Image.fromarray(md).save(p); Image.fromarray(mc).save(pc)
        src_depth = MultiFaceDataset.read_depth(p)
        std_const = torch.ones_like(src_depth) * 1e-3                                    # multiface.py:307
        std_const[src_depth == 0] = 0                                                    # :311
        std_conf = MultiFaceDataset.read_depth(pc)
        std_conf = (-1.582e-2 * std_conf + 1.649e-2).clip(min=0)                         # :310
        std_conf[src_depth == 0] = 0
        out["mf_depth_u16"], out["mf_conf_u16"] = md[None], mc[None]
        out["mf_depth"], out["mf_std_const"], out["mf_std_conf"] = src_depth[None].numpy(), std_const[None].numpy(), std_conf[None].numpy()
    path = os.path.join(ROOT, "tests", "golden", "depth_format.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items() if hasattr(v, "shape")})


if __name__ == "__main__":
    main()
