"""Print the round-2 parity metrics (tests/parity_metrics.py) on a B200 without asserting -- used to choose and to
re-check the tolerances written into tests/test_gpu_parity_r2.py.

    python tools/parity_probe.py [golden] [shapes] [train] > gpurun_out/parity_probe.json
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import parity_metrics as PM  # noqa: E402
from conftest import load_golden  # noqa: E402


def main():
    what = sys.argv[1:] or ["golden", "shapes", "train"]
    dev = torch.device("cuda", 0)
    res = {}
    if "golden" in what:
        for name in ("tiny", "tiny3", "plumbing64"):
            g, scene, rays, _ = load_golden(name)
            for prec in ("bf16", "split"):
                out, m = PM.render_on_golden_z(scene, rays, g, dev, prec)
                res[f"render.{name}.{prec}"] = dict(
                    rgb=PM.error_metrics(out.fine.rgb, torch.from_numpy(g["rgb"])),
                    depth=PM.error_metrics(out.fine.depth, torch.from_numpy(g["depth"])),
                    weights=PM.error_metrics(out.fine.weights, torch.from_numpy(g["weights"])))
                m.to_device("cpu")
                res[f"grads.{name}.{prec}"] = PM.golden_gradient_metrics(scene, rays, g, dev, prec)
            print(json.dumps({k: v for k, v in res.items() if name in k}), flush=True)
    if "shapes" in what:
        for wl in ("facescape", "dtu", "multiface"):
            res[f"shape.{wl}"] = PM.config_shape_parity(wl, dev)
            print(json.dumps({f"shape.{wl}": res[f"shape.{wl}"]}), flush=True)
    if "train" in what:
        for prec in ("bf16", "split"):
            res[f"train.{prec}"] = PM.train_shape_gradient_metrics(dev, prec)
            print(json.dumps({f"train.{prec}": res[f"train.{prec}"]}), flush=True)


if __name__ == "__main__":
    main()
