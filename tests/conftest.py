import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def load_golden(name):
    """Golden vectors written by oracle/make_golden.py (outputs of the unmodified reference)."""
    from tools.synthetic import make_scene
    g = dict(np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"), allow_pickle=False))
    scene = make_scene(str(g["workload"]), seed=int(g["seed"]))
    rays = scene.rays
    if int(g["ray_start"]) >= 0:
        rays = rays[:, int(g["ray_start"]):int(g["ray_stop"])].contiguous()
    noise = dict(u_coarse=torch.from_numpy(g["u_coarse"]),
                 n_gauss=torch.from_numpy(g["n_gauss"]),
                 u_fill=torch.from_numpy(g["u_fill"]))
    return g, scene, rays, noise


@pytest.fixture(scope="session")
def golden_cache():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_golden(name)
        return cache[name]
    return get
