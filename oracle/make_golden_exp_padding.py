"""Generate tests/golden/exp_padding.npz from the UNMODIFIED reference (build container only).

    python -m oracle.make_golden_exp_padding

Runs the reference's `exponential_padding` and its `grid_sample(padding_mode="exponential")`
(src/util/torch_helpers.py:100-160, the body of SpatialEncoder.index_depth_std, image_encoder.py:153-180) on
random std maps whose BORDER pixels are non-zero -- the case the synthetic scenes never produce, because their
background (std = 0) surrounds the object -- with the production constants pad = 100, double_width = 12.
Test infrastructure.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_shim  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "exp_padding.npz")


def main():
    ref_shim.load()
    from src.util.torch_helpers import exponential_padding, grid_sample
    g = torch.Generator().manual_seed(21)
    out = {}
    for name, (N, H, W) in dict(a=(3, 37, 50), b=(2, 64, 80)).items():
        std = torch.rand(N, 1, H, W, generator=g) * 0.02 + 0.001
        std[:, :, H // 3: H // 2, W // 4: W // 2] = 0.0           # a hole (std = 0 must stay 0)
        std[0, :, 0, : W // 2] = 0.0                              # part of the border is background
        padded = exponential_padding(std, 100, 12)
        uv = torch.rand(N, 4000, 1, 2, generator=g) * 3.2 - 1.6   # inside, in the padded ring and beyond it
        taps = grid_sample(std, uv, mode="nearest", padding_mode="exponential", align_corners=False,
                           pad_size=100, pad_double_width=12, exp_padding_mode="zeros")
        out[f"{name}_std"] = std.numpy()
        out[f"{name}_padded"] = padded.numpy()
        out[f"{name}_uv"] = uv.numpy()
        out[f"{name}_taps"] = taps.numpy()
        print(name, tuple(std.shape), "->", tuple(padded.shape), "taps", tuple(taps.shape),
              "non-zero taps", int((taps != 0).sum()))
    np.savez_compressed(GOLDEN, **out)
    print("wrote", GOLDEN, f"{os.path.getsize(GOLDEN) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
