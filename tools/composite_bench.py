"""Stand-alone bandwidth measurement of the compositing kernels through the C ABI (development tool).

Algorithmic bytes per ray as in SURVEY.md 8(d): forward 1008 B, backward 1776 B (with d_weights) at K = 40.  Inputs are
larger than L2 or L2 is flushed between launches; CUDA events on the launching stream.
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from diner_b200 import _lib

K = 40
dev = "cuda"
peak = 6544.0
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
p = _lib.ptr
out = {}
for n in (10923, 65536, 1 << 20):
    torch.manual_seed(0)
    z = torch.sort(torch.rand(n, K, device=dev) * 1.5 + 1.0, -1).values.contiguous()
    rays = torch.zeros(n, 8, device=dev); rays[:, 6] = 1.0; rays[:, 7] = 2.5
    rs = torch.rand(n, K, 4, device=dev); rs[..., 3] = torch.relu(torch.randn(n, K, device=dev) * 30)
    w = torch.empty(n, K, device=dev); rgb = torch.empty(n, 3, device=dev); dep = torch.empty(n, device=dev)
    g_rgb = torch.randn(n, 3, device=dev); g_d = torch.randn(n, device=dev); g_w = torch.randn(n, K, device=dev)
    d = torch.empty_like(rs)
    st = _lib.stream_ptr()
    fwd = lambda: _lib.call("dnr_composite_fwd", p(rs), p(z), p(rays), n, K, 1, p(w), p(rgb), p(dep), st)
    bwd = lambda: _lib.call("dnr_composite_bwd", p(rs), p(z), p(rays), p(g_rgb), p(g_d), p(g_w), n, K, 1, p(d), st)
    for name, fn, bytes_per_ray in (("fwd", fwd, 32 + K * 24 + 16), ("bwd", bwd, 16 + K * 28 + K * 16)):
        ts = []
        for it in range(8):
            if n * 1008 < (256 << 20):      # the 1 M-ray case streams 1 GB: larger than L2 by itself
                flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            if it >= 3:
                ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        gbs = n * bytes_per_ray / ms / 1e6
        out[f"{name}_{n}"] = dict(us=round(ms * 1e3, 2), gbs=round(gbs, 1), frac=round(gbs / peak, 3))
print(json.dumps(out, indent=1))
