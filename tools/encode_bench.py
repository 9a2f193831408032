"""Time dnr_encode_fwd alone on a workload's frame (CUDA events, L2 flushed), for ncu captures of the gather kernel:

    python tools/encode_bench.py [workload] [max_rays] [reps]
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from diner_b200 import _lib, sample_rays
from diner_b200.field import _encode_args
from diner_b200.scene import get_scene
from tools.synthetic import make_scene

workload = sys.argv[1] if len(sys.argv) > 1 else "facescape"
max_rays = int(sys.argv[2]) if len(sys.argv) > 2 else 0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
dev = torch.device("cuda", 0)
scene = make_scene(workload, seed=0, calibrate=False)
wl = scene.workload
m = scene.model.to_device(dev)
rays = scene.rays.to(dev)
if max_rays and rays.shape[1] > max_rays:
    mid = rays.shape[1] // 2
    rays = rays[:, mid - max_rays // 2: mid + max_rays // 2].contiguous()
sc = get_scene(m)
K = wl.n_samples
z = sample_rays(sc, rays, K, wl.n_candidates, wl.n_gaussian, seed=3)
n = rays.shape[0] * rays.shape[1]
mv = _lib.load().dnr_view_row_tiles(n * K, sc.NV)
lat = torch.empty(mv * 128 * 512, device=dev, dtype=torch.bfloat16)
code = torch.empty(mv * 128 * 64, device=dev, dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
p, st = _lib.ptr, _lib.stream_ptr()
r2 = rays.view(-1, 8)
times = []
for i in range(reps + 2):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.call("dnr_encode_fwd", p(r2), p(z), p(sc.poses), p(sc.focal), p(sc.cxy), p(sc.depth), p(sc.latent_nhwc),
              p(sc.freqs), p(sc.phases), *_encode_args(sc, rays.shape[1], K, 0, n), p(lat), p(code), st)
    e1.record()
    torch.cuda.synchronize()
    if i >= 2:
        times.append(e0.elapsed_time(e1))
rows = n * K * sc.NV
ms = min(times)
# footprint statistics: how often consecutive samples of a (ray, view) share a bilinear footprint
print(json.dumps(dict(workload=workload, rays=n, view_rows=rows, ms=ms, rows_per_us=rows / ms / 1e3,
                      hbm_write_gbs=rows * 1152 / ms / 1e6, latent_mb=sc.latent_nhwc.numel() * 2 / 1e6)))
