"""Development tool: throughput of the drop-in renderer when the CALLER chunks the frame the way the reference does
(DINER.predict_imgs_from_batch splits the rays into ray_batch_size = 4096 pieces, src/models/diner.py:57,86-92), against
one full-frame call.  CUDA events, L2 flushed between frames."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from diner_b200 import NeRFRendererDGS
from tools.synthetic import make_scene

scene = make_scene("facescape", seed=0)
wl = scene.workload
dev = torch.device("cuda", 0)
model = scene.model.to_device(dev)
rays = scene.rays.to(dev)
r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
out = {}
for chunk in (65536, 16384, 4096, 1024):
    def frame():
        with torch.no_grad():
            return [r(model, c) for c in rays.split(chunk, dim=1)]
    for _ in range(3):
        frame()
    torch.cuda.synchronize()
    ts, cpu = [], []
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(); frame(); e1.record()
        cpu.append(time.perf_counter() - t0)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sum(ts) / len(ts)
    out[str(chunk)] = dict(ms_per_frame=ms, rays_per_s=rays.shape[1] / ms * 1e3, host_ms_per_frame=1e3 * sum(cpu) / len(cpu),
                           calls=len(rays.split(chunk, dim=1)))
print(json.dumps(out, indent=1))
