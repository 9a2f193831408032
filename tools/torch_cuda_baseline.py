"""Development tool (test infrastructure, not product): time the reference ALGORITHM in stock PyTorch CUDA ops on
the GPU box -- the restatement in oracle/dgs_oracle.py executed on CUDA tensors, chunked like the reference
(4096 rays per chunk, src/models/diner.py:57; eval_batch_size 100000).  This is the "reference on a GPU today"
figure of SURVEY.md 8(d); the CUDA path of this repo never touches it.

    python tools/torch_cuda_baseline.py [workload] [n_chunks]
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tools.synthetic import make_scene
from oracle import dgs_oracle as O

workload = sys.argv[1] if len(sys.argv) > 1 else "facescape"
n_chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
scene = make_scene(workload, seed=0)
wl = scene.workload
scene.model.to_device("cuda")
rays = scene.rays.cuda()
mid = rays.shape[1] // 2
chunks = [rays[:, mid + i * 4096: mid + (i + 1) * 4096].contiguous() for i in range(-(n_chunks // 2), n_chunks - n_chunks // 2)]
with torch.no_grad():
    O.render(scene.model, chunks[0], wl.n_samples, wl.n_candidates, wl.n_gaussian, wl.white_bkgd)      # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for c in chunks:
        O.render(scene.model, c, wl.n_samples, wl.n_candidates, wl.n_gaussian, wl.white_bkgd)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
n = sum(c.shape[0] * c.shape[1] for c in chunks)
print(json.dumps(dict(workload=workload, impl="reference algorithm, torch CUDA ops (fp32)", rays=n, seconds=dt,
                      rays_per_s=n / dt, peak_mem_gb=torch.cuda.max_memory_allocated() / 1e9)))
