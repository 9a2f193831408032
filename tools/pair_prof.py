"""Cycle-counter breakdown of the cluster MLP kernels (development tool).

The counters are compiled in only in a profiling build of the library:

    DNR_CLUSTER_PROF=1 python tools/pair_prof.py        # rebuilds diner_b200/libdiner_b200.so with -DDNR_CLUSTER_PROF=1

(the production build reads no clocks; the next import without the variable rebuilds it).
"""
import os, sys
if os.environ.get("DNR_CLUSTER_PROF", "0") != "1":
    sys.exit("set DNR_CLUSTER_PROF=1: the cycle counters are not part of the production build")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from diner_b200 import NeRFRendererDGS, field as F_
from tools.synthetic import make_scene

scene = make_scene(os.environ.get("DNR_WORKLOAD", "facescape"), seed=0)
if os.environ.get("DNR_KERNEL"):
    F_.FORWARD_KERNEL = os.environ["DNR_KERNEL"]
wl = scene.workload
m = scene.model.to_device("cuda")
r = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=True)
rays = scene.rays.cuda()
with torch.no_grad():
    r(m, rays)
    F_.PAIR_PROF = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
    r(m, rays)
torch.cuda.synchronize()
p = F_.PAIR_PROF.view(148, 16).double().cpu()
p = p[p[:, 4] > 0]            # CTAs that ran (the quad kernel may leave SMs idle)
print('CTAs reporting:', p.shape[0]); lead = p[p[:, 0] > 0]
names = ["mma_total", "mma_wait_ring", "mma_wait_own", "mma_wait_peer", "epi_total", "epi_wait_accf", "-", "-", "loop_EPI_X(4/tile)", "loop_EPI_T(5/tile)", "loop_EPI_PARK(1)", "loop_EPI_HEAD(1)", "epi_merge", "wait_peer_full", "ring_lat", "pfull_lat"]
MMA_SIDE = {0, 1, 2, 3, 13, 14, 15}           # counters of the MMA-issuing thread (even CTA of each pair only)
for i, n in enumerate(names):
    rows = lead if i in MMA_SIDE else p
    ref = lead[:, 0].mean() if i in MMA_SIDE else p[:, 4].mean()
    print(f"{n:16s} mean {rows[:, i].mean():14.0f} cyc  ({100 * rows[:, i].mean() / max(ref, 1):5.1f}% of "
          f"{'mma_total' if i in MMA_SIDE else 'epi_total'})  min {rows[:, i].min():.0f} max {rows[:, i].max():.0f}")
