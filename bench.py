"""Benchmark of the DINER per-ray rendering hot path (BASELINE.json metric: render rays/s + Msamples/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload facescape|dtu|multiface|...] [--impl reference]

A step is one full-frame eval render of the workload's target view through the drop-in
`diner_b200.NeRFRendererDGS.forward` (sampler -> projection/gather -> ResnetFC MLP -> compositing) on one GPU.
With N > 1 (launched by torch.distributed.run, one rank per GPU) every rank renders its own frame of a camera
sweep around the same scene (the reference's create_cam_sweep loop, src/models/diner.py:178-199): target rays
shard across ranks with no data-path collective, so scaling is weak and `value` is all ranks' rays / max time.

Numbers on the JSON line
  value      rays/s, inputs resident in HBM, CUDA-event time per step (L2 flushed between steps), max over ranks
  e2e        the same through the same call with HOST (pinned) rays in and HOST rgb/depth out, copies timed
  roofline   the tcgen05 MLP GEMM kernel: algorithmic FLOPs / CUDA-event time of its launches vs measured bf16 peak
  cpu_baseline  the CPU oracle (torch-CPU restatement of the reference renderer, bit-pinned to the reference's
             outputs) timed on this box's host cores on a bounded sample (rank 0, N=1 only)
`--impl reference` times that CPU implementation alone (the reference is pure PyTorch; its CPU path is this).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def mlp_flops_per_sample(nv: int) -> float:
    """forward FLOPs of ResnetFC per sample point."""
    return 2.0 * (nv * (55 * 512 + 9 * 512 * 512) + 4 * 512 * 512 + 512 * 4)


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=p["hbm_gbs"], bf16_tflops=p["bf16_tflops"],
                    bf16_tflops_sustained=p.get("bf16_tflops_sustained", p["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if t0 is not None and not (t0 <= ts <= t1 + 0.2):
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except Exception:
                continue
            for nme, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def cpu_oracle_rate(workload: str, n_rays: int, repeats: int = 1):
    """rays/s of the CPU oracle (reference algorithm in torch-CPU ops) on `n_rays` rays of the workload."""
    from diner_b200.synthetic import make_scene
    from oracle import dgs_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    scene = make_scene(workload, seed=0)
    wl = scene.workload
    rays = scene.rays
    mid = rays.shape[1] // 2
    lo = max(0, mid - n_rays // 2)
    sub = rays[:, lo:lo + n_rays].contiguous()
    with torch.no_grad():
        O.render(scene.model, sub[:, :256].contiguous(), wl.n_samples, wl.n_candidates, wl.n_gaussian, wl.white_bkgd)
        times = []
        for _ in range(repeats):
            t0 = time.perf_counter()
            O.render(scene.model, sub, wl.n_samples, wl.n_candidates, wl.n_gaussian, wl.white_bkgd)
            times.append(time.perf_counter() - t0)
    n = sub.shape[0] * sub.shape[1]
    return n / min(times), n, torch.get_num_threads(), times


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (torch-CPU oracle, all host threads)."""
    if rank != 0:
        return
    from diner_b200.synthetic import WORKLOADS
    wl = WORKLOADS[args.workload]
    chunk = 4096      # the reference's own ray_batch_size (src/models/diner.py:57)
    rates = []
    t_begin = time.time()
    for i in range(args.warmup + args.steps):
        rate, n, threads, _ = cpu_oracle_rate(args.workload, chunk)
        if i >= args.warmup:
            rates.append(rate)
        if time.time() - t_begin > 180 and rates:      # keep the whole run within a few minutes on slow hosts
            break
    if not rates:
        rates = [rate]
    value = statistics.mean(rates)
    line = dict(impl="reference", metric="render_rays_per_s", value=value, unit="rays/s", n_gpus=args.gpus,
                steps=len(rates), warmup=args.warmup, ms_per_step=1e3 * chunk / value, higher_is_better=True,
                scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=args.workload + "_eval", rays_per_gpu_per_step=wl.tgt_H * wl.tgt_W * wl.SB,
                            samples_per_ray=wl.n_samples, n_candidates=wl.n_candidates, n_gaussian=wl.n_gaussian,
                            n_views=wl.NV, source_hw=[wl.H, wl.W],
                            sample="bounded: one 4096-ray chunk (the reference's ray_batch_size) of the frame per step; "
                                   "rays/s is per ray, so it compares with the full-frame figure"),
                msamples_per_s=value * wl.n_samples / 1e6,
                cpu_baseline=dict(value=value, unit="rays/s", cores=threads, kind="port",
                                  sample=f"{chunk} rays of the {args.workload} frame per step, torch-CPU oracle"),
                e2e=dict(value=value, unit="rays/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="facescape")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-rays", type=int, default=4096)
    ap.add_argument("--train", action="store_true", help="time forward+backward of the workload's ray batch instead")
    ap.add_argument("--shard-frame", action="store_true",
                    help="N > 1: all ranks render ONE frame (contiguous ray slices, all-gather of rgb/depth): strong scaling")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    args.warmup = max(args.warmup, 3)

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from types import SimpleNamespace
    from diner_b200 import NeRFRendererDGS, _lib
    from diner_b200.sharding import render_sharded
    from diner_b200.synthetic import make_scene, target_rays

    scene = make_scene(args.workload, seed=0)
    wl = scene.workload
    model = scene.model.to_device(dev)
    # every rank renders its own frame of a camera sweep around the scene
    shard = args.shard_frame and world > 1 and not args.train
    rays_cpu = target_rays(scene, azimuth_deg=2.0 + (0.0 if shard else 6.0 * rank), elevation_deg=8.0)
    n_rays = rays_cpu.shape[0] * rays_cpu.shape[1]
    renderer = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
    rays_dev = rays_cpu.to(dev)
    rays_host = rays_cpu.pin_memory()
    rays_stage = torch.empty_like(rays_dev)
    rgb_host = torch.empty(rays_cpu.shape[0], rays_cpu.shape[1], 3).pin_memory()
    depth_host = torch.empty(rays_cpu.shape[0], rays_cpu.shape[1]).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2

    enc_bucket = None
    if args.train:
        model.encoder.latent = model.encoder.latent.detach().requires_grad_(True)
        if world > 1:
            # DDP of the reference all-reduces every parameter gradient (configs/train_diner_facescape.yaml:87,
            # 11.67 M fp32 = 46.7 MB).  The MLP's 3.45 M gradients come from our backward; the
            # encoder trunk is stock PyTorch and not part of this bench, so a bucket of its size stands in for it.
            from diner_b200.sharding import allreduce_gradients
            n_mlp = sum(p.numel() for p in model.mlp_fine.parameters())
            enc_bucket = torch.zeros(max(0, 11_670_000 - n_mlp), device=dev)

    def train_allreduce():
        if world > 1:
            allreduce_gradients(list(model.mlp_fine.parameters()))
            dist.all_reduce(enc_bucket)

    def step_device():
        if args.train:
            out = renderer(model, rays_dev)
            loss = out.fine.rgb.square().mean()
            loss.backward()
            train_allreduce()
            return out
        if shard:        # one frame for all ranks: diner_b200.sharding splits the rays and all-gathers rgb / depth (NCCL)
            rgb, depth = render_sharded(renderer, model, rays_dev)
            return SimpleNamespace(fine=SimpleNamespace(rgb=rgb, depth=depth))
        with torch.no_grad():
            return renderer(model, rays_dev)

    def step_e2e():
        rays_stage.copy_(rays_host, non_blocking=True)
        if shard:
            rgb, depth = render_sharded(renderer, model, rays_stage)
            rgb_host.copy_(rgb, non_blocking=True)
            depth_host.copy_(depth, non_blocking=True)
            return None
        if args.train:
            out = renderer(model, rays_stage)
            out.fine.rgb.square().mean().backward()
            train_allreduce()
        else:
            with torch.no_grad():
                out = renderer(model, rays_stage)
        rgb_host.copy_(out.fine.rgb, non_blocking=True)
        depth_host.copy_(out.fine.depth, non_blocking=True)
        return out

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        evs = []
        n0 = _lib.launch_count
        t0 = time.time()
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t1 = time.time()
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        launches = _lib.launch_count - n0
        if world > 1:
            t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        return total_ms, launches, (t0, t1)

    clocks = ClockSampler(local_rank)
    clocks.start()
    total_ms, launches, window = timed(step_device, args.steps, args.warmup)
    clk = clocks.stop(*window)
    e2e_ms, _, _ = timed(step_e2e, args.steps, 2)

    # ---- per-kernel CUDA-event timing (separate instrumented steps; not part of the numbers above) ----
    _lib.profile = []
    prof_steps = 3
    for _ in range(prof_steps):
        flush.zero_()
        step_device()
    torch.cuda.synchronize()
    per = {}
    for name, e0, e1, tag in _lib.profile:
        if tag == "overlapped":              # runs on a second stream beside another kernel: not additive
            name += " (overlapped)"
        d = per.setdefault(name, [0.0, 0])
        d[0] += e0.elapsed_time(e1); d[1] += 1
    _lib.profile = None
    breakdown = {k: dict(ms_per_step=v[0] / prof_steps, launches_per_step=v[1] // prof_steps) for k, v in per.items()}

    peaks = load_peaks()
    n_local = n_rays // world if shard else n_rays          # rays this rank renders per step
    n_samples = n_local * wl.n_samples
    gemm_flops = 2.0 * (wl.NV * (55 * 512 + 9 * 512 * 512) + 4 * 512 * 512) * n_samples * (3.0 if args.train else 1.0)
    # every C-ABI call whose kernel runs the MLP's tcgen05 GEMMs (inference: one fused launch per chunk)
    gemm_names = ("dnr_mlp_quad_fwd", "dnr_mlp_pair_fwd", "dnr_mlp_fused_fwd", "dnr_mlp_gemm", "dnr_mlp_wgrad")
    kernel_name = ("mlp_quad_fwd_kernel (tcgen05 cta_group::2, 4-CTA clusters, whole ResnetFC per launch; "
                   "mlp_pair_fwd_kernel on the spare SMs, overlapped)" if "dnr_mlp_quad_fwd" in breakdown
                   else "mlp_pair_fwd_kernel (tcgen05, CTA pair, whole ResnetFC per launch)" if "dnr_mlp_pair_fwd" in breakdown
                   else "mlp_fused_fwd_kernel (tcgen05)" if "dnr_mlp_fused_fwd" in breakdown
                   else "mlp_gemm_kernel + mlp_wgrad_kernel (tcgen05)")
    gemm_ms = sum(breakdown[k]["ms_per_step"] for k in gemm_names if k in breakdown)
    gemm_launches = sum(breakdown[k]["launches_per_step"] for k in gemm_names if k in breakdown)
    achieved = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(args.workload)
    roofline = dict(bound="tensor", kernel=kernel_name, achieved=achieved,
                    peak=peaks["bf16_tflops_sustained"], unit="TFLOP/s",
                    frac=achieved / peaks["bf16_tflops_sustained"], traffic=traffic,
                    peak_source=f"{peaks['source']} sustained bf16 (kernel timed inside a long step)",
                    launches_per_step=gemm_launches, kernel_ms_per_step=gemm_ms,
                    algorithmic_flops_per_step=gemm_flops)

    # the other stages against the memory roofline (SURVEY.md 8d algorithmic bytes; HBM level and L2 level)
    def _gbs(nbytes, key):
        ms = breakdown.get(key, {}).get("ms_per_step", 0.0)
        return nbytes / (ms * 1e-3) / 1e9 if ms > 0 else None
    rows = n_samples * wl.NV
    stage_rooflines = dict(
        encode=dict(kernel="encode_fwd_kernel", hbm_gbs=_gbs(rows * 1152.0, "dnr_encode_fwd"),
                    l2_gbs=_gbs(rows * (4100.0 + 1152.0), "dnr_encode_fwd"), hbm_peak_gbs=peaks["hbm_gbs"],
                    note="HBM: 1152 B written per view row (operands of the MLP); L2: + 4 bilinear taps x 1 KB read "
                         "from the L2-resident bf16 latent; the L2-level figure is what the L2 -> SM fabric (12-13 TB/s chip-wide, DESIGN.md) bounds"),
        composite=dict(kernel="composite_fwd3_kernel", hbm_gbs=_gbs(n_local * 1008.0, "dnr_composite_fwd"),
                       hbm_peak_gbs=peaks["hbm_gbs"], note="1008 B per ray; ~10 us launches, latency-bound at this size"),
        sampler=dict(kernel="sampler_kernel", hbm_gbs=_gbs(n_local * 192.0, "dnr_sampler_fwd"),
                     l2_gbs=_gbs(n_local * wl.NV * wl.n_candidates * 20.0, "dnr_sampler_fwd"),
                     note="not memory-bound: NV*C candidate-view evaluations per ray on cache-resident maps"))

    if rank == 0:
        ms_per_step = total_ms / args.steps
        frames = 1 if shard else world
        value = n_rays * frames / (ms_per_step * 1e-3)
        e2e_value = n_rays * frames / (e2e_ms / args.steps * 1e-3)
        line = dict(metric="render_rays_per_s", value=value, unit="rays/s", n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=ms_per_step, higher_is_better=True,
                    scaling="strong" if shard else "weak", vs_baseline=None, dtype="bf16", data="synthetic",
                    config=dict(workload=args.workload + ("_train_fwd_bwd" if args.train else "_eval"),
                                rays_per_gpu_per_step=n_rays // world if shard else n_rays, samples_per_ray=wl.n_samples,
                                n_candidates=wl.n_candidates, n_gaussian=wl.n_gaussian, n_views=wl.NV,
                                source_hw=[wl.H, wl.W], l2="flushed between timed steps (256 MiB write)",
                                parallelism=(f"one frame, rays sharded over {world} ranks, all-gather of rgb/depth" if shard
                                             else f"rays sharded, dp{world}") + (
                                    ", NCCL all-reduce of 46.7 MB gradients per step (MLP grads + encoder-sized bucket)"
                                    if args.train and world > 1 else "")),
                    msamples_per_s=value * wl.n_samples / 1e6,
                    e2e=dict(value=e2e_value, unit="rays/s", h2d_bytes_per_step=rays_host.numel() * 4,
                             d2h_bytes_per_step=(rgb_host.numel() + depth_host.numel()) * 4,
                             ms_per_step=e2e_ms / args.steps),
                    gpu_launches=launches, clocks=clk, roofline=roofline, stage_rooflines=stage_rooflines,
                    breakdown_ms=breakdown)
        if world == 1 and not args.no_cpu_baseline:
            rate, n, threads, times = cpu_oracle_rate(args.workload, args.cpu_rays)
            line["cpu_baseline"] = dict(value=rate, unit="rays/s", cores=threads, kind="port",
                                        sample=f"{n} rays (one reference chunk) of the same frame, torch-CPU oracle, "
                                               f"{times[0]:.1f} s")
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
