"""Benchmark of the DINER per-ray rendering hot path (BASELINE.json metric: render rays/s + Msamples/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload facescape|dtu|multiface|...] [--impl reference]

Headline (the JSON line's metric / value / e2e / roofline): one full-frame eval render of the FaceScape-shape target
view (BASELINE.json config 3, the configuration the metric is quoted on) through the drop-in
`diner_b200.NeRFRendererDGS.forward` (sampler -> projection/gather -> ResnetFC MLP -> compositing).

N = 1   the frame on one GPU.  The same invocation also measures, under the key "workloads" (so that they carry the
        driver's clock record): the config-4 training step (forward + backward, SB = 2 x 4 096 rays), the DTU-shape frame
        (config 2), a 131 072-ray slice of the Multiface stress frame (config 5) and the reference ALGORITHM on stock
        torch CUDA ops on the same GPU.
N > 1   (torch.distributed.run, one rank per GPU) the partition north_star states: ONE frame's rays sharded over the
        ranks by diner_b200.sharding.render_sharded, NCCL all-gather of rgb / depth inside the timed region,
        "scaling": "strong"; the scene (latent + maps + cameras + weights) is broadcast from rank 0 once per scene
        (`scene_broadcast` reports its time).  "workloads" then holds the replica sweep (one frame per rank, weak) and
        the training step with the 46.7 MB gradient all-reduce.

Numbers on the JSON line
  value      rays/s, inputs resident in HBM, CUDA-event time per step (L2 flushed between steps), max over ranks
  e2e        the same through the same call with HOST (pinned) rays in and HOST rgb/depth out, copies timed
  roofline   the tcgen05 MLP kernel(s): algorithmic FLOPs / CUDA-event time of their launches vs measured bf16 peak
  cpu_baseline  the CPU oracle (torch-CPU restatement of the reference renderer, bit-pinned to the reference's
             outputs) timed on this box's host cores on a bounded sample (rank 0, N = 1 only)
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
    """forward FLOPs of ResnetFC per sample point (the 4-wide lin_out is left out: it runs as fp32
    FMAs inside the epilogue, not on the tensor pipe)."""
    return 2.0 * (nv * (55 * 512 + 9 * 512 * 512) + 4 * 512 * 512)


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


# ---------------------------------------------------------------------------------------------------------------
# CPU arms
# ---------------------------------------------------------------------------------------------------------------

class CpuOracle:
    """The reference algorithm in torch-CPU ops (oracle/dgs_oracle.py) on a bounded sample of a workload's frame; the
    synthetic scene is built ONCE."""

    def __init__(self, workload: str):
        from tools.synthetic import make_scene
        torch.set_num_threads(os.cpu_count() or 1)
        self.scene = make_scene(workload, seed=0)
        self.threads = torch.get_num_threads()

    def rate(self, n_rays: int, warm: bool = True):
        from oracle import dgs_oracle as O
        wl, rays = self.scene.workload, self.scene.rays
        mid = rays.shape[1] // 2
        lo = max(0, mid - n_rays // 2)
        sub = rays[:, lo:lo + n_rays].contiguous()
        args = (wl.n_samples, wl.n_candidates, wl.n_gaussian, wl.white_bkgd)
        with torch.no_grad():
            if warm:
                O.render(self.scene.model, sub[:, :256].contiguous(), *args)
            t0 = time.perf_counter()
            O.render(self.scene.model, sub, *args)
            dt = time.perf_counter() - t0
        n = sub.shape[0] * sub.shape[1]
        return n / dt, n, dt


L2_NOTE = ("GPU arm: flushed between timed steps (256 MiB write); CPU reference arm: the working set of a step "
           "(GBs of activations) is far beyond the last-level cache")


def eval_config(wl, workload, n_rays_frame, extra):
    return dict(workload=workload + "_eval", rays_per_frame=n_rays_frame, samples_per_ray=wl.n_samples,
                n_candidates=wl.n_candidates, n_gaussian=wl.n_gaussian, n_views=wl.NV, source_hw=[wl.H, wl.W],
                l2=L2_NOTE, **extra)


def parallelism_note(world: int, shard: bool, train: bool) -> str:
    if world == 1:
        return "single GPU"
    if shard:
        return (f"one frame, rays sharded contiguously over {world} ranks (render_sharded), NCCL all-gather of rgb/depth "
                "in the timed region, scene broadcast once per scene")
    if train:
        return "SB=2 x 4096 rays per rank, NCCL all-reduce of 46.7 MB gradients per step"
    return f"one independent frame per rank ({world} replicas), no data-plane collective"


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (torch-CPU oracle, all host threads)."""
    if rank != 0:
        return
    cpu = CpuOracle(args.workload)
    wl = cpu.scene.workload
    chunk = 4096      # the reference's own ray_batch_size (src/models/diner.py:57)
    rates, t_begin = [], time.time()
    for i in range(args.warmup + args.steps):
        rate, n, _ = cpu.rate(chunk, warm=(i == 0))
        if i >= args.warmup:
            rates.append(rate)
        if time.time() - t_begin > 150 and rates:      # keep the whole run within a few minutes on slow hosts
            break
    if not rates:
        rates = [rate]
    value = statistics.mean(rates)
    n_frame = wl.tgt_H * wl.tgt_W * wl.SB
    line = dict(impl="reference", metric="render_rays_per_s", value=value, unit="rays/s", n_gpus=args.gpus,
                steps=len(rates), warmup=args.warmup, ms_per_step=1e3 * chunk / value, higher_is_better=True,
                scaling="strong" if world > 1 else "weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=eval_config(wl, args.workload, n_frame, dict(
                    rays_per_gpu_per_step=n_frame // world if world > 1 else n_frame,
                    parallelism=parallelism_note(world, world > 1, False))),
                msamples_per_s=value * wl.n_samples / 1e6,
                cpu_baseline=dict(value=value, unit="rays/s", cores=cpu.threads, kind="port",
                                  sample=f"bounded: one {chunk}-ray chunk (the reference's ray_batch_size, diner.py:57) of "
                                         f"the {args.workload} frame per step -- rays/s is per ray, so it compares with "
                                         "the full-frame figure; torch-CPU oracle, bit-pinned to the unmodified "
                                         "reference's outputs (tests/test_oracle.py)"),
                e2e=dict(value=value, unit="rays/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    emit(line)


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------

GEMM_CALLS = ("dnr_mlp_quad_fwd", "dnr_mlp_pair_fwd", "dnr_mlp_fused_fwd", "dnr_mlp_gemm", "dnr_mlp_wgrad")


class Harness:
    def __init__(self, dev, rank, world):
        self.dev, self.rank, self.world = dev, rank, world
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)      # > 126 MB L2
        self.peaks = load_peaks()
        self.scenes = {}

    def scene(self, workload):
        if workload not in self.scenes:
            from tools.synthetic import make_scene
            self.scenes = {}                       # one scene resident at a time (the Multiface latent is 1.9 GB fp32)
            s = make_scene(workload, seed=0)
            s.model.to_device(self.dev)
            self.scenes[workload] = s
        return self.scenes[workload]

    def timed(self, fn, steps, warmup):
        """W untimed steps, then `steps` steps with the L2 flushed before each; CUDA events; barrier + synchronize on both
        sides; max over ranks.  Returns (total ms, C-ABI launches, wall-clock window)."""
        from diner_b200 import _lib
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        evs = []
        n0 = _lib.launch_count
        t0 = time.time()
        for _ in range(steps):
            self.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t1 = time.time()
        total_ms = sum(a.elapsed_time(b) for a, b in evs)
        launches = _lib.launch_count - n0
        if self.world > 1:
            t = torch.tensor([total_ms], device=self.dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms = float(t.item())
        return total_ms, launches, (t0, t1)

    @staticmethod
    def _library_timings():
        """(quad ms, pair ms, timed launch pairs) accumulated by dnr_render_fwd on this device (dnr_render_timings)."""
        import ctypes as C
        from diner_b200 import _lib
        torch.cuda.synchronize()
        q, pr, n = C.c_double(0), C.c_double(0), C.c_longlong(0)
        lib = _lib.load()
        if lib.dnr_render_timings(C.byref(q), C.byref(pr), C.byref(n)) != 0:
            return 0.0, 0.0, 0
        return q.value, pr.value, n.value

    def kernel_breakdown(self, fn, prof_steps=3):
        """Per-C-ABI-call CUDA-event timing on separate instrumented steps (not part of the headline numbers)."""
        from diner_b200 import _lib
        _lib.profile = []
        for _ in range(prof_steps):
            self.flush.zero_()
            fn()
        torch.cuda.synchronize()
        per = {}
        for name, e0, e1, tag in _lib.profile:
            if tag == "overlapped":              # runs on a second stream beside another kernel: not additive
                name += " (overlapped)"
            d = per.setdefault(name, [0.0, 0])
            d[0] += e0.elapsed_time(e1); d[1] += 1
        _lib.profile = None
        return {k: dict(ms_per_step=v[0] / prof_steps, launches_per_step=v[1] // prof_steps) for k, v in per.items()}

    def run(self, workload, mode, steps, warmup, shard=False, max_rays=None, with_e2e=True, allreduce=False,
            clocks=None):
        """mode "eval": one frame (or its central `max_rays` rays; with `shard` the frame is split over the ranks,
        otherwise every rank renders its own frame of a camera sweep); mode "train": forward + scalar loss + backward
        of the workload's ray batch on every rank (+ gradient all-reduce with `allreduce`)."""
        from diner_b200 import NeRFRendererDGS, field as F_
        from diner_b200.sharding import allreduce_gradients, render_sharded
        from tools.synthetic import target_rays
        dev, world, rank = self.dev, self.world, self.rank
        scene = self.scene(workload)
        wl, model = scene.workload, scene.model
        train = mode == "train"
        azimuth = 2.0 if (shard or world == 1) else 2.0 + 6.0 * rank     # replicas: a camera sweep, one frame per rank
        rays_cpu = target_rays(scene, azimuth_deg=azimuth, elevation_deg=8.0)
        if max_rays is not None and rays_cpu.shape[1] > max_rays:
            mid = rays_cpu.shape[1] // 2
            rays_cpu = rays_cpu[:, mid - max_rays // 2: mid + max_rays // 2].contiguous()
        n_rays = rays_cpu.shape[0] * rays_cpu.shape[1]
        renderer = NeRFRendererDGS(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
        rays_dev, rays_host = rays_cpu.to(dev), rays_cpu.pin_memory()
        rays_stage = torch.empty_like(rays_dev)
        rgb_host = torch.empty(rays_cpu.shape[0], rays_cpu.shape[1], 3).pin_memory()
        depth_host = torch.empty(rays_cpu.shape[0], rays_cpu.shape[1]).pin_memory()
        enc_bucket = None
        model.encoder.latent = model.encoder.latent.detach().requires_grad_(train)
        if train and allreduce and world > 1:
            # DDP of the reference all-reduces every parameter gradient (configs/train_diner_facescape.yaml:87,
            # 11.67 M fp32 = 46.7 MB).  The MLP's 3.45 M gradients come from our backward; the
            # encoder trunk is stock PyTorch and not part of this bench, so a bucket of its size stands in for it.
            n_mlp = sum(p.numel() for p in model.mlp_fine.parameters())
            enc_bucket = torch.zeros(max(0, 11_670_000 - n_mlp), device=dev)

        def render(rays):
            if train:
                out = renderer(model, rays)
                out.fine.rgb.square().mean().backward()
                if enc_bucket is not None:
                    allreduce_gradients(list(model.mlp_fine.parameters()))
                    dist.all_reduce(enc_bucket)
                return out.fine.rgb.detach(), out.fine.depth.detach()
            if shard:    # one frame for all ranks: contiguous ray slices, all-gather of rgb / depth (NCCL)
                return render_sharded(renderer, model, rays)
            with torch.no_grad():
                out = renderer(model, rays)
            return out.fine.rgb, out.fine.depth

        def step_device():
            render(rays_dev)

        def step_e2e():
            rays_stage.copy_(rays_host, non_blocking=True)
            rgb, depth = render(rays_stage)
            rgb_host.copy_(rgb, non_blocking=True)
            depth_host.copy_(depth, non_blocking=True)

        if clocks is not None:
            clocks.start()
        lib_t0 = self._library_timings()
        total_ms, launches, window = self.timed(step_device, steps, warmup)
        lib_t1 = self._library_timings()
        clk = clocks.stop(*window) if clocks is not None else None
        e2e_ms = self.timed(step_e2e, steps, 2)[0] if with_e2e else None
        F_.LAST_SPLIT.clear()                                            # (a training step never sets it: no stale record)
        breakdown = self.kernel_breakdown(step_device)
        split = dict(F_.LAST_SPLIT)

        rays_total = n_rays * (1 if shard else world)                    # all ranks' rays per step
        ms = total_ms / steps
        n_local = n_rays // world if shard else n_rays                   # rays this rank renders per step
        n_samples = n_local * wl.n_samples
        peaks = self.peaks
        flops = mlp_flops_per_sample(wl.NV) * n_samples * (3.0 if train else 1.0)
        gemm_ms = sum(breakdown[k]["ms_per_step"] for k in GEMM_CALLS if k in breakdown)
        gemm_launches = sum(breakdown[k]["launches_per_step"] for k in GEMM_CALLS if k in breakdown)
        achieved = flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
        kernel_name = ("mlp_quad_fwd_kernel (tcgen05 cta_group::2, 4-CTA clusters, whole ResnetFC per launch; "
                       "mlp_pair_fwd_kernel on the spare SMs, overlapped)" if "dnr_mlp_quad_fwd" in breakdown
                       else "mlp_pair_fwd_kernel (tcgen05, CTA pair, whole ResnetFC per launch)" if "dnr_mlp_pair_fwd" in breakdown
                       else "mlp_fused_fwd_kernel (tcgen05)" if "dnr_mlp_fused_fwd" in breakdown
                       else "mlp_gemm_kernel (fwd + dgrad) + mlp_wgrad2_kernel (tcgen05)")
        timing = ("CUDA events around every C-ABI call on 3 instrumented steps right after the timed region (call-by-call "
                  "sequence: same kernels, arguments and order)")
        if not train and lib_t1[2] > lib_t0[2] and "dnr_mlp_quad_fwd" in breakdown:
            # evaluation through dnr_render_fwd: the library times its cluster-kernel launches itself (CUDA events on the
            # launching stream), so the dominant kernel's duration comes from INSIDE the timed region, warm-up steps
            # included in the average (they run the same launches)
            n_pairs = lib_t1[2] - lib_t0[2]
            per_launch = (lib_t1[0] - lib_t0[0]) / n_pairs
            gemm_launches = breakdown["dnr_mlp_quad_fwd"]["launches_per_step"]
            gemm_ms = per_launch * gemm_launches
            achieved = flops / (gemm_ms * 1e-3) / 1e12
            timing = (f"CUDA events recorded by the library around each dnr_mlp_quad_fwd launch inside the timed region "
                      f"(dnr_render_timings: {n_pairs} launches averaged; the 2-CTA-cluster kernel beside it averaged "
                      f"{(lib_t1[1] - lib_t0[1]) / n_pairs:.3f} ms per launch); the instrumented pass after the region "
                      f"measured {breakdown['dnr_mlp_quad_fwd']['ms_per_step']:.3f} ms per step for the same kernel")
        roofline = dict(bound="tensor", kernel=kernel_name, achieved=achieved, peak=peaks["bf16_tflops_sustained"],
                        unit="TFLOP/s", frac=achieved / peaks["bf16_tflops_sustained"], traffic=None,
                        frac_of_burst_peak=achieved / peaks["bf16_tflops"],
                        peak_source=f"{peaks['source']} sustained bf16 (kernel timed inside a long step)",
                        launches_per_step=gemm_launches, kernel_ms_per_step=gemm_ms, algorithmic_flops_per_step=flops,
                        tiles_per_launch=split or None,
                        timing=timing)
        self._attach_traffic(roofline, workload, mode, split)

        def _gbs(nbytes, key):
            t = breakdown.get(key, {}).get("ms_per_step", 0.0)
            return nbytes / (t * 1e-3) / 1e9 if t > 0 else None
        rows = n_samples * wl.NV
        hbm = peaks["hbm_gbs"]
        enc = _gbs(rows * 1152.0, "dnr_encode_fwd")
        comp = _gbs(n_local * 1008.0, "dnr_composite_fwd")
        stage = dict(
            encode=dict(kernel="encode_fwd_kernel", hbm_gbs=enc, hbm_frac=enc / hbm if enc else None,
                        ms_per_step=breakdown.get("dnr_encode_fwd", {}).get("ms_per_step"),
                        note="algorithmic HBM bytes: 1152 B written per view row (the MLP's operands); the taps are read "
                             "from the bf16 latent (L2-resident up to ~100 MB; 0.94 GB at the Multiface shape)"),
            composite=dict(kernel="composite_fwd3_kernel", hbm_gbs=comp, hbm_frac=comp / hbm if comp else None,
                           launches_per_step=breakdown.get("dnr_composite_fwd", {}).get("launches_per_step"),
                           note="1008 B per ray, one launch per renderer call"),
            sampler=dict(kernel="sampler_kernel", hbm_gbs=_gbs(n_local * 192.0, "dnr_sampler_fwd"),
                         l2_gbs=_gbs(n_local * wl.NV * wl.n_candidates * 20.0, "dnr_sampler_fwd"),
                         ms_per_step=breakdown.get("dnr_sampler_fwd", {}).get("ms_per_step"),
                         note="not memory-bound: NV*C candidate-view evaluations per ray on cache-resident maps"))
        if train:
            bwd = _gbs(n_local * 1616.0, "dnr_composite_bwd")
            stage["composite_bwd"] = dict(kernel="composite_bwd2_kernel", hbm_gbs=bwd, hbm_frac=bwd / hbm if bwd else None)
            stage["encode_bwd"] = dict(kernel="encode_bwd_kernel",
                                       ms_per_step=breakdown.get("dnr_encode_bwd", {}).get("ms_per_step"))
        value = rays_total / (ms * 1e-3)
        res = dict(value=value, unit="rays/s", ms_per_step=ms, msamples_per_s=value * wl.n_samples / 1e6,
                   rays_per_step=rays_total, rays_per_gpu_per_step=n_local, gpu_launches=launches,
                   roofline=roofline, stage_rooflines=stage, breakdown_ms=breakdown, clocks=clk,
                   config=eval_config(wl, workload, n_rays, {}) | dict(
                       workload=workload + ("_train_fwd_bwd" if train else "_eval")))
        if with_e2e:
            e2e_step = e2e_ms / steps
            res["e2e"] = dict(value=rays_total / (e2e_step * 1e-3), unit="rays/s", h2d_bytes_per_step=rays_host.numel() * 4,
                              d2h_bytes_per_step=(rgb_host.numel() + depth_host.numel()) * 4, ms_per_step=e2e_step)
        model.encoder.latent = model.encoder.latent.detach()
        for p in model.mlp_fine.parameters():
            p.grad = None
        return res

    def _attach_traffic(self, roofline, workload, mode, split):
        """dram bytes of ONE launch of the dominant kernel from the committed ncu --set full capture -- only when that
        capture was taken at the tile count the timed launches actually have."""
        path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
        if not os.path.exists(path):
            return
        with open(path) as f:
            rec = json.load(f).get(f"{workload}_{mode}")
        if not isinstance(rec, dict):
            return
        # the capture must be of the launch geometry that was timed: same tiles per chunk, and the quad kernel's share of
        # them (the split with the pair kernel adapts at run time) within 2 %
        same = (rec.get("tiles_chunk") == split.get("tiles") and split.get("tiles_main") and
                abs(rec.get("tiles", 0) - split["tiles_main"]) <= 0.02 * split["tiles_main"])
        if same:
            roofline["traffic"] = rec["dram_bytes"]
            roofline["traffic_over_algorithmic"] = rec["dram_bytes"] / rec["algorithmic_bytes"]
            roofline["traffic_tiles"] = rec["tiles"]
            roofline["traffic_source"] = rec.get("source")
        else:
            roofline["traffic_note"] = (f"profiles/roofline_traffic.json holds a capture at {rec.get('tiles')} of "
                                        f"{rec.get('tiles_chunk')} tiles per launch, this run launches "
                                        f"{split.get('tiles_main')} of {split.get('tiles')}: not reported")

    def torch_cuda_reference(self, workload, n_chunks=4):
        """The reference ALGORITHM (oracle/dgs_oracle.py) in stock torch CUDA ops on this GPU, chunked like the reference
        (4096 rays per call): the "reference on a GPU today" figure of SURVEY.md 8(d).  Never on the product path."""
        from oracle import dgs_oracle as O
        scene = self.scene(workload)
        wl = scene.workload
        rays = scene.rays.to(self.dev)
        mid = rays.shape[1] // 2
        chunks = [rays[:, mid + i * 4096: mid + (i + 1) * 4096].contiguous()
                  for i in range(-(n_chunks // 2), n_chunks - n_chunks // 2)]
        args = (wl.n_samples, wl.n_candidates, wl.n_gaussian, wl.white_bkgd)
        with torch.no_grad():
            O.render(scene.model, chunks[0], *args)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for c in chunks:
                O.render(scene.model, c, *args)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        n = sum(c.shape[0] * c.shape[1] for c in chunks)
        return dict(value=n / dt, unit="rays/s", impl="reference algorithm, stock torch CUDA ops (fp32), 4096-ray calls",
                    rays=n, seconds=dt)

    def novel_variants(self, steps=4, warmup=3, n_vertices=5023):
        """the NOVEL and NOVEL_PE renderer variants (diner_b200.novel / novel_pe) on the FaceScape frame
        against a FLAME-sized target mesh -- evaluation frame and the training step of the config-4 shape."""
        from diner_b200.novel import NeRFRendererDGS as NovelRenderer
        from tools.synthetic import add_novel_inputs, add_novel_pe_inputs, make_scene, novel_to_device
        out = {}
        self.scenes = {}
        for tag, pe in (("novel", False), ("novel_pe", True)):
            for mode, wl_name in (("eval", "facescape"), ("train_fwd_bwd", "facescape_train")):
                scene = add_novel_inputs(make_scene(wl_name, seed=0), n_vertices=n_vertices, seed=0, gen_hw=(192, 192))
                if pe:
                    scene = add_novel_pe_inputs(scene, seed=0)
                novel_to_device(scene, self.dev)
                wl, m = scene.workload, scene.model
                rays = scene.rays.to(self.dev)
                r = NovelRenderer(wl.n_samples, wl.n_candidates, wl.n_gaussian, white_bkgd=wl.white_bkgd)
                args = (m, rays, scene.target_vertices, scene.tgt_in_offsets, scene.tgt_gen_offsets)
                if mode == "eval":
                    def step():
                        with torch.no_grad():
                            r(*args)
                else:
                    m.encoder.latent = m.encoder.latent.detach().requires_grad_(True)
                    m.gen_latent = m.gen_latent.detach().requires_grad_(True)

                    def step():
                        for q in (*m.parameters(), m.gen_latent, m.encoder.latent):
                            q.grad = None
                        r(*args).fine.rgb.square().mean().backward()
                total_ms, launches, _ = self.timed(step, steps, warmup)
                n = rays.shape[0] * rays.shape[1]
                out[f"{tag}_{mode}"] = dict(value=n / (total_ms / steps * 1e-3), unit="rays/s", ms_per_step=total_ms / steps,
                                            rays_per_step=n, n_vertices=n_vertices, gpu_launches=launches)
                del scene, m, rays, r, args
                torch.cuda.empty_cache()
        return out

    def scene_broadcast(self, workload, repeats=3):
        """ms of one diner_b200.sharding.broadcast_scene (latent + maps + cameras + MLP weights in one flat bucket)."""
        from diner_b200.sharding import broadcast_scene
        model = self.scene(workload).model
        nbytes = sum(getattr(model.encoder, n).numel() * 4 for n in ("latent", "depths", "depths_std", "normals"))
        nbytes += sum(p.numel() * 4 for p in model.mlp_fine.parameters())
        broadcast_scene(model, src=0)
        torch.cuda.synchronize()
        times = []
        for _ in range(repeats):
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            broadcast_scene(model, src=0)
            e1.record()
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1))
        t = torch.tensor([min(times)], device=self.dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        return dict(ms_per_scene=ms, bytes=nbytes, gbs=nbytes / (ms * 1e-3) / 1e9,
                    note="once per scene, outside the per-frame timed region; rank 0 -> all, one flat fp32 bucket, NCCL")


_REAL_STDOUT = None


def quiet_stdout():
    """stdout must carry ONE JSON line and nothing else, but libraries write there at C level (NCCL prints its version
    banner on communicator creation): point fd 1 at stderr for the whole run and keep the real stdout for emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="facescape")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the additional workloads (key 'workloads')")
    ap.add_argument("--cpu-rays", type=int, default=4096)
    ap.add_argument("--train", action="store_true", help="headline = forward+backward of the workload's ray batch")
    ap.add_argument("--replicas", action="store_true",
                    help="N > 1: headline = one independent frame per rank (weak scaling) instead of one sharded frame")
    ap.add_argument("--max-rays", type=int, default=None)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    args.warmup = max(args.warmup, 3)

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from diner_b200 import _lib
    _lib.load()            # fails loudly when the CUDA library is missing (no fallback)

    H = Harness(dev, rank, world)
    shard = world > 1 and not args.train and not args.replicas
    bcast = H.scene_broadcast(args.workload) if world > 1 else None
    main_res = H.run(args.workload, "train" if args.train else "eval", args.steps, args.warmup, shard=shard,
                     max_rays=args.max_rays, allreduce=True, clocks=ClockSampler(local_rank))

    extra = {}
    if not args.no_extra and not args.train and args.workload == "facescape":
        es, ew = max(3, args.steps // 4), 3
        if world == 1:
            extra["torch_cuda_reference"] = H.torch_cuda_reference("facescape")
            extra["facescape_train_fwd_bwd"] = H.run("facescape_train", "train", max(5, args.steps // 2), ew)
            extra["dtu_eval"] = H.run("dtu", "eval", es, ew)
            extra["multiface_eval_131072_rays"] = H.run("multiface", "eval", es, ew, max_rays=131072)
            extra.update(H.novel_variants())
        else:
            extra["facescape_eval_replicas"] = H.run("facescape", "eval", es, ew, shard=False, with_e2e=False) | dict(
                scaling="weak", note="one independent frame of a camera sweep per rank, no data-plane collective")
            extra["facescape_train_fwd_bwd_allreduce"] = H.run(
                "facescape_train", "train", max(5, args.steps // 2), ew, allreduce=True, with_e2e=False) | dict(
                scaling="weak", note="SB = 2 x 4096 rays per rank; NCCL all-reduce of 46.7 MB of gradients per step "
                                     "(MLP grads in one flat bucket + an encoder-sized bucket) inside the timed region")

    if rank == 0:
        par = parallelism_note(world, shard, args.train)
        line = dict(metric="render_rays_per_s", value=main_res["value"], unit="rays/s", n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=main_res["ms_per_step"], higher_is_better=True,
                    scaling="strong" if shard else "weak", vs_baseline=None, dtype="bf16", data="synthetic",
                    config=main_res["config"] | dict(rays_per_gpu_per_step=main_res["rays_per_gpu_per_step"],
                                                     parallelism=par),
                    msamples_per_s=main_res["msamples_per_s"], e2e=main_res["e2e"],
                    gpu_launches=main_res["gpu_launches"], clocks=main_res["clocks"], roofline=main_res["roofline"],
                    stage_rooflines=main_res["stage_rooflines"], breakdown_ms=main_res["breakdown_ms"])
        if bcast is not None:
            line["scene_broadcast"] = bcast
        if extra:
            line["workloads"] = extra
        if world == 1 and not args.no_cpu_baseline:
            cpu = CpuOracle(args.workload)
            rate, n, dt = cpu.rate(args.cpu_rays)
            line["cpu_baseline"] = dict(value=rate, unit="rays/s", cores=cpu.threads, kind="port",
                                        sample=f"{n} rays (one reference chunk) of the same frame, torch-CPU oracle "
                                               f"(bit-pinned to the unmodified reference), {dt:.1f} s")
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
