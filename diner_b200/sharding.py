"""Multi-GPU use of the renderer: one process per GPU, target rays sharded across ranks.

The render path has no exchange step (rays are independent), so the only collectives are
  * per scene: broadcast of the encoded scene (feature map, depth/std/normal maps, cameras) and of the MLP
    weights from the rank that ran the encoder             -> `broadcast_scene`
  * end of frame: all-gather of the rank-local rgb / depth slices (<= 1 MB)         -> `render_sharded`
  * training: sum of the parameter / feature-map gradients (what DDP does for the reference,
    configs/train_diner_facescape.yaml:87)                 -> `allreduce_gradients`
all through torch.distributed (NCCL over NVLink on the GPU box; gloo in the CPU tests).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced split of n rays: the first n % world ranks get one extra ray."""
    base, extra = divmod(n, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def _world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def render_sharded(renderer, model, rays: torch.Tensor, group=None, **kw):
    """Render rays (SB,B,8) with every rank taking a contiguous slice of B; returns the full-frame
    (rgb (SB,B,3), depth (SB,B)) on every rank.  Forward only (eval)."""
    rank, world = _world(group)
    SB, B, _ = rays.shape
    lo, hi = shard_range(B, rank, world)
    with torch.no_grad():
        out = renderer(model, rays[:, lo:hi].contiguous(), **kw) if hi > lo else None
    if world == 1:
        return out.fine.rgb, out.fine.depth
    width = (B + world - 1) // world                     # equal-sized slots for all_gather
    slot = torch.zeros(SB, width, 4, device=rays.device, dtype=torch.float32)
    if hi > lo:
        slot[:, :hi - lo, :3] = out.fine.rgb
        slot[:, :hi - lo, 3] = out.fine.depth
    parts = [torch.empty_like(slot) for _ in range(world)]
    dist.all_gather(parts, slot, group=group)
    full = torch.cat([parts[r][:, :shard_range(B, r, world)[1] - shard_range(B, r, world)[0]] for r in range(world)], 1)
    return full[..., :3].contiguous(), full[..., 3].contiguous()


_SCENE_ENCODER_TENSORS = ("latent", "depths", "depths_std", "normals")
_SCENE_MODEL_TENSORS = ("poses", "focal", "c", "image_shape")


def broadcast_scene(model, src: int = 0, group=None, with_weights: bool = True) -> None:
    """Broadcast the encoded scene (and the MLP parameters) from rank `src` in place.  Every rank must already
    hold tensors of the right shape (e.g. from its own encode() of dummy inputs or torch.empty).

    All tensors travel in ONE flat fp32 bucket (NVSwitch: one launch instead of 8 + 30 small ones; the latent
    dominates: 151 MB at FaceScape shape), and the renderer's per-scene / per-weights caches are dropped afterwards:
    they key on the tensors' version counters, which neither a c10d broadcast nor a `.data` write bumps."""
    _, world = _world(group)
    if world == 1:
        return
    rank = dist.get_rank(group)
    slots = [(model.encoder, n) for n in _SCENE_ENCODER_TENSORS] + [(model, n) for n in _SCENE_MODEL_TENSORS]
    tensors = [getattr(o, n).detach() for o, n in slots]
    if with_weights:
        tensors += [p.data for p in model.mlp_fine.parameters()]
    dev = tensors[0].device
    sizes = [t.numel() for t in tensors]
    flat = torch.empty(sum(sizes), device=dev, dtype=torch.float32)
    if rank == src:
        torch.cat([t.reshape(-1).to(torch.float32) for t in tensors], out=flat)
    dist.broadcast(flat, src, group=group)
    off = 0
    with torch.no_grad():
        for i, (t, n) in enumerate(zip(tensors, sizes)):
            piece = flat[off:off + n].view(t.shape).to(t.dtype)
            off += n
            if i < len(slots):
                o, name = slots[i]
                setattr(o, name, piece if rank != src else getattr(o, name))   # fresh tensors on the receivers
            elif rank != src:
                t.copy_(piece)
    from . import scene
    scene.invalidate()


def allreduce_gradients(tensors, group=None, average: bool = True) -> None:
    """Sum (or average) gradients across ranks in one flat bucket (NVSwitch: size for launch latency, not links)."""
    _, world = _world(group)
    grads = [t.grad for t in tensors if t is not None and t.grad is not None]
    if world == 1 or not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, group=group)
    if average:
        flat /= world
    off = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[off:off + n].view_as(g))
        off += n
