"""Batch sharding of the sampler over the GPUs of one node.

Samples never interact (no BatchNorm, attention is within-image, one scalar t per step), so
the N images are split contiguously over the ranks with no collective on the step path; the
only communication is one gather of the finished images (NCCL over NVLink on GPUs, gloo in
the CPU tests).  To stay seed-compatible with the single-process reference (ViT.py:224), every
rank draws the FULL [N,3,H,W] noise from the same-seeded global CPU RNG and keeps its slice,
so each sample's result is independent of the world size.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous split: rank r owns [r*ceil(n/world), ...) clipped to n (may be empty)."""
    per = -(-n // world)
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def sample_sharded(local_sampler, n: int, img_shape, noise: torch.Tensor | None = None, group=None,
                   gather: bool = True, gather_device=None):
    """Run ``local_sampler(noise_slice) -> CPU or device tensor [m,3,H,W]`` on this rank's
    slice and gather all slices on every rank (rank order == sample order).

    ``noise``: the full [N,3,H,W] tensor (same on every rank); drawn here from the global CPU
    RNG when None.  ``gather_device``: device to stage the collective on (cuda for NCCL).
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if noise is None:
        noise = torch.normal(0, 1, (n,) + tuple(img_shape))
    lo, hi = shard_bounds(n, world, rank)
    per = -(-n // world)
    local = local_sampler(noise[lo:hi]) if hi > lo else torch.empty((0,) + tuple(img_shape))
    if not gather or world == 1:
        return local
    dev = torch.device(gather_device) if gather_device is not None else local.device
    pad = torch.zeros((per,) + tuple(img_shape), dtype=torch.float32, device=dev)
    pad[: hi - lo] = local.to(dev)
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    parts = []
    for r in range(world):
        a, b = shard_bounds(n, world, r)
        parts.append(bufs[r][: b - a])
    return torch.cat(parts, dim=0)
