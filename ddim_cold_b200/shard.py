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
    local = local_sampler(noise[lo:hi]) if hi > lo else None
    if gather_device is not None:
        dev = torch.device(gather_device)
    elif local is not None:
        dev = local.device
    else:       # an empty slice (n < world * ceil(n / world)) has no tensor to take the device from: NCCL needs CUDA
        dev = torch.device("cuda", torch.cuda.current_device()) \
            if dist.is_initialized() and dist.get_backend(group) == "nccl" else torch.device("cpu")
    if local is None:
        local = torch.empty((0,) + tuple(img_shape), dtype=torch.float32, device=dev)
    if not gather or world == 1:
        return local
    pad = torch.zeros((per,) + tuple(img_shape), dtype=torch.float32, device=dev)
    pad[: hi - lo] = local.to(dev)
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    parts = []
    for r in range(world):
        a, b = shard_bounds(n, world, r)
        parts.append(bufs[r][: b - a])
    return torch.cat(parts, dim=0)


class SharedHostImages:
    """One page-locked host tensor [N,3,H,W] shared by all ranks of a node (a /dev/shm file mapped by every process).

    With one process per GPU the finished images leave the devices over eight PCIe links at once -- each rank copies
    its own slice straight into the common buffer -- instead of being funnelled through rank 0's single link after a
    gather (8 x 50 MB through one link cost 4 % of the 8-GPU run).  Rank 0 (or any rank) then reads all N images from
    its own mapping.  Collective calls: one broadcast of the file name at construction, one barrier per `wait_all`.
    """

    def __init__(self, n: int, img_shape, group=None):
        import os
        import uuid
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        name = [f"/dev/shm/ddimcold_{os.getpid()}_{uuid.uuid4().hex}.f32" if self.rank == 0 else None]
        numel = n
        for d in img_shape:
            numel *= int(d)
        self.path = None
        if self.world > 1:
            if self.rank == 0:
                with open(name[0], "wb") as f:
                    f.truncate(numel * 4)
            dist.broadcast_object_list(name, src=0, group=group)
            self.path = name[0]
            self.flat = torch.from_file(self.path, shared=True, size=numel, dtype=torch.float32)
        else:
            self.flat = torch.empty(numel, dtype=torch.float32)
        self.tensor = self.flat.view((n,) + tuple(int(d) for d in img_shape))
        self._registered = False
        if torch.cuda.is_available():
            rc = torch.cuda.cudart().cudaHostRegister(self.flat.data_ptr(), numel * 4, 0)
            self._registered = int(rc) == 0
        if self.world > 1:
            dist.barrier(group=group)            # every rank has mapped the file: the name can go
            if self.rank == 0:
                os.unlink(self.path)

    def put(self, lo: int, images: torch.Tensor):
        """Asynchronous copy of this rank's finished images (device or host) into rows [lo, lo + len)."""
        self.tensor[lo:lo + images.shape[0]].copy_(images, non_blocking=True)

    def wait_all(self):
        """Every rank's `put` has landed (device sync + one barrier)."""
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=self.group)
        return self.tensor

    def close(self):
        if self._registered:
            torch.cuda.cudart().cudaHostUnregister(self.flat.data_ptr())
            self._registered = False
