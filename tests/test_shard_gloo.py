"""CPU, world_size 2 over gloo: the batch-sharding host logic of the sampler (no step-path
collective, one gather of the finished images, results independent of the world size).
The per-rank "sampler" here is a deterministic per-image function standing in for the CUDA
loop -- what is under test is the slicing, the seed-compatible noise draw and the gather."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_local_sampler(noise):
    # per-image, order-independent: any cross-sample mixing or misplaced slice changes the result
    return torch.tanh(noise * 0.5 + noise.flatten(1).mean(dim=1).view(-1, 1, 1, 1))


def _worker(rank, world, port, n, out_dir):
    sys.path.insert(0, ROOT)
    from ddim_cold_b200.shard import sample_sharded, shard_bounds
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(1234)                                  # same seed on every rank, like the reference's single process
    got = sample_sharded(_fake_local_sampler, n, (3, 8, 8))
    lo, hi = shard_bounds(n, world, rank)
    # the parallel-download path: every rank drops its slice into one host buffer shared through /dev/shm
    from ddim_cold_b200.shard import SharedHostImages
    shared = SharedHostImages(n, (3, 8, 8))
    if hi > lo:
        shared.put(lo, got[lo:hi])
    host_all = shared.wait_all().clone()
    shared.close()
    torch.save({"all": got, "lo": lo, "hi": hi, "shared": host_all}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [7, 8, 1])
def test_sharded_sampling_matches_single_process(tmp_path, n):
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n, str(tmp_path)), nprocs=world, join=True)
    torch.manual_seed(1234)
    noise = torch.normal(0, 1, (n, 3, 8, 8))                 # the draw ViT.py:224 would make
    want = _fake_local_sampler(noise)
    r0 = torch.load(tmp_path / "r0.pt")
    r1 = torch.load(tmp_path / "r1.pt")
    assert torch.equal(r0["all"], want) and torch.equal(r1["all"], want)
    assert r0["lo"] == 0 and r0["hi"] == r1["lo"] and r1["hi"] == n
    assert torch.equal(r0["shared"], want) and torch.equal(r1["shared"], want)


def test_single_process_path_needs_no_process_group():
    from ddim_cold_b200.shard import sample_sharded
    torch.manual_seed(5)
    got = sample_sharded(_fake_local_sampler, 3, (3, 4, 4))
    torch.manual_seed(5)
    assert torch.equal(got, _fake_local_sampler(torch.normal(0, 1, (3, 3, 4, 4))))
