"""GPU: the allocation-free forward (`ddc_forward_fixed_scratch`) against `ddc_forward`.

Some hosts cannot allocate device memory when they need a prediction -- their memory belongs to
another framework's pool, they sit in a latency-critical loop where cudaMalloc's implicit device
synchronisation is not acceptable, or they are inside a region where allocation is forbidden.
For them the library keeps one fixed 32 MiB activation buffer of its own.  This test drives that
entry point through the C ABI with caller-owned inputs and outputs only, twice in a row (the
buffer keeps the previous call's contents, so the second call also proves nothing stale leaks
through), and compares with the ordinary entry point that uses the plan's work arena.  Both run
the same kernels, so the results must agree to the last bit; the oversize path must fail with a
message instead of running.

The file name sorts last on purpose: if it fails, `pytest -x` has already run everything else.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(sh, seed):
    from ddim_cold_b200 import DiffusionVisionTransformer
    from oracle import ddim_oracle as O
    sd = O.synth_state_dict(sh, seed=seed, rich=True)
    m = DiffusionVisionTransformer(img_size=[sh.img_h, sh.img_w], patch_size=sh.patch, embed_dim=sh.embed,
                                   depth=sh.depth, num_heads=sh.heads, total_steps=sh.total_steps)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()}, strict=True)
    return m.eval().cuda(), sd


@pytest.mark.parametrize("shape_name,batch", [("smoke", 5), ("s1", 3)])
def test_fixed_scratch_forward_matches_arena_forward(shape_name, batch):
    from ddim_cold_b200 import _lib
    from oracle import ddim_oracle as O
    lib = _lib.load()
    sh = {"smoke": O.Shape(32, 32, 4, 128, 2, 4), "s1": O.S1}[shape_name]
    model, sd = _model(sh, seed=4)
    dev = torch.device("cuda", torch.cuda.current_device())
    plan = model._plan(dev, batch)                      # packs the weights; owns the ordinary work arena
    st = torch.cuda.current_stream().cuda_stream
    for call in range(2):                                # two calls in a row, different inputs
        x = torch.from_numpy(O.synth_noise((batch, 3, sh.img_h, sh.img_w), seed=30 + call)).cuda()
        t = torch.tensor([1999, 700, 19, 1, 0][:batch], dtype=torch.int32, device="cuda")
        ref = torch.empty_like(x)
        got = torch.full_like(x, float("nan"))
        _lib.check(lib.ddc_forward(plan.handle, x.data_ptr(), t.data_ptr(), batch, ref.data_ptr(), st))
        _lib.check(lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), batch, got.data_ptr(), st))
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert torch.equal(got, ref), float((got - ref).abs().max())
        # and both agree with the CPU oracle within the bf16 tolerance of the other GPU tests
        y = O.forward(sd, sh, x.cpu().numpy(), t.cpu().numpy().astype(np.int64))
        assert np.abs(got.cpu().numpy() - y).max() < 2e-2


def test_fixed_scratch_rejects_oversize_batch():
    from ddim_cold_b200 import _lib
    from oracle import ddim_oracle as O
    lib = _lib.load()
    model, _ = _model(O.S1, seed=4)
    dev = torch.device("cuda", torch.cuda.current_device())
    plan = model._plan(dev, 1)
    x = torch.zeros(1, 3, 64, 64, device="cuda")
    t = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), 4096, x.data_ptr(), None)
    assert rc == _lib.DDC_E_INVALID
    msg = lib.ddc_last_error().decode()
    assert "built-in array" in msg and "at most" in msg
