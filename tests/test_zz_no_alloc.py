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


def _largest_fitting_batch(lib, plan, x, t):
    """Ask the library: an absurd batch is refused with a message that names the largest batch that fits."""
    import re
    rc = lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), 1 << 20, x.data_ptr(), None)
    assert rc != 0
    return int(re.search(r"at most (\d+) images", lib.ddc_last_error().decode()).group(1))


def test_fixed_scratch_capacity_boundary_and_regrowth():
    """The largest batch the built-in 32 MiB array can hold runs and matches the arena path; one image more is
    refused.  The small call first makes the entry point lay the array out for fewer images than the big call needs,
    so the big call also exercises the re-layout."""
    from ddim_cold_b200 import _lib
    from oracle import ddim_oracle as O
    lib = _lib.load()
    sh = O.S1
    model, _ = _model(sh, seed=4)
    dev = torch.device("cuda", torch.cuda.current_device())
    st = torch.cuda.current_stream().cuda_stream
    probe_x = torch.zeros(1, 3, 64, 64, device="cuda")
    probe_t = torch.zeros(1, dtype=torch.int32, device="cuda")
    plan = model._plan(dev, 128)
    fit = _largest_fitting_batch(lib, plan, probe_x, probe_t)
    assert 32 <= fit <= 128, fit                           # about 64 S1 images (include/ddimcold.h)
    for batch in (3, fit):
        x = torch.from_numpy(O.synth_noise((batch, 3, 64, 64), seed=50 + batch)).cuda()
        t = torch.randint(0, 2000, (batch,), dtype=torch.int32).cuda()
        ref, got = torch.empty_like(x), torch.full_like(x, float("nan"))
        _lib.check(lib.ddc_forward(plan.handle, x.data_ptr(), t.data_ptr(), batch, ref.data_ptr(), st))
        _lib.check(lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), batch, got.data_ptr(), st))
        torch.cuda.synchronize()
        assert torch.equal(got, ref), (batch, float((got - ref).abs().max()))
    x = torch.zeros(fit + 1, 3, 64, 64, device="cuda")
    t = torch.zeros(fit + 1, dtype=torch.int32, device="cuda")
    rc = lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), fit + 1, x.data_ptr(), st)
    assert rc == _lib.DDC_E_INVALID and f"at most {fit} images" in lib.ddc_last_error().decode()


def test_fixed_scratch_leaves_the_plans_sampling_graph_alone():
    """A sampling loop (which caches a CUDA graph of one step inside the plan), then a fixed-scratch forward on the
    same plan, then the same sampling loop again: the two loops must agree bit for bit, i.e. the detour through the
    built-in array disturbed neither the plan's buffers nor its cached graph."""
    from ddim_cold_b200 import _lib
    from oracle import ddim_oracle as O
    lib = _lib.load()
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    model, _ = _model(sh, seed=4)
    x_T = torch.from_numpy(O.synth_noise((4, 3, 32, 32), seed=77))
    a = model.sampler("cuda", 200, 4, noise=x_T)
    dev = torch.device("cuda", torch.cuda.current_device())
    plan = model._plan(dev, 4)
    x = x_T.cuda()
    t = torch.full((4,), 999, dtype=torch.int32, device="cuda")
    ref, got = torch.empty_like(x), torch.empty_like(x)
    st = plan.stream.cuda_stream
    _lib.check(lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), 4, got.data_ptr(), st))
    _lib.check(lib.ddc_forward(plan.handle, x.data_ptr(), t.data_ptr(), 4, ref.data_ptr(), st))
    plan.stream.synchronize()
    assert torch.equal(got, ref)
    b = model.sampler("cuda", 200, 4, noise=x_T)
    assert torch.equal(a, b)
