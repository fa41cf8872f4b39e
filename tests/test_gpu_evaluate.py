"""GPU: trainer-side forward reuse (reference multi_gpu_trainer.py:32-45) -- the smooth-L1 reduction kernel through the
C ABI against torch, and `evaluate()` against the losses the reference's own function returned (tests/golden/evaluate.npz)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 5, 4096, 6 * 3 * 32 * 32, 1024 * 3 * 64 * 64 + 3])
def test_smooth_l1_kernel_vs_torch(n):
    from ddim_cold_b200 import _lib
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(n % 1000 + 1)
    a = torch.randn(n, device="cuda", generator=g) * 1.5          # |a - b| on both sides of beta = 1
    b = torch.randn(n, device="cuda", generator=g)
    scratch = torch.zeros(1024, device="cuda")
    out = torch.full((2,), float("nan"), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    vals = []
    for _ in range(3):                                             # the ticket counter must come back to zero every time
        _lib.check(lib.ddc_k_smooth_l1(a.data_ptr(), b.data_ptr(), n, scratch.data_ptr(), out.data_ptr(), st))
        torch.cuda.synchronize()
        vals.append(float(out[0]))
    assert vals[0] == vals[1] == vals[2], vals                     # fixed summation order
    assert torch.isnan(out[1])
    ref = float(torch.nn.functional.smooth_l1_loss(a.double(), b.double()))
    assert abs(vals[0] - ref) <= 2e-6 * max(1.0, abs(ref)), (vals[0], ref)
    assert lib.ddc_k_smooth_l1(a.data_ptr(), b.data_ptr(), 0, scratch.data_ptr(), out.data_ptr(), st) == _lib.DDC_E_INVALID


def test_evaluate_vs_reference_fixture(golden):
    """The reference's evaluate() ran on CPU in fp32; here the forward runs with 16-bit tensor-core operands, so the
    per-batch losses agree to the forward's tolerance (|x0 err| < 1e-2 -> a mean-loss error far below 1e-3)."""
    from ddim_cold_b200 import DiffusionVisionTransformer, evaluate
    from oracle import ddim_oracle as O
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    sd = O.synth_state_dict(sh, seed=3, rich=True, gain=2.0)
    m = DiffusionVisionTransformer(img_size=[sh.img_h, sh.img_w], patch_size=sh.patch, embed_dim=sh.embed, depth=sh.depth,
                                   num_heads=sh.heads, total_steps=sh.total_steps)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()}, strict=True)
    m = m.cuda()
    batches = [tuple(torch.from_numpy(a) for a in b) for b in O.synth_eval_batches(sh)]
    g = golden("evaluate")
    got = evaluate(m, batches, torch.device("cuda"))
    print(f"evaluate(): {got:.6f} vs reference {float(g['mean']):.6f}")
    assert abs(got - float(g["mean"])) < 5e-4
    assert not m.training                                          # evaluate() switched the model to eval like the reference
    m.operand_format = "fp32"
    got32 = evaluate(m, batches, torch.device("cuda"))
    assert abs(got32 - float(g["mean"])) < 2e-6
