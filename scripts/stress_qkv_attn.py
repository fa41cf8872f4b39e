"""Randomised stress of the fused qkv + attention kernel (ddc_k_qkv_attn) against fp32 torch: random batch sizes (so every
work-unit split, partial tile pairs, single tiles), embed dims, both operand formats, repeated launches on the same
buffers (stale shared-memory / tensor-memory state from the previous launch must not leak).
    python scripts/stress_qkv_attn.py [cases] [seed]"""
import math
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ddim_cold_b200 import _lib  # noqa: E402

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
lib = _lib.load()
st = torch.cuda.current_stream().cuda_stream
L, d = 65, 32
worst = 0.0
for case in range(cases):
    Cd = rng.choice([64, 128, 192, 256, 384])
    heads = Cd // d
    B = rng.choice([1, 2, 3, 4, 5, 7, 8, 63, 64, 65, 147, 148, 149, 295, 296, 297, 300, 511, 600, 1024, rng.randint(1, 1200)])
    fmt = rng.choice([0, 1])
    dt = torch.float16 if fmt else torch.bfloat16
    M = B * L
    Mp = (M + 127) // 128 * 128
    g = torch.Generator(device="cuda").manual_seed(case)
    xn = torch.zeros(Mp, Cd, device="cuda", dtype=dt)
    xn[:M] = torch.randn(M, Cd, device="cuda", generator=g).to(dt)
    W = (torch.randn(3 * Cd, Cd, device="cuda", generator=g) * (1.5 / math.sqrt(Cd))).to(dt)
    b = torch.randn(3 * Cd, device="cuda", generator=g) * 0.2
    out = torch.full((M + 3, Cd), 7.0, device="cuda", dtype=dt)
    w_hp, b_hp = torch.empty_like(W), torch.empty_like(b)
    cls = torch.empty(B, 3 * Cd, device="cuda", dtype=dt)
    outs = []
    for rep in range(3):
        _lib.check(lib.ddc_k_qkv_attn(fmt, xn.data_ptr(), W.data_ptr(), b.data_ptr(), out.data_ptr(), B, Cd, heads, w_hp.data_ptr(),
                                      b_hp.data_ptr(), cls.data_ptr(), st))
        torch.cuda.synchronize()
        outs.append(out.clone())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2]), f"case {case}: launches differ (B={B} C={Cd} fmt={fmt})"
    qkv = (xn[:M].float() @ W.float().T + b).to(dt).float()
    x = qkv.reshape(B, L, 3, heads, d).permute(2, 0, 3, 1, 4)
    pr = ((x[0] @ x[1].transpose(-2, -1)) * d ** -0.5).softmax(dim=-1)
    ref = (pr @ x[2]).transpose(1, 2).reshape(M, Cd)
    err = (out[:M].float() - ref).abs().max().item() / max(1.0, ref.abs().max().item())
    worst = max(worst, err)
    assert err < 2e-2, f"case {case}: err {err} (B={B} C={Cd} fmt={fmt})"
    assert bool((out[M:].float() == 7.0).all()), f"case {case}: rows beyond M written"
print(f"{cases} cases ok, worst relative error {worst:.2e}")
