"""Small run of every step-path kernel family for compute-sanitizer (one tool per gpurun call):

    compute-sanitizer --tool memcheck  --log-file gpurun_out/memcheck.log  python scripts/sanitize_smoke.py
    compute-sanitizer --tool racecheck --log-file gpurun_out/racecheck.log python scripts/sanitize_smoke.py

Shapes are fixture-sized so the instrumented kernels finish in seconds: the fused block tail + CTA-pair GEMMs + the
65-token TMA attention (32x32 / p4 / C128), the separate-kernel fallback (TINY, C = 64), the cp.async attention for
L > 128 (64x64 / p4 / C128: L = 257), and the allocation-free forward.  Results are checked against the CPU oracle so a
silent corruption under instrumentation would also show.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from ddim_cold_b200 import DiffusionVisionTransformer, _lib, synth  # noqa: E402
from oracle import ddim_oracle as O  # noqa: E402  (checker)


def model_of(sh, seed):
    sd = synth.synth_state_dict(sh, seed=seed, rich=True)
    m = DiffusionVisionTransformer(img_size=[sh.img_h, sh.img_w], patch_size=sh.patch, embed_dim=sh.embed, depth=sh.depth,
                                   num_heads=sh.heads)
    m.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()}, strict=True)
    return m.eval().cuda(), sd


def main():
    cases = [("fused-path", synth.Shape(32, 32, 4, 128, 2, 4), 3), ("tiny", synth.TINY, 2), ("L257", synth.Shape(64, 64, 4, 128, 1, 2), 2)]
    for fmt in ("fp16", "bf16"):
        for name, sh, n in cases:
            m, sd = model_of(sh, 3)
            m.operand_format = fmt
            x_T = torch.from_numpy(synth.synth_noise((n, 3, sh.img_h, sh.img_w), seed=5))
            img = m.sampler("cuda", 500, n, noise=x_T).numpy()
            ref = O.sampler(sd, sh, x_T.numpy(), 500)
            err = float(np.abs(img - ref).max())
            print(f"\n{name} {fmt}: sampler k=500 N={n} max|cuda - oracle| = {err:.3e}")
            assert err < 2e-2
    # allocation-free forward twice, against the arena forward
    sh = cases[0][1]
    m, _ = model_of(sh, 4)
    lib = _lib.load()
    plan = m._plan(torch.device("cuda", 0), 3)
    st = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        x = torch.from_numpy(synth.synth_noise((3, 3, 32, 32), seed=30 + call)).cuda()
        t = torch.tensor([1999, 700, 19], dtype=torch.int32, device="cuda")
        a, b = torch.empty_like(x), torch.empty_like(x)
        _lib.check(lib.ddc_forward(plan.handle, x.data_ptr(), t.data_ptr(), 3, a.data_ptr(), st))
        _lib.check(lib.ddc_forward_fixed_scratch(plan.handle, x.data_ptr(), t.data_ptr(), 3, b.data_ptr(), st))
        torch.cuda.synchronize()
        assert torch.equal(a, b)
    print("sanitize_smoke: ok")


if __name__ == "__main__":
    main()
