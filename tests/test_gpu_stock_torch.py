"""GPU: stock PyTorch on the same B200 -- "the reference's own GPU kernel path" (BASELINE.md).

The reference class cannot travel to the GPU box, so this runs the oracle's torch port (the same ATen ops in the same
order, pinned bit-identical to the reference on CPU) on the device, in fp32 and under bf16 autocast, times the same
k-strided loop, and writes what it measured to ``gpurun_out/stock_torch_gpu.json`` (DESIGN.md quotes it).  It is a
reported baseline; the assertions only check that the hand-written path beats both and agrees with the fp32 run.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N, K = 512, 100          # 20 denoise steps: seconds, not minutes, for the eager arm


def _eager_loop(OT, sd, sh, x, k, autocast):
    T = sh.total_steps
    x0 = None
    for t in range(T - 1, 0, -k):
        steps = torch.full((x.shape[0],), t, dtype=torch.long, device=x.device)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            y = OT.forward(sd, sh, x, steps)
        x0 = torch.clamp(y.float(), -1, 1)
        a_tk = 1 - math.sqrt((t + 1 - k) / T)
        a_t = 1 - math.sqrt((t + 1) / T) + 1e-5
        eps = (x - math.sqrt(a_t) * x0) / math.sqrt(1 - a_t)
        x = math.sqrt(a_tk) * (x / math.sqrt(a_t) + (math.sqrt((1 - a_tk) / a_tk) - math.sqrt((1 - a_t) / a_t)) * eps)
    return (x0 + 1) / 2


def _time(fn, reps=3):
    fn()                                       # warm-up (cuBLAS heuristics, allocator)
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best, out


def test_hand_written_path_beats_stock_pytorch_on_the_same_gpu():
    from oracle import ddim_oracle as O
    from oracle import ddim_oracle_torch as OT
    from test_gpu_model import build, psnr

    sh = O.S1
    sd_np = O.synth_state_dict(sh, seed=0)
    sd = {k: v.cuda() for k, v in OT.to_torch_sd(sd_np).items()}
    x_T = torch.from_numpy(O.synth_noise((N, 3, sh.img_h, sh.img_w), seed=123)).cuda()
    model = build(sh, sd_np)

    with torch.no_grad():
        t_fp32, img_fp32 = _time(lambda: _eager_loop(OT, sd, sh, x_T.clone(), K, False))
        t_bf16, _ = _time(lambda: _eager_loop(OT, sd, sh, x_T.clone(), K, True))
        t_ours, img_ours = _time(lambda: model.sampler("cuda", K, N, noise=x_T.cpu(), return_on_device=True))
    report = {
        "workload": f"S1 N={N} k={K} ({len(range(1999, 0, -K))} steps), one B200",
        "stock_torch_fp32_images_per_s": N / t_fp32,
        "stock_torch_bf16_autocast_images_per_s": N / t_bf16,
        "ddim_cold_b200_images_per_s_incl_h2d": N / t_ours,
        "psnr_ours_vs_stock_fp32_db": psnr(img_ours.cpu().numpy(), img_fp32.cpu().numpy()),
        "torch": torch.__version__,
    }
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/stock_torch_gpu.json", "w") as f:
        json.dump(report, f, indent=1)
    print(json.dumps(report))
    assert report["psnr_ours_vs_stock_fp32_db"] > 40.0
    assert t_ours < t_bf16 and t_ours < t_fp32
