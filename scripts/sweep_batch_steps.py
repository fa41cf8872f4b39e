"""Config #4 on one GPU: images/s and ms per denoise step of sampler() over batch sizes and step jumps (S1 shape,
random-init weights, host buffers in and out).  Writes gpurun_out/sweep_s1.json."""
import json, os, sys, time, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddim_cold_b200 import DiffusionVisionTransformer

torch.manual_seed(0)
m = DiffusionVisionTransformer(img_size=[64, 64], patch_size=8, embed_dim=384, depth=7, num_heads=12).eval().to("cuda")
rows = []
for k in (50, 10, 1):
    for n in (1, 2, 8, 32, 128, 512, 2048, 8192):
        if k == 1 and n > 512:
            continue
        steps = len(range(1999, 0, -k))
        noise = torch.randn(n, 3, 64, 64).pin_memory()
        with contextlib.redirect_stdout(sys.stderr):
            m.sampler("cuda", k, n, noise=noise)           # warm-up: plan, graph
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            img = m.sampler("cuda", k, n, noise=noise)
            dt = time.perf_counter() - t0
        assert torch.isfinite(img).all()
        rows.append({"N": n, "k": k, "steps": steps, "seconds": dt, "images_per_s": n / dt, "ms_per_step": dt / steps * 1e3})
        print(rows[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/sweep_s1.json", "w"), indent=1)
