"""ms per denoise step of sampler() at small batches (S1 shape, k = 10: 200 steps), for comparing kernel choices in the
latency-bound regime.   python scripts/latency_probe.py [N ...]"""
import contextlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddim_cold_b200 import DiffusionVisionTransformer

ns = [int(a) for a in sys.argv[1:]] or [1, 8, 32, 64, 128, 256]
torch.manual_seed(0)
m = DiffusionVisionTransformer(img_size=[64, 64], patch_size=8, embed_dim=384, depth=7, num_heads=12).eval().to("cuda")
for n in ns:
    noise = torch.randn(n, 3, 64, 64).pin_memory()
    best = 1e9
    with contextlib.redirect_stdout(sys.stderr):
        m.sampler("cuda", 10, n, noise=noise)
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            m.sampler("cuda", 10, n, noise=noise)
            best = min(best, time.perf_counter() - t0)
    print(f"QKV_ATTN={os.environ.get('DDC_QKV_ATTN', '1')} N={n}: {best / 200 * 1e3:.4f} ms per step, {n / best:.1f} images/s", flush=True)
