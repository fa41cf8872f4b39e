"""Probe: what HBM delivers for pure writes, pure reads and the qkv GEMM's 1:3 read:write mix (torch fill / sum / copy)."""
import torch
def t(fn, n=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
big = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")   # 1 GiB: every pass misses L2
w = big[:153 * 10**6 * 4].view(torch.bfloat16)
us = t(lambda: w.fill_(1.0)); print("write 612 MB: %.1f us -> %.2f TB/s" % (us, 612e6 / us / 1e6))
r = big[:612 * 10**6].view(torch.float32)
us = t(lambda: r.sum()); print("read 612 MB: %.1f us -> %.2f TB/s" % (us, 612e6 / us / 1e6))
a = big[:300 * 10**6].view(torch.float32); b = big[512 * 2**20:512 * 2**20 + 300 * 10**6].view(torch.float32)
us = t(lambda: b.copy_(a)); print("copy 300 MB (600 MB moved): %.1f us -> %.2f TB/s" % (us, 600e6 / us / 1e6))
