"""Probe: call the attention kernel alone at the S1 benchmark shape (for ncu captures and quick timing)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from ddim_cold_b200 import _lib

B, L, C, heads = 1024, 65, 384, 12
if len(sys.argv) > 1:
    B, L, C, heads = (int(v) for v in sys.argv[1:5])
lib = _lib.load()
qkv = (torch.randn(B * L + 128, 3 * C, device="cuda") * 0.5).to(torch.bfloat16)
out = torch.empty(B * L + 128, C, device="cuda", dtype=torch.bfloat16)
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    _lib.check(lib.ddc_k_attention(0, qkv.data_ptr(), out.data_ptr(), B, L, C, heads, st))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    _lib.check(lib.ddc_k_attention(0, qkv.data_ptr(), out.data_ptr(), B, L, C, heads, st))
e1.record()
torch.cuda.synchronize()
print("attention B=%d L=%d C=%d heads=%d: %.1f us" % (B, L, C, heads, e0.elapsed_time(e1) * 100))
