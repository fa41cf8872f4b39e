"""Times ddc_k_attention alone (CUDA events, L2 flushed between runs by the size of the tensors).
    python scripts/att_micro.py B L C heads [fmt] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ddim_cold_b200 import _lib  # noqa: E402

B, L, C, H = (int(a) for a in sys.argv[1:5])
fmt = int(sys.argv[5]) if len(sys.argv) > 5 else 1
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 5
lib = _lib.load()
qkv = (torch.randn(B * L, 3 * C, device="cuda") * 1.5).to(torch.float16 if fmt else torch.bfloat16)
out = torch.zeros(B * L, C, device="cuda", dtype=qkv.dtype)
st = torch.cuda.current_stream().cuda_stream
ts = []
for i in range(reps + 2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.check(lib.ddc_k_attention(fmt, qkv.data_ptr(), out.data_ptr(), B, L, C, H, st))
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
d = C // H
print(f"attention B={B} L={L} C={C} heads={H} fmt={fmt} FA={os.environ.get('DDC_ATT_FA', '1')}: "
      f"{min(ts[2:]):.1f} us (best of {reps}), {4.0 * B * H * L * L * d / min(ts[2:]) / 1e6:.0f} TFLOP/s useful")
