"""Clock-stamp timeline of CTA 0 of the fused qkv + attention kernel (tuning aid; see qa_stamp in csrc/qkv_attn.cuh).
    python scripts/qa_trace.py [B] [C] [heads]
Prints, per head-pair iteration, when each role passed its stamps (cycles since the first stamp), then the kernel time."""
import ctypes as C
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ddim_cold_b200 import _lib  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
Cd = int(sys.argv[2]) if len(sys.argv) > 2 else 384
H = int(sys.argv[3]) if len(sys.argv) > 3 else 12
fmt = 1
lib = _lib.load()
L, M = 65, B * 65
Mp = (M + 127) // 128 * 128
xn = torch.zeros(Mp, Cd, device="cuda", dtype=torch.float16)
xn[:M] = torch.randn(M, Cd, device="cuda").half()
W = (torch.randn(3 * Cd, Cd, device="cuda") * (1.5 / math.sqrt(Cd))).half()
b = torch.randn(3 * Cd, device="cuda") * 0.2
out = torch.zeros(M, Cd, device="cuda", dtype=torch.float16)
w_hp, b_hp = torch.empty_like(W), torch.empty_like(b)
cls = torch.empty(B, 3 * Cd, device="cuda", dtype=torch.float16)
trace = torch.zeros(6 * 32 * 8, device="cuda", dtype=torch.int64)
st = torch.cuda.current_stream().cuda_stream
lib.ddc_dbg_qkv_attn_trace.argtypes = [C.c_void_p]


def run():
    _lib.check(lib.ddc_k_qkv_attn(fmt, xn.data_ptr(), W.data_ptr(), b.data_ptr(), out.data_ptr(), B, Cd, H, w_hp.data_ptr(),
                                  b_hp.data_ptr(), cls.data_ptr(), st))


for _ in range(3):
    run()
torch.cuda.synchronize()
lib.ddc_dbg_qkv_attn_trace(C.c_void_p(trace.data_ptr()))
run()
torch.cuda.synchronize()
lib.ddc_dbg_qkv_attn_trace(C.c_void_p(0))
t = trace.cpu().reshape(6, 32, 8)
t0 = int(t[t > 0].min())
if os.environ.get("DDC_QKV_ATTN", "1") == "2":        # warp-level attention inside the fused kernel
    names = {0: ["mma:acc free", "mma:first W", "mma:issued"], 1: ["epi:acc full", "epi:tiles free", "epi:drained", "epi:ready", "epi:cls done"],
             2: ["att:ready", "att:attended", "att:stored"]}
else:                                                  # tcgen05 attention
    names = {0: ["mma:acc free", "mma:first W", "mma:issued"], 1: ["epi:acc full", "epi:tiles free", "epi:drained", "epi:c0", "epi:c1", "epi:c2"],
             4: ["cls:tiles free", "cls:drained", "cls:done", "cls:c3", "cls:c4", "cls:c5"],
             3: ["-", "S issued"],
             2: ["smx:S full", "smx:P written", "smx:O full", "smx:stored"]}
n_it = 2 * (Cd // 64) + 2
for it in range(n_it):
    parts = []
    for role in names:
        for ev, nm in enumerate(names[role]):
            v = int(t[role, it, ev])
            parts.append(f"{nm}={v - t0 if v else -1}")
    print(f"it {it:2d}: " + "  ".join(parts))
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    run()
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print(f"repack + class-row GEMM + fused kernel: best {min(ts):.1f} us")
