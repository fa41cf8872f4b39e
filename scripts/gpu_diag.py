"""Bring-up diagnostics: run each GEMM configuration in a fresh process (a trapped kernel
poisons its CUDA context) and print error structure, not just pass/fail."""
import subprocess
import sys

CASES = [(0, 3, 128, 64, 64), (0, 3, 128, 128, 64), (0, 3, 128, 128, 128), (0, 3, 200, 128, 384), (1, 3, 128, 64, 64),
         (1, 3, 300, 384, 384), (0, 0, 300, 1152, 384), (1, 1, 300, 384, 384), (0, 2, 300, 384, 384), (1, 3, 200, 192, 384),
         (1, 3, 200, 48, 256)]

CHILD = r'''
import sys, torch, ctypes as C
sys.path.insert(0, ".")
from ddim_cold_b200 import _lib
lib = _lib.load()
amode, epi, M, N, K = map(int, sys.argv[1:6])
torch.manual_seed(0)
dev = "cuda"
W = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
bias = torch.randn(N, device=dev) * 0.1
Mp = (M + 127) // 128 * 128
if amode == 0:
    A = torch.zeros(Mp, K, device=dev, dtype=torch.bfloat16); A[:M] = torch.randn(M, K, device=dev).to(torch.bfloat16)
    Ae = A[:M].float(); g = b = None
else:
    A = torch.zeros(Mp, K, device=dev); A[:M] = torch.randn(M, K, device=dev) * 1.7 + 0.3
    g = 1 + 0.1 * torch.randn(K, device=dev); b = 0.1 * torch.randn(K, device=dev)
    Ae = torch.nn.functional.layer_norm(A[:M], (K,), g, b, 1e-5).to(torch.bfloat16).float()
ref = Ae @ W.float().T + bias
if epi == 1: ref = torch.nn.functional.gelu(ref)
out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16 if epi in (0, 1) else torch.float32)
if epi == 2:
    out = torch.randn(M, N, device=dev); ref = ref + out
rc = lib.ddc_k_gemm(amode, epi, M, N, K, A.data_ptr(), g.data_ptr() if amode else None, b.data_ptr() if amode else None,
                    W.data_ptr(), bias.data_ptr(), out.data_ptr(), N, torch.cuda.current_stream().cuda_stream)
if rc: print("rc", rc, lib.ddc_last_error()); sys.exit(1)
torch.cuda.synchronize()
err = (out.float() - ref).abs()
print(f"max err {err.max().item():.4e}  ref max {ref.abs().max().item():.3f}  out max {out.float().abs().max().item():.3f}")
if err.max().item() > 3e-2:
    rb = err.reshape(-1, N)[: (M // 8) * 8].reshape(M // 8, 8, N).amax(dim=(1, 2))
    print("  err by 8-row group (first 16):", [f"{v:.2f}" for v in rb[:16].tolist()])
    cb = err[:, : (N // 8) * 8].reshape(M, N // 8, 8).amax(dim=(0, 2))
    print("  err by 8-col group (first 16):", [f"{v:.2f}" for v in cb[:16].tolist()])
    print("  out[0,:8]", out[0, :8].float().tolist()); print("  ref[0,:8]", ref[0, :8].tolist())
'''

for case in CASES:
    print("== gemm amode=%d epi=%d M=%d N=%d K=%d" % case, flush=True)
    try:
        r = subprocess.run([sys.executable, "-c", CHILD] + [str(c) for c in case], capture_output=True, text=True,
                           timeout=120)
        print(r.stdout.strip())
        if r.returncode != 0:
            print("  exit", r.returncode, r.stderr.strip()[-600:])
    except subprocess.TimeoutExpired:
        print("  TIMEOUT")
