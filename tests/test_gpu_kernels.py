"""GPU: each CUDA kernel through its C-ABI entry point against a plain fp32 torch restatement
of the same op (operands pre-rounded to bf16 where the kernel consumes bf16, so the
tolerance only has to cover fp32 accumulation order and the bf16 rounding of the output)."""
import ctypes as C
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    from ddim_cold_b200 import _lib
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return _lib.load()


def _check(code):
    from ddim_cold_b200 import _lib
    _lib.check(code)


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _pad_rows(t, mult=128):
    m = t.shape[0]
    mp = (m + mult - 1) // mult * mult
    out = torch.zeros((mp,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
    out[:m] = t
    return out


GEMM_SHAPES = [(200, 128, 64), (130, 192, 384), (1000, 1152, 384), (257 * 3, 768, 256), (65 * 40, 384, 192),
               (128, 64, 64), (4097, 384, 384), (66560 // 4, 1152, 384), (40000, 256, 256)]


def _op(t, fmt):
    """Round to the 16-bit operand format; fp16 tensors travel as raw 16-bit words."""
    return t.to(torch.float16 if fmt else torch.bfloat16)


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("epi", [0, 1, 2])
@pytest.mark.parametrize("fmt", [0, 1])
def test_gemm_stream(lib, M, N, K, epi, fmt):
    """qkv / fc1 / head GEMMs: persistent tcgen05 kernel, TMA in, TMA-store (bf16) or direct (fp32) out."""
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N * 3 + K + epi)
    dev = "cuda"
    W = _op(torch.randn(N, K, device=dev, generator=g) * 0.05, fmt)
    bias = torch.randn(N, device=dev, generator=g) * 0.1
    A = _pad_rows(_op(torch.randn(M, K, device=dev, generator=g), fmt))
    ref = A[:M].float() @ W.float().T + bias
    if epi == 1:
        ref = torch.nn.functional.gelu(ref)
    out = torch.full((M + 3, N), 7.0, device=dev, dtype=torch.float32 if epi == 2 else W.dtype)
    before = out.clone()
    _check(lib.ddc_k_gemm(fmt, epi, M, N, K, A.data_ptr(), W.data_ptr(), bias.data_ptr(), out.data_ptr(), N, _stream()))
    torch.cuda.synchronize()
    tol = 3e-3 if epi == 2 else (4e-3 if fmt else 2e-2)          # fp32 output vs 16-bit output rounding
    err = (out[:M].float() - ref).abs().max().item()
    assert err < tol * max(1.0, ref.abs().max().item()), (err, ref.abs().max().item())
    assert torch.equal(out[M:], before[M:]), "rows beyond M were written"


@pytest.mark.parametrize("B,L,C,P", [(3, 65, 384, 192), (2, 257, 256, 48), (1, 626, 384, 192), (5, 17, 64, 48),
                                     (130, 65, 384, 192), (7, 65, 128, 48), (2, 100, 512, 192)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_embed_gemm(lib, B, L, C, P, fmt):
    """Patch-embedding GEMM with the token assembly fused in (gemm2_kernel<R_EMBED>; ViT.py:150,154,199-206) and the
    first LayerNorm: the all-zero cls slot of every image must come out as cls + pos[0] + time[t], the position rows
    wrap around inside 128-row tiles (L = 65, 257, 626, 17, 100 are all coprime to 128), t differs per image."""
    g = torch.Generator(device="cuda").manual_seed(B * 131 + L + C)
    dev = "cuda"
    T = 50
    Kp = (P + 63) // 64 * 64
    rnd = lambda *sh, s=1.0: torch.randn(*sh, device=dev, generator=g) * s
    W = torch.zeros(C, Kp, device=dev)
    W[:, :P] = rnd(C, P, s=0.07)
    W = _op(W, fmt)
    cls, pos, conv_b, time_tab = rnd(C, s=0.3), rnd(L, C, s=0.3), rnd(C, s=0.1), rnd(T, C, s=0.3)
    ln_g, ln_b = 1 + rnd(C, s=0.1), rnd(C, s=0.1)
    M = B * L
    ap = torch.zeros(M, Kp, device=dev)
    ap[:, :P] = rnd(M, P)
    ap.view(B, L, Kp)[:, 0] = 0                                         # cls slots (ddc_k_patchify writes them as zeros)
    ap = _pad_rows(_op(ap, fmt))
    l_idx = torch.arange(L + 127, device=dev) % L
    posb = (pos[l_idx] + torch.where((l_idx == 0)[:, None], cls[None], conv_b[None])).contiguous()
    t_img = torch.randint(0, T, (B,), device=dev, generator=g, dtype=torch.int32)
    x = torch.full((ap.shape[0], C), 5.0, device=dev)
    xn = torch.full((M + 5, C), 3.0, device=dev, dtype=W.dtype)
    _check(lib.ddc_k_embed(fmt, B, L, C, Kp, T, ap.data_ptr(), W.data_ptr(), posb.data_ptr(), time_tab.data_ptr(),
                           t_img.data_ptr(), ln_g.data_ptr(), ln_b.data_ptr(), x.data_ptr(), xn.data_ptr(), _stream()))
    torch.cuda.synchronize()
    tok = (ap[:M].float() @ W.float().T).view(B, L, C)
    tok = tok + torch.cat([(cls + pos[0])[None], pos[1:] + conv_b], dim=0)[None] + time_tab[t_img.long()][:, None]
    x_ref = tok.reshape(M, C)
    xn_ref = torch.nn.functional.layer_norm(x_ref, (C,), ln_g, ln_b, 1e-5)
    assert (x[:M] - x_ref).abs().max().item() < 3e-3 * max(1.0, x_ref.abs().max().item())
    cls_rows = x[:M].view(B, L, C)[:, 0]
    assert (cls_rows - (cls + pos[0] + time_tab[t_img.long()])).abs().max().item() < 1e-5      # no GEMM term at all
    assert (xn[:M].float() - xn_ref).abs().max().item() < (4e-3 if fmt else 2e-2) * max(1.0, xn_ref.abs().max().item())
    assert bool((x[M:] == 5.0).all()) and bool((xn[M:].float() == 3.0).all()), "rows beyond M were written"


@pytest.mark.parametrize("M,C,K", [(200, 64, 64), (130, 384, 384), (1000, 256, 256), (4097, 384, 384), (300, 128, 192),
                                   (66560 // 2, 384, 384), (515, 512, 512)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_gemm_resid_layernorm(lib, M, C, K, fmt):
    """proj / fc2: x += A W^T + b in place (fp32) and the fused LayerNorm of the new row -> bf16."""
    g = torch.Generator(device="cuda").manual_seed(M + C + K)
    dev = "cuda"
    W = _op(torch.randn(C, K, device=dev, generator=g) * 0.05, fmt)
    bias = torch.randn(C, device=dev, generator=g) * 0.1
    A = _pad_rows(_op(torch.randn(M, K, device=dev, generator=g), fmt))
    x = _pad_rows(torch.randn(M, C, device=dev, generator=g) * 1.3 + 0.2)
    x[M:] = 5.0
    ln_g = 1 + 0.1 * torch.randn(C, device=dev, generator=g)
    ln_b = 0.1 * torch.randn(C, device=dev, generator=g)
    xn = torch.full((M + 5, C), 3.0, device=dev, dtype=W.dtype)
    x_ref = x[:M] + A[:M].float() @ W.float().T + bias
    xn_ref = torch.nn.functional.layer_norm(x_ref, (C,), ln_g, ln_b, 1e-5)
    _check(lib.ddc_k_gemm_resid_ln(fmt, M, C, K, A.data_ptr(), W.data_ptr(), bias.data_ptr(), x.data_ptr(), ln_g.data_ptr(),
                                   ln_b.data_ptr(), xn.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert (x[:M] - x_ref).abs().max().item() < 3e-3 * max(1.0, x_ref.abs().max().item())
    assert (xn[:M].float() - xn_ref).abs().max().item() < (4e-3 if fmt else 2e-2) * max(1.0, xn_ref.abs().max().item())
    assert bool((x[M:] == 5.0).all()) and bool((xn[M:].float() == 3.0).all()), "rows beyond M were written"


@pytest.mark.parametrize("M,C", [(200, 128), (130, 384), (1000, 256), (4097, 384), (66560 // 2, 384), (300, 384)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_block_tail(lib, M, C, fmt):
    """Fused proj + LayerNorm2 + fc1 + GELU + fc2 + residuals + next LayerNorm against plain torch.  The kernel only
    normalises: the affine parts are folded into the consuming weights by the caller (here in torch, as ddc_pack_weights
    does on the device), so the reference is the reference's own arithmetic with its LayerNorm affine in place."""
    g = torch.Generator(device="cuda").manual_seed(M + C + fmt)
    dev = "cuda"
    F = torch.nn.functional
    rnd = lambda *sh, s=1.0: torch.randn(*sh, device=dev, generator=g) * s
    Wp, W2 = (_op(rnd(C, C, s=0.05), fmt) for _ in range(2))
    W1f = rnd(C, C, s=0.05)
    bp, b1, b2 = (rnd(C, s=0.1) for _ in range(3))
    g2 = 1 + rnd(C, s=0.1)
    be2 = rnd(C, s=0.1)
    W1 = _op(W1f * g2[None, :], fmt)                       # gamma2 folded into fc1, beta2 into its bias
    b1f = b1 + W1f @ be2
    O = _pad_rows(_op(rnd(M, C), fmt))
    x = _pad_rows(rnd(M, C, s=1.3) + 0.2)
    x[M:] = 5.0
    xn = torch.full((M + 5, C), 3.0, device=dev, dtype=Wp.dtype)
    x1 = x[:M] + O[:M].float() @ Wp.float().T + bp
    h = _op(F.gelu(F.layer_norm(x1, (C,), g2, be2, 1e-5) @ W1f.T + b1), fmt).float()
    x_ref = x1 + h @ W2.float().T + b2
    xn_ref = F.layer_norm(x_ref, (C,), None, None, 1e-5)
    _check(lib.ddc_k_block_tail(fmt, M, C, O.data_ptr(), Wp.data_ptr(), bp.data_ptr(), W1.data_ptr(), b1f.data_ptr(),
                                W2.data_ptr(), b2.data_ptr(), x.data_ptr(), xn.data_ptr(), _stream()))
    torch.cuda.synchronize()
    tol16 = 4e-3 if fmt else 2e-2
    assert (x[:M] - x_ref).abs().max().item() < tol16 * max(1.0, x_ref.abs().max().item())
    assert (xn[:M].float() - xn_ref).abs().max().item() < 2 * tol16 * max(1.0, xn_ref.abs().max().item())
    assert bool((x[M:] == 5.0).all()) and bool((xn[M:].float() == 3.0).all()), "rows beyond M were written"


ATT_SHAPES = [(3, 65, 384, 12), (2, 257, 256, 4), (1, 626, 384, 12), (5, 17, 64, 2), (2, 33, 128, 2), (4, 16, 64, 1),
              (300, 65, 384, 12), (3, 65, 128, 4), (2, 100, 256, 8), (2, 65, 256, 4),
              # the tcgen05 kernel (L > 128): several items per CTA, odd tile counts, remainder rows, exact multiples
              (40, 257, 256, 4), (30, 626, 384, 12), (3, 300, 128, 2), (2, 130, 128, 4), (2, 256, 64, 1), (1, 640, 64, 2),
              (2, 129, 64, 2), (3, 385, 192, 3), (2, 200, 128, 4)]


@pytest.mark.parametrize("B,L,Cdim,heads", ATT_SHAPES)
@pytest.mark.parametrize("fmt", [0, 1])
def test_attention(lib, B, L, Cdim, heads, fmt):
    g = torch.Generator(device="cuda").manual_seed(L + heads)
    d = Cdim // heads
    qkv = _op(torch.randn(B * L, 3 * Cdim, device="cuda", generator=g) * 1.5, fmt)
    out = torch.zeros(B * L, Cdim, device="cuda", dtype=qkv.dtype)
    _check(lib.ddc_k_attention(fmt, qkv.data_ptr(), out.data_ptr(), B, L, Cdim, heads, _stream()))
    torch.cuda.synchronize()
    x = qkv.float().reshape(B, L, 3, heads, d).permute(2, 0, 3, 1, 4)
    q, k, v = x[0], x[1], x[2]
    p = ((q @ k.transpose(-2, -1)) * d ** -0.5).softmax(dim=-1)
    ref = (p @ v).transpose(1, 2).reshape(B * L, Cdim)
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("B,L,Cdim,heads", [(3, 65, 384, 12), (300, 65, 384, 12), (5, 17, 64, 2), (2, 100, 256, 8), (4, 16, 64, 1),
                                            (2, 128, 128, 2), (9, 65, 128, 2)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_attention_flash_tcgen05_short_sequences(lib, B, L, Cdim, heads, fmt, monkeypatch):
    """attention_fa.cuh forced onto short sequences too (DDC_ATT_FA=2): a single, partly filled query tile and a
    single key chunk, masked keys, head pairs of 32 and single heads of 64."""
    monkeypatch.setenv("DDC_ATT_FA", "2")
    test_attention(lib, B, L, Cdim, heads, fmt)


@pytest.mark.parametrize("B,L,Cdim,heads", [(2, 257, 256, 4), (40, 257, 256, 4), (2, 256, 64, 1), (1, 640, 64, 1), (2, 129, 64, 1),
                                            (3, 385, 192, 3), (3, 300, 128, 2)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_attention_flash_tcgen05_head_dim_64(lib, B, L, Cdim, heads, fmt, monkeypatch):
    """Long sequences with head_dim 64 take the mma.sync kernel by default (it measures faster there); DDC_ATT_FA=3
    sends them through attention_fa.cuh: two query tiles per item, an idle second tile when their number is odd,
    remainder rows on the CUDA cores (257 = 2 * 128 + 1)."""
    monkeypatch.setenv("DDC_ATT_FA", "3")
    test_attention(lib, B, L, Cdim, heads, fmt)


@pytest.mark.parametrize("B,L,Cdim,heads", [(2, 257, 256, 4), (1, 626, 384, 12)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_attention_legacy_long_sequence_kernel(lib, B, L, Cdim, heads, fmt, monkeypatch):
    """The cp.async + mma.sync kernel that the tcgen05 one replaced for L > 128 stays selectable (DDC_ATT_FA=0)."""
    monkeypatch.setenv("DDC_ATT_FA", "0")
    test_attention(lib, B, L, Cdim, heads, fmt)


@pytest.mark.parametrize("B,H,W,p", [(3, 64, 64, 8), (2, 64, 64, 4), (1, 200, 200, 8), (2, 16, 16, 4), (2, 32, 48, 4)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_patchify(lib, B, H, W, p, fmt):
    x = torch.randn(B, 3, H, W, device="cuda")
    P = 3 * p * p
    Kp = (P + 63) // 64 * 64
    n = (H // p) * (W // p)
    L = n + 1
    ap = torch.full((B * L, Kp), 9.0, device="cuda", dtype=torch.float16 if fmt else torch.bfloat16)
    _check(lib.ddc_k_patchify(fmt, x.data_ptr(), ap.data_ptr(), B, H, W, p, Kp, _stream()))
    torch.cuda.synchronize()
    ref = x.reshape(B, 3, H // p, p, W // p, p).permute(0, 2, 4, 1, 3, 5).reshape(B, n, P).to(ap.dtype)
    got = ap.reshape(B, L, Kp)
    assert torch.equal(got[:, 1:, :P], ref)
    assert torch.count_nonzero(got[:, 0]) == 0 and torch.count_nonzero(got[:, :, P:]) == 0


@pytest.mark.parametrize("B,H,W,p", [(3, 64, 64, 8), (2, 64, 64, 4), (1, 200, 200, 8), (2, 16, 16, 4)])
@pytest.mark.parametrize("t,k", [(1999, 20), (999, 20), (19, 20), (1, 1)])
def test_head_ddim(lib, B, H, W, p, t, k):
    from ddim_cold_b200 import _lib
    from oracle import ddim_oracle as O
    sh = O.Shape(H, W, p, 64, 1, 2)
    P = 3 * p * p
    L = sh.tokens
    y = torch.randn(B * L, P, device="cuda") * 0.8
    x = torch.randn(B, 3, H, W, device="cuda")
    x_before = x.clone()
    x0 = torch.empty_like(x)
    eps = torch.empty_like(x)
    tab, n = _lib.ddim_table(k, 2000, t)
    _check(lib.ddc_k_head_ddim(y.data_ptr(), P, x.data_ptr(), B, H, W, p, C.byref(tab[0]), x0.data_ptr(),
                               eps.data_ptr(), _stream()))
    torch.cuda.synchronize()
    y_np = y.cpu().numpy().reshape(B, L, P)
    x0_ref = np.clip(O.unpatchify(y_np[:, 1:], sh), -1, 1)
    eps_ref, xn_ref = O.ddim_update(x_before.cpu().numpy().astype(np.float64), x0_ref.astype(np.float64), t, k)
    assert np.array_equal(x0.cpu().numpy(), x0_ref)                      # pure data movement + clamp: bit exact
    assert np.abs(x.cpu().numpy() - xn_ref).max() < 1e-5 * max(1.0, np.abs(xn_ref).max())
    assert np.abs(eps.cpu().numpy() - eps_ref).max() < 1e-5 * max(1.0, np.abs(eps_ref).max())


@pytest.mark.parametrize("N,per_image_draft", [(5, False), (3, True)])
def test_mix_matches_reference_expression_bitwise(lib, N, per_image_draft):
    """Draft->drawing start state (ViT_draft2drawing.py:395-396): noise * sqrt(1 - a) + draft * sqrt(a) with the reference's
    three separate roundings -- bit-identical to the torch CPU expression."""
    a = 1 - math.sqrt(1599 / 2000)
    z = torch.randn(N, 3, 16, 24)
    draft = torch.rand((N, 3, 16, 24) if per_image_draft else (3, 16, 24)) * 2 - 1
    want = z * math.sqrt(1 - a) + draft * math.sqrt(a)
    zc, dc = z.cuda(), draft.cuda()
    out = torch.empty_like(zc)
    _check(lib.ddc_k_mix(zc.data_ptr(), dc.data_ptr(), out.data_ptr(), zc.numel(), dc.numel(), math.sqrt(1 - a), math.sqrt(a),
                         _stream()))
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), want)


@pytest.mark.parametrize("B,Cdim,heads", [(4, 384, 12), (5, 128, 4), (1, 64, 2), (301, 384, 12), (130, 256, 8), (2, 192, 6)])
@pytest.mark.parametrize("fmt", [0, 1])
def test_qkv_attention_fused(lib, B, Cdim, heads, fmt):
    """qkv Linear + attention in one kernel (qkv_attn.cuh) for 65-token images: against fp32 torch, with q/k/v rounded to
    the operand format as the kernel does.  Batches that are not a multiple of four leave part of the last CTA pair empty;
    class-token rows take their q/k/v from the separate class-row GEMM."""
    L, d = 65, 32
    g = torch.Generator(device="cuda").manual_seed(B * 7 + Cdim)
    M = B * L
    xn = _pad_rows(_op(torch.randn(M, Cdim, device="cuda", generator=g), fmt))
    W = _op(torch.randn(3 * Cdim, Cdim, device="cuda", generator=g) * (1.5 / math.sqrt(Cdim)), fmt)
    b = torch.randn(3 * Cdim, device="cuda", generator=g) * 0.2
    out = torch.full((M + 3, Cdim), 7.0, device="cuda", dtype=W.dtype)
    before = out.clone()
    w_hp = torch.empty_like(W)
    b_hp = torch.empty_like(b)
    qkv_cls = torch.empty(B, 3 * Cdim, device="cuda", dtype=W.dtype)
    _check(lib.ddc_k_qkv_attn(fmt, xn.data_ptr(), W.data_ptr(), b.data_ptr(), out.data_ptr(), B, Cdim, heads, w_hp.data_ptr(),
                              b_hp.data_ptr(), qkv_cls.data_ptr(), _stream()))
    torch.cuda.synchronize()
    qkv = _op(xn[:M].float() @ W.float().T + b, fmt).float()
    x = qkv.reshape(B, L, 3, heads, d).permute(2, 0, 3, 1, 4)
    q, k, v = x[0], x[1], x[2]
    pr = ((q @ k.transpose(-2, -1)) * d ** -0.5).softmax(dim=-1)
    ref = (pr @ v).transpose(1, 2).reshape(M, Cdim)
    err = (out[:M].float() - ref).abs().max().item()
    assert err < 2e-2 * max(1.0, ref.abs().max().item()), err
    assert torch.equal(out[M:], before[M:]), "rows beyond M were written"
