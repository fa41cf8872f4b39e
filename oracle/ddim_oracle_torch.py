"""Torch-CPU port of the reference sampler -- TEST / BASELINE INFRASTRUCTURE ONLY.

Same role and same rules as ``oracle/ddim_oracle.py`` (read its header): this is the
checker and the *timed CPU baseline*, never the product.  It exists because the reference
(`/root/reference`) does not travel to the GPU box, and a numpy oracle would not be a fair
stand-in for the reference's CPU speed: the reference runs on ATen's CPU kernels (oneDNN /
MKL, all host threads), so the timed port calls exactly those library ops --
``F.conv2d``, ``F.linear``, ``F.layer_norm``, ``F.gelu``, ``softmax``, ``matmul`` -- in the
same order and with the same per-step scalar arithmetic as ViT.py:199-237, including the
per-step ``torch.tensor([t]*N)`` construction (ViT.py:227).  It is a functional restatement
over a plain state dict, not a copy of the reference's module classes.

Pinned against the imported reference by ``oracle/make_golden.py`` (bit-for-bit on CPU for
the same torch build) and against the committed fixtures by ``tests/test_oracle_golden.py``.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .ddim_oracle import Shape


def to_torch_sd(sd, dtype=torch.float32):
    return {k: torch.as_tensor(v).to(dtype) for k, v in sd.items()}


def forward(sd, sh: Shape, x: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
    """ViT.py:199-218 on library ops (eval semantics: every Dropout/DropPath is identity)."""
    B = x.shape[0]
    C, h = sh.embed, sh.heads
    tok = F.conv2d(x, sd["patch_embed.proj.weight"], sd["patch_embed.proj.bias"], stride=sh.patch)
    tok = tok.flatten(2).transpose(1, 2)                                     # ViT.py:154
    tok = torch.cat((sd["cls_token"].expand(B, -1, -1), tok), dim=1)         # ViT.py:202-203
    tok = tok + sd["pos_embed"] + F.embedding(t, sd["time_embed.weight"]).unsqueeze(1)  # :204-205
    for i in range(sh.depth):
        p = f"blocks.{i}."
        y = F.layer_norm(tok, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], 1e-5)
        L = y.shape[1]
        qkv = F.linear(y, sd[p + "attn.qkv.weight"], sd[p + "attn.qkv.bias"])
        qkv = qkv.reshape(B, L, 3, h, C // h).permute(2, 0, 3, 1, 4)         # ViT.py:107
        q, k, v = qkv[0], qkv[1], qkv[2]
        attn = (q @ k.transpose(-2, -1)) * ((C // h) ** -0.5)                # ViT.py:110
        attn = attn.softmax(dim=-1)
        y = (attn @ v).transpose(1, 2).reshape(B, L, C)                      # ViT.py:114
        tok = tok + F.linear(y, sd[p + "attn.proj.weight"], sd[p + "attn.proj.bias"])
        y = F.layer_norm(tok, (C,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], 1e-5)
        y = F.gelu(F.linear(y, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"]))
        tok = tok + F.linear(y, sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])
    tok = F.layer_norm(tok, (C,), sd["norm.weight"], sd["norm.bias"], 1e-5)
    y = F.linear(tok, sd["head.weight"], sd["head.bias"])[:, 1:, :]          # ViT.py:213-214
    hp, wp, ps = sh.img_h // sh.patch, sh.img_w // sh.patch, sh.patch
    img = y.reshape(-1, hp, wp, ps, ps, sh.in_chans).permute(0, 5, 1, 3, 2, 4).contiguous()
    return img.view(-1, sh.in_chans, sh.img_h, sh.img_w)                     # ViT.py:215-217


@torch.no_grad()
def denoise_loop(sd, sh: Shape, x_T: torch.Tensor, k: int, t_start: int | None = None,
                 taps: bool = False, teacher=None):
    """ViT.py:226-235 (and ViT_draft2drawing.py:399-407 via ``t_start``)."""
    x = x_T
    N = x.shape[0]
    T = sh.total_steps
    trace = []
    x0 = None
    t0 = T - 1 if t_start is None else t_start
    for j, t in enumerate(range(t0, 0, -k)):
        if teacher is not None:
            x = teacher[j]
        steps = torch.tensor([t] * N).long()
        x0 = torch.clamp(forward(sd, sh, x, steps), -1, 1)
        alpha_tk = 1 - math.sqrt((t + 1 - k) / T)
        alpha_t = 1 - math.sqrt((t + 1) / T) + 1e-5
        noise = (x - math.sqrt(alpha_t) * x0) / math.sqrt(1 - alpha_t)
        x_next = math.sqrt(alpha_tk) * (
            x / math.sqrt(alpha_t)
            + (math.sqrt((1 - alpha_tk) / alpha_tk) - math.sqrt((1 - alpha_t) / alpha_t)) * noise)
        if taps:
            trace.append({"t": t, "x_t": x, "x0": x0, "eps": noise, "x_next": x_next})
        x = x_next
    return x0, trace


@torch.no_grad()
def sampler(sd, sh: Shape, x_T: torch.Tensor, k: int = 10) -> torch.Tensor:
    """ViT.py:220-237 given the initial noise: (last clamped x0 + 1) / 2."""
    x0, _ = denoise_loop(sd, sh, x_T, k)
    return (x0 + 1) / 2
