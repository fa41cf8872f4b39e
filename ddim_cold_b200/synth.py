"""Deterministic synthetic workload of the sampling path: shapes, weights, noise, draft, FLOP count.

There is no network on the GPU box and the reference's checkpoints are absent, so
benchmarks, the smoke test and the parity tests all use random-init weights of the reference's
architectures.  They travel as a *seed*: every machine regenerates bit-identical numpy arrays with
the reference's key names, shapes and order (ViT.py:169-183) and its init distributions
(ViT.py:184-196).  Pure numpy, no CUDA: ``bench.py`` and the CLI use it for inputs, the oracle and
the golden-fixture generator import the same functions so that fixtures and GPU runs agree on the
weights.  Nothing here computes the model -- that is the CUDA library's job (and, for checking, the
oracle's).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class Shape:
    """Constructor arguments that fix every tensor shape (ViT.py:160-162)."""

    img_h: int = 64
    img_w: int = 64
    patch: int = 8
    embed: int = 256
    depth: int = 3
    heads: int = 4
    in_chans: int = 3
    total_steps: int = 2000

    @property
    def n_patches(self) -> int:  # ViT.py:146
        return (self.img_h // self.patch) * (self.img_w // self.patch)

    @property
    def tokens(self) -> int:  # ViT.py:174 (num_patches + 1)
        return self.n_patches + 1

    @property
    def patch_dim(self) -> int:  # ViT.py:183 (in_chans * patch_size**2)
        return self.in_chans * self.patch * self.patch


S1 = Shape(64, 64, 8, 384, 7, 12)     # 20220822.yaml:10-15, ViT_draft2drawing.py:342
S2 = Shape(64, 64, 4, 256, 6, 4)      # ViT.py:277
S3 = Shape(200, 200, 8, 384, 7, 12)   # declared stand-in for the unrecorded 200x200 model
TINY = Shape(16, 16, 4, 64, 2, 2)     # fixture-sized model (not in the reference)


def state_dict_spec(sh: Shape) -> list[tuple[str, tuple[int, ...]]]:
    """Key names / shapes / order of the reference state_dict (ViT.py:169-183)."""
    C, L, P, p = sh.embed, sh.tokens, sh.patch_dim, sh.patch
    spec = [
        ("cls_token", (1, 1, C)),
        ("pos_embed", (1, L, C)),
        ("patch_embed.proj.weight", (C, sh.in_chans, p, p)),
        ("patch_embed.proj.bias", (C,)),
        ("time_embed.weight", (sh.total_steps, C)),
    ]
    for i in range(sh.depth):
        b = f"blocks.{i}."
        spec += [
            (b + "norm1.weight", (C,)), (b + "norm1.bias", (C,)),
            (b + "attn.qkv.weight", (3 * C, C)), (b + "attn.qkv.bias", (3 * C,)),
            (b + "attn.proj.weight", (C, C)), (b + "attn.proj.bias", (C,)),
            (b + "norm2.weight", (C,)), (b + "norm2.bias", (C,)),
            (b + "mlp.fc1.weight", (C, C)), (b + "mlp.fc1.bias", (C,)),
            (b + "mlp.fc2.weight", (C, C)), (b + "mlp.fc2.bias", (C,)),
        ]
    spec += [("norm.weight", (C,)), ("norm.bias", (C,)), ("head.weight", (P, C)), ("head.bias", (P,))]
    return spec


def synth_state_dict(sh: Shape, seed: int = 0, rich: bool = True, gain: float = 1.0):
    """Deterministic numpy weights of the reference's shapes, identical on every machine.

    Distributions follow the reference init (trunc-normal std .02 for Linear / pos / cls /
    time, ViT.py:184-196; Conv2d left at PyTorch's kaiming-uniform default) and, with
    ``rich``, also randomise what the init leaves at constants (biases, LN affine) so that
    every bias / gain code path is exercised.  ``gain`` scales the 2-D weights ("hot" set,
    SURVEY.md §8c).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    sd = {}
    for name, shape in state_dict_spec(sh):
        if name == "patch_embed.proj.weight":
            bound = 1.0 / math.sqrt(sh.patch_dim)
            a = rng.uniform(-bound, bound, size=shape) * gain
        elif name == "patch_embed.proj.bias":
            bound = 1.0 / math.sqrt(sh.patch_dim)
            a = rng.uniform(-bound, bound, size=shape)
        elif name.endswith("norm1.weight") or name.endswith("norm2.weight") or name == "norm.weight":
            a = 1.0 + (0.1 * rng.standard_normal(shape) if rich else 0.0) * np.ones(shape)
        elif name.endswith(".bias"):
            a = (0.02 * rng.standard_normal(shape)) if rich else np.zeros(shape)
        else:
            a = np.clip(rng.standard_normal(shape), -2.0, 2.0) * 0.02
            if a.ndim == 2 and name != "time_embed.weight":
                a = a * gain
        sd[name] = np.ascontiguousarray(a, dtype=np.float32)
    return sd


def synth_noise(shape, seed: int):
    """Deterministic N(0,1) test input (stands in for ViT.py:224's global-RNG draw)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.standard_normal(shape).astype(np.float32)


def synth_draft(sh: Shape, seed: int = 7):
    """Synthetic draft image in [-1,1]: low-frequency colour field + hard-edged boxes
    (Saved_Models/draft.jpg is absent from the reference)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    yy, xx = np.meshgrid(np.linspace(0, 1, sh.img_h), np.linspace(0, 1, sh.img_w), indexing="ij")
    img = np.stack([np.sin(2 * np.pi * (f[0] * xx + f[1] * yy) + f[2])
                    for f in rng.uniform(0.3, 1.5, size=(3, 3))]) * 0.6
    for _ in range(3):
        y0, x0 = rng.integers(0, sh.img_h // 2), rng.integers(0, sh.img_w // 2)
        h, w = rng.integers(sh.img_h // 8, sh.img_h // 2), rng.integers(sh.img_w // 8, sh.img_w // 2)
        img[:, y0:y0 + h, x0:x0 + w] = rng.uniform(-1, 1, size=(3, 1, 1))
    return np.clip(img, -1, 1).astype(np.float32)


def flops_per_image_step(sh: Shape) -> float:
    """Algorithmic FLOPs of one forward for one image (SURVEY.md §8d formula)."""
    n, P, C, L, h = sh.n_patches, sh.patch_dim, sh.embed, sh.tokens, sh.heads
    d = C // h
    per_block = 2 * L * C * 3 * C + 4 * h * L * L * d + 2 * L * C * C + 4 * L * C * C
    return float(2 * n * P * C + sh.depth * per_block + 2 * L * C * P)
