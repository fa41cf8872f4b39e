"""CPU oracle for the DDIM-COLD sampling hot path -- TEST INFRASTRUCTURE ONLY.

This file is a numpy restatement of the reference algorithm, written to be the
*checker* for the CUDA path.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  The product
(``ddim_cold_b200``) never does; it fails loudly when the CUDA library is missing.

Parity pin: the reference ships no tests or golden vectors.  The oracle is
pinned by running the *reference itself* (``/root/reference/ViT.py`` and
``ViT_draft2drawing.py`` imported unmodified, ``.eval()``, fp32, CPU) in the authoring
container on seeded inputs and committing its outputs under ``tests/golden/``
(``oracle/make_golden.py`` is the generating script).  ``tests/test_oracle_golden.py``
checks this oracle against those fixtures on every run.

Every function cites the reference lines (``/root/reference/...``) it restates.  The
state dict is the reference's own: a mapping ``name -> ndarray`` with the key names and
shapes of ``DiffusionVisionTransformer.state_dict()`` (ViT.py:169-183).

``dtype`` selects the arithmetic: ``np.float32`` mirrors the reference's fp32 tensors,
``np.float64`` gives a higher-precision "truth" used to tell reference rounding noise
from a real deviation.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

try:  # scipy is present in the image; keep a slow exact fallback so the oracle never lies
    from scipy.special import erf as _erf
except Exception:  # pragma: no cover
    _erf = np.vectorize(math.erf)


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


def steps_of(k: int, total_steps: int = 2000, t_start: int | None = None) -> list[int]:
    """The jumped timesteps the loop visits: ``range(T-1, 0, -k)`` (ViT.py:226)."""
    t0 = total_steps - 1 if t_start is None else t_start
    return list(range(t0, 0, -k))


def ddim_scalars(t: int, k: int, total_steps: int = 2000) -> tuple[float, float]:
    """alpha_{t-k}, alpha_t in Python doubles (ViT.py:231-232).

    Note the asymmetric ``+1e-5`` on alpha_t only.  ``math.sqrt`` raises ValueError for
    ``t + 1 - k < 0`` exactly like the reference does mid-loop.
    """
    alpha_tk = 1 - math.sqrt((t + 1 - k) / total_steps)
    alpha_t = 1 - math.sqrt((t + 1) / total_steps) + 1e-5
    return alpha_tk, alpha_t


def ddim_update(x_t: np.ndarray, x0: np.ndarray, t: int, k: int, total_steps: int = 2000):
    """eps and x_{t-k} in the reference's literal operation order (ViT.py:233-234).

    ``x0`` must already be clamped (ViT.py:229).  Scalars are doubles applied to arrays of
    ``x_t.dtype`` (numpy weak-scalar promotion == torch's scalar handling).
    """
    alpha_tk, alpha_t = ddim_scalars(t, k, total_steps)
    noise = (x_t - math.sqrt(alpha_t) * x0) / math.sqrt(1 - alpha_t)
    if alpha_tk == 0.0:
        raise ZeroDivisionError("float division by zero")  # same failure as ViT.py:234
    x_next = math.sqrt(alpha_tk) * (
        x_t / math.sqrt(alpha_t)
        + (math.sqrt((1 - alpha_tk) / alpha_tk) - math.sqrt((1 - alpha_t) / alpha_t)) * noise
    )
    return noise.astype(x_t.dtype, copy=False), x_next.astype(x_t.dtype, copy=False)


def ddim_ab(t: int, k: int, total_steps: int = 2000) -> tuple[float, float, float, float]:
    """Collapsed coefficients: x_{t-k} = A*x_t + B*x0, eps = (x_t - sa*x0) * inv_s1a.

    Algebraically identical to ViT.py:233-234; used to check the packed
    per-step table the CUDA path reads.  Returns (A, B, sqrt(alpha_t), 1/sqrt(1-alpha_t)).
    """
    alpha_tk, alpha_t = ddim_scalars(t, k, total_steps)
    a = math.sqrt(1 - alpha_tk) / math.sqrt(1 - alpha_t)
    b = math.sqrt(alpha_tk) - a * math.sqrt(alpha_t)
    return a, b, math.sqrt(alpha_t), 1.0 / math.sqrt(1 - alpha_t)


# --------------------------------------------------------------------------- forward


def _linear(x, w, b):
    """nn.Linear: x @ w.T + b (ViT.py:79,81,100,102,183)."""
    return x @ w.T + b


def layer_norm(x, g, b, eps: float = 1e-5):
    """nn.LayerNorm(C), biased variance, eps 1e-5 (ViT.py:124,128,182)."""
    mu = x.mean(axis=-1, keepdims=True)
    xc = x - mu
    var = (xc * xc).mean(axis=-1, keepdims=True)
    return xc / np.sqrt(var + x.dtype.type(eps)) * g + b


def gelu_erf(x):
    """nn.GELU() default, approximate='none' (ViT.py:75,80)."""
    return (0.5 * x * (1.0 + _erf(x / math.sqrt(2.0)))).astype(x.dtype, copy=False)


def patchify(x, p: int):
    """im2col for Conv2d(kernel=stride=p): [B,3,H,W] -> [B, n, 3*p*p] with column
    index c*p*p + p1*p + p2, row index ph*Wp + pw (ViT.py:150,154)."""
    B, C, H, W = x.shape
    hp, wp = H // p, W // p
    x = x.reshape(B, C, hp, p, wp, p).transpose(0, 2, 4, 1, 3, 5)
    return x.reshape(B, hp * wp, C * p * p)


def unpatchify(y, sh: Shape):
    """[B, n, p*p*3] -> [B,3,H,W]; channel is the fastest head feature (ViT.py:215-217)."""
    B = y.shape[0]
    hp, wp, p = sh.img_h // sh.patch, sh.img_w // sh.patch, sh.patch
    img = y.reshape(B, hp, wp, p, p, sh.in_chans).transpose(0, 5, 1, 3, 2, 4)
    return np.ascontiguousarray(img).reshape(B, sh.in_chans, sh.img_h, sh.img_w)


def prepare_tokens(sd, sh: Shape, x, t):
    """Patch embed + cls + pos + time embedding (ViT.py:199-206; dropout is identity in
    eval, ViT.py:175,206)."""
    B = x.shape[0]
    w = sd["patch_embed.proj.weight"].reshape(sh.embed, sh.patch_dim)
    tok = patchify(x, sh.patch) @ w.T + sd["patch_embed.proj.bias"]
    cls = np.broadcast_to(sd["cls_token"], (B, 1, sh.embed))
    tok = np.concatenate([cls, tok], axis=1)
    time_step = sd["time_embed.weight"][np.asarray(t).reshape(B)][:, None, :]
    return tok + sd["pos_embed"] + time_step


def attention(sd, sh: Shape, i: int, x):
    """Attention.forward (ViT.py:105-117); attn_drop / proj_drop are identity in eval."""
    B, L, C = x.shape
    h, d = sh.heads, C // sh.heads
    qkv = _linear(x, sd[f"blocks.{i}.attn.qkv.weight"], sd[f"blocks.{i}.attn.qkv.bias"])
    qkv = qkv.reshape(B, L, 3, h, d).transpose(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    s = (q @ k.transpose(0, 1, 3, 2)) * x.dtype.type(d ** -0.5)
    s = s - s.max(axis=-1, keepdims=True)
    e = np.exp(s)
    p = e / e.sum(axis=-1, keepdims=True)
    o = (p @ v).transpose(0, 2, 1, 3).reshape(B, L, C)
    return _linear(o, sd[f"blocks.{i}.attn.proj.weight"], sd[f"blocks.{i}.attn.proj.bias"])


def mlp(sd, i: int, x):
    """Mlp.forward with hidden == C because mlp_ratio defaults to 1.0 (ViT.py:84-90,161)."""
    hdn = gelu_erf(_linear(x, sd[f"blocks.{i}.mlp.fc1.weight"], sd[f"blocks.{i}.mlp.fc1.bias"]))
    return _linear(hdn, sd[f"blocks.{i}.mlp.fc2.weight"], sd[f"blocks.{i}.mlp.fc2.bias"])


def block(sd, sh: Shape, i: int, x):
    """Pre-LN block; DropPath is identity in eval (ViT.py:132-138)."""
    x = x + attention(sd, sh, i, layer_norm(x, sd[f"blocks.{i}.norm1.weight"], sd[f"blocks.{i}.norm1.bias"]))
    x = x + mlp(sd, i, layer_norm(x, sd[f"blocks.{i}.norm2.weight"], sd[f"blocks.{i}.norm2.bias"]))
    return x


def cast_sd(sd, dtype):
    return {k: np.asarray(v, dtype=dtype) for k, v in sd.items()}


def forward(sd, sh: Shape, x, t, dtype=np.float32):
    """DiffusionVisionTransformer.forward (ViT.py:208-218): unclamped x0 prediction."""
    if next(iter(sd.values())).dtype != dtype:
        sd = cast_sd(sd, dtype)
    x = np.asarray(x, dtype=dtype)
    tok = prepare_tokens(sd, sh, x, t)
    for i in range(sh.depth):
        tok = block(sd, sh, i, tok)
    tok = layer_norm(tok, sd["norm.weight"], sd["norm.bias"])
    y = _linear(tok, sd["head.weight"], sd["head.bias"])
    return unpatchify(y[:, 1:, :], sh)


# --------------------------------------------------------------------------- loops


def denoise_loop(sd, sh: Shape, x_T, k: int, t_start: int | None = None, dtype=np.float32,
                 taps: bool = False, teacher=None):
    """The DDIM loop body shared by sampler / diffusion_sequence / draft->drawing
    (ViT.py:226-235, ViT_draft2drawing.py:399-407).

    Returns ``(x0_last, trace)``.  With ``taps`` the trace holds, per step, the x_t fed in,
    the clamped x0, the derived eps and x_{t-k}.  ``teacher`` (a list of x_t arrays) forces
    each step's input instead of the running state (teacher-forced parity).
    """
    sd = cast_sd(sd, dtype)
    x = np.asarray(x_T, dtype=dtype)
    B = x.shape[0]
    trace = []
    x0 = None
    for j, t in enumerate(steps_of(k, sh.total_steps, t_start)):
        if teacher is not None:
            x = np.asarray(teacher[j], dtype=dtype)
        x0 = np.clip(forward(sd, sh, x, np.full((B,), t, dtype=np.int64), dtype), -1, 1)
        eps, x_next = ddim_update(x, x0, t, k, sh.total_steps)
        if taps:
            trace.append({"t": t, "x_t": x, "x0": x0, "eps": eps, "x_next": x_next})
        x = x_next
    return x0, trace


def sampler(sd, sh: Shape, x_T, k: int = 10, dtype=np.float32):
    """DiffusionVisionTransformer.sampler given its initial noise (ViT.py:220-237):
    returns ``(x0_last + 1) / 2`` -- the last step's clamped prediction, not x_{t-k}."""
    x0, _ = denoise_loop(sd, sh, x_T, k, dtype=dtype)
    return (x0 + 1) / 2


def diffusion_sequence(sd, sh: Shape, x_T, k: int = 100, dtype=np.float32):
    """ViT.py:239-256: element 0 is (x_T+1)/2 unclamped, then every step's (x0+1)/2."""
    _, trace = denoise_loop(sd, sh, x_T, k, dtype=dtype, taps=True)
    return [(np.asarray(x_T, dtype=dtype) + 1) / 2] + [(s["x0"] + 1) / 2 for s in trace]


def cold_sequence(sd, sh: Shape, x_init, dtype=np.float32):
    """cold_sampler / cold_diffusion_sequence loop (ViT_draft2drawing.py:272-288,
    298-308): t = 6..1, x <- clamp(forward(x, t)).  ``x_init`` is the spatially constant
    expansion of normal(0,1,(N,3)) (ViT_draft2drawing.py:265,295)."""
    sd = cast_sd(sd, dtype)
    x = np.asarray(x_init, dtype=dtype)
    B = x.shape[0]
    out = [(x + 1) / 2]
    for t in range(6, 0, -1):
        x = np.clip(forward(sd, sh, x, np.full((B,), t, dtype=np.int64), dtype), -1, 1)
        out.append((x + 1) / 2)
    return out


def d2d_init(noise, draft, t_start: int, total_steps: int = 2000):
    """Draft->drawing partial noising (ViT_draft2drawing.py:395-396): alpha has no +1 and
    no +1e-5 here."""
    alpha = 1 - math.sqrt(t_start / total_steps)
    return noise * math.sqrt(1 - alpha) + draft * math.sqrt(alpha)


def d2d(sd, sh: Shape, noise, draft, t_start: int, k: int = 10, dtype=np.float32):
    """Draft->drawing loop (ViT_draft2drawing.py:391-409), batched over noise draws."""
    x = d2d_init(np.asarray(noise, dtype=dtype), np.asarray(draft, dtype=dtype), t_start, sh.total_steps)
    x0, _ = denoise_loop(sd, sh, x.astype(dtype), k, t_start=t_start, dtype=dtype)
    return (x0 + 1) / 2


# --------------------------------------------------------------------------- weights


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
