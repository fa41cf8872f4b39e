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
import numpy as np

try:  # scipy is present in the image; keep a slow exact fallback so the oracle never lies
    from scipy.special import erf as _erf
except Exception:  # pragma: no cover
    _erf = np.vectorize(math.erf)


# Shapes and the deterministic synthetic weights / inputs live in the product package (pure numpy, no
# model arithmetic) so that bench.py and the CLI need not import the checker; re-exported here.
from ddim_cold_b200.synth import (S1, S2, S3, TINY, Shape, flops_per_image_step, state_dict_spec,  # noqa: E402,F401
                                  synth_draft, synth_noise, synth_state_dict)


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


def synth_eval_batches(sh, n_batches=3, batch=6):
    """Synthetic (noisy_img, img, t) batches for the trainer-side evaluate() fixture (multi_gpu_trainer.py:36):
    images in [-1, 1], noised to the sqrt-schedule level of their own t; travels as seeds."""
    out = []
    for i in range(n_batches):
        img = np.clip(synth_noise((batch, 3, sh.img_h, sh.img_w), seed=400 + i) * 0.6, -1, 1).astype(np.float32)
        noise = synth_noise((batch, 3, sh.img_h, sh.img_w), seed=500 + i)
        t = (np.arange(batch) * 397 + 131 * i + 5) % sh.total_steps
        alpha = (1 - np.sqrt((t + 1) / sh.total_steps)).astype(np.float32)[:, None, None, None]
        noisy = (np.sqrt(alpha) * img + np.sqrt(1 - alpha) * noise).astype(np.float32)
        out.append((noisy, img, t.astype(np.int64)))
    return out



def smooth_l1_mean(pred, target):
    """F.smooth_l1_loss(pred, target) with its defaults (beta = 1, mean), in float64 (multi_gpu_trainer.py:42)."""
    d = np.abs(np.asarray(pred, dtype=np.float64) - np.asarray(target, dtype=np.float64))
    return float(np.where(d < 1, 0.5 * d * d, d - 0.5).mean())

