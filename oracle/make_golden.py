"""Generate tests/golden/*.npz by running the UNMODIFIED reference on CPU.

Run in the authoring container only (``/root/reference`` does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden

What it does
  * imports ``/root/reference/ViT.py`` and ``ViT_draft2drawing.py`` as they are (the latter
    needs a stub ``turtle`` module because tkinter is absent -- SURVEY.md §8c),
  * builds the reference model, ``.eval()``, fp32, and loads the deterministic numpy weights
    of ``oracle.ddim_oracle.synth_state_dict`` with ``strict=True`` (so the weights travel as
    a seed, not as a 28 MB file),
  * records the reference's outputs for forward / sampler / diffusion_sequence /
    cold_sampler / cold_diffusion_sequence and for a restated draft->drawing loop that calls
    the reference's own ``forward`` (that loop lives in a ``__main__`` block,
    ViT_draft2drawing.py:391-409, and cannot be imported),
  * checks the numpy oracle (fp32 and fp64) and the torch port against those outputs and
    stores the observed deviations in ``tests/golden/PIN.json``.
"""
from __future__ import annotations

import contextlib
import io
import json
import math
import os
import sys
import types

import numpy as np
import torch

from . import ddim_oracle as O
from . import ddim_oracle_torch as OT

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def import_reference():
    sys.dont_write_bytecode = True
    sys.path.insert(0, REF)
    turtle = types.ModuleType("turtle")
    turtle.mode = None
    sys.modules.setdefault("turtle", turtle)
    import ViT  # noqa: E402
    import ViT_draft2drawing as D2D  # noqa: E402
    return ViT, D2D


def ref_model(mod, sh: O.Shape, sd_np):
    m = mod.DiffusionVisionTransformer(img_size=[sh.img_h, sh.img_w], total_steps=sh.total_steps,
                                       patch_size=sh.patch, embed_dim=sh.embed, depth=sh.depth,
                                       num_heads=sh.heads)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd_np.items()}, strict=True)
    return m.eval()


def quiet(fn, *a, **kw):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **kw)


def maxabs(a, b):
    return float(np.max(np.abs(np.asarray(a, dtype=np.float64) - np.asarray(b, dtype=np.float64))))


def ref_trace(model, x_T, k, t_start=None):
    """Instrumented restatement of ViT.py:226-235 around the reference's own forward."""
    T = model.total_steps
    x = x_T.clone()
    N = x.shape[0]
    out = []
    t0 = T - 1 if t_start is None else t_start
    with torch.no_grad():
        for t in range(t0, 0, -k):
            steps = torch.tensor([t] * N).long()
            x0 = torch.clamp(model.forward(x, steps), -1, 1)
            a_tk = 1 - math.sqrt((t + 1 - k) / T)
            a_t = 1 - math.sqrt((t + 1) / T) + 1e-5
            eps = (x - math.sqrt(a_t) * x0) / math.sqrt(1 - a_t)
            xn = math.sqrt(a_tk) * (x / math.sqrt(a_t)
                                    + (math.sqrt((1 - a_tk) / a_tk) - math.sqrt((1 - a_t) / a_t)) * eps)
            out.append({"t": t, "x_t": x, "x0": x0, "eps": eps, "x_next": xn})
            x = xn
    return out


def gen_tiny(ViT, D2D, record):
    # ------------------------------------------------------------------ TINY: whole API
    sh = O.TINY
    sd = O.synth_state_dict(sh, seed=11, rich=True, gain=4.0)   # gain 4 -> clamp is exercised
    m = ref_model(ViT, sh, sd)
    md = ref_model(D2D, sh, sd)
    fix = {}
    # the two class copies share one state_dict layout
    assert list(m.state_dict().keys()) == [k for k, _ in O.state_dict_spec(sh)]
    assert list(md.state_dict().keys()) == list(m.state_dict().keys())

    # forward at assorted t, with per-sample-different t (forward() allows it, ViT.py:204)
    x = torch.from_numpy(O.synth_noise((5, 3, sh.img_h, sh.img_w), seed=1))
    t = torch.tensor([1999, 1000, 19, 1, 0])
    with torch.no_grad():
        y = m.forward(x, t)
    fix["fwd_x"], fix["fwd_t"], fix["fwd_y"] = x.numpy(), t.numpy(), y.numpy()
    record("tiny.forward.np32", maxabs(O.forward(sd, sh, x.numpy(), t.numpy(), np.float32), y))
    record("tiny.forward.np64", maxabs(O.forward(sd, sh, x.numpy(), t.numpy(), np.float64), y))
    record("tiny.forward.torchport", maxabs(OT.forward(OT.to_torch_sd(sd), sh, x, t), y))

    # sampler: same global-RNG draw as ViT.py:224
    for k in (100, 500, 2000):
        torch.manual_seed(123)
        x_T = torch.normal(0, 1, (3, 3, sh.img_h, sh.img_w))
        torch.manual_seed(123)
        img = quiet(m.sampler, "cpu", k, 3)
        fix[f"samp{k}_xT"], fix[f"samp{k}_img"] = x_T.numpy(), img.numpy()
        tr = ref_trace(m, x_T, k)
        assert torch.equal((tr[-1]["x0"] + 1) / 2, img), "instrumented loop != reference sampler"
        record(f"tiny.sampler.k{k}.np32", maxabs(O.sampler(sd, sh, x_T.numpy(), k, np.float32), img))
        record(f"tiny.sampler.k{k}.np64", maxabs(O.sampler(sd, sh, x_T.numpy(), k, np.float64), img))
        record(f"tiny.sampler.k{k}.torchport", maxabs(OT.sampler(OT.to_torch_sd(sd), sh, x_T, k), img))
        if k == 100:
            for key in ("x_t", "x0", "eps", "x_next"):
                fix[f"trace100_{key}"] = np.stack([s[key].numpy() for s in tr])
            # teacher-forced oracle check per step
            _, otr = O.denoise_loop(sd, sh, x_T.numpy(), k, taps=True, teacher=[s["x_t"].numpy() for s in tr])
            record("tiny.trace.k100.x0.np32", max(maxabs(a["x0"], b["x0"]) for a, b in zip(otr, tr)))
            record("tiny.trace.k100.eps.np32", max(maxabs(a["eps"], b["eps"]) for a, b in zip(otr, tr)))
            record("tiny.trace.k100.xnext.np32", max(maxabs(a["x_next"], b["x_next"]) for a, b in zip(otr, tr)))

    # bad k raises exactly like the reference
    for k in (3, 7, 30, 4000):
        try:
            quiet(m.sampler, "cpu", k, 1)
            raise AssertionError(f"reference accepted k={k}")
        except ValueError as e:
            assert "math domain error" in str(e)
        try:
            O.sampler(sd, sh, np.zeros((1, 3, sh.img_h, sh.img_w), np.float32), k)
            raise AssertionError
        except ValueError:
            pass
    fix["bad_k"] = np.array([3, 7, 30, 4000])

    # diffusion_sequence (ViT.py:239-256)
    torch.manual_seed(5)
    x_T = torch.normal(0, 1, (2, 3, sh.img_h, sh.img_w))
    torch.manual_seed(5)
    with torch.no_grad():
        seq = quiet(m.diffusion_sequence, "cpu", 400, 2)
    fix["seq400_xT"], fix["seq400_imgs"] = x_T.numpy(), torch.stack(seq).numpy()
    oseq = O.diffusion_sequence(sd, sh, x_T.numpy(), 400)
    record("tiny.diffusion_sequence.k400.np32", maxabs(np.stack(oseq), fix["seq400_imgs"]))

    # cold sampler / cold sequence (ViT_draft2drawing.py:259-309)
    torch.manual_seed(9)
    z = torch.normal(0, 1, (4, 3))
    x_init = z.unsqueeze(2).unsqueeze(3).expand(-1, -1, sh.img_h, sh.img_w)
    torch.manual_seed(9)
    cimg = quiet(md.cold_sampler, "cpu", 10, 4)
    torch.manual_seed(9)
    cseq = quiet(md.cold_diffusion_sequence, "cpu", 4)
    fix["cold_z"], fix["cold_img"], fix["cold_seq"] = z.numpy(), cimg.numpy(), torch.stack(cseq).numpy()
    oc = O.cold_sequence(sd, sh, x_init.numpy())
    record("tiny.cold_sampler.np32", maxabs(oc[-1], cimg))
    record("tiny.cold_sequence.np32", maxabs(np.stack(oc), fix["cold_seq"]))

    # draft -> drawing (restated loop around the reference forward)
    draft = torch.from_numpy(O.synth_draft(sh))
    noise = torch.from_numpy(O.synth_noise((2, 3, sh.img_h, sh.img_w), seed=21))
    for t_start in (1599, 1999):
        alpha = 1 - math.sqrt(t_start / 2000)
        x = noise * math.sqrt(1 - alpha) + draft * math.sqrt(alpha)
        tr = ref_trace(md, x, 10, t_start=t_start)
        img = (tr[-1]["x0"] + 1) / 2
        fix[f"d2d{t_start}_img"] = img.numpy()
        record(f"tiny.d2d.{t_start}.np32", maxabs(O.d2d(sd, sh, noise.numpy(), draft.numpy(), t_start), img))
    fix["d2d_noise_seed"] = np.array(21)
    np.savez_compressed(os.path.join(OUT, "tiny.npz"), **fix)


def gen_real(ViT, D2D, record):
    # ------------------------------------------------------------------ S1 / S2: real shapes
    for tag, sh, wseed in (("s1", O.S1, 0), ("s2", O.S2, 1)):
        sd = O.synth_state_dict(sh, seed=wseed, rich=True)
        m = ref_model(ViT, sh, sd)
        fix = {}
        torch.manual_seed(123)
        x_T = torch.normal(0, 1, (2, 3, sh.img_h, sh.img_w))
        torch.manual_seed(123)
        img = quiet(m.sampler, "cpu", 100, 2)
        tr = ref_trace(m, x_T, 100)
        assert torch.equal((tr[-1]["x0"] + 1) / 2, img)
        fix["xT"], fix["samp100_img"] = x_T.numpy(), img.numpy()
        # teacher-forced taps for sample 0 at a handful of steps (keeps the fixture small)
        idx = [0, 1, 9, 18, 19]
        fix["tap_idx"] = np.array(idx)
        fix["tap_t"] = np.array([tr[i]["t"] for i in idx])
        for key in ("x_t", "x0", "eps", "x_next"):
            fix[f"tap_{key}"] = np.stack([tr[i][key][:1].numpy() for i in idx])
        record(f"{tag}.sampler.k100.np32", maxabs(O.sampler(sd, sh, x_T.numpy(), 100, np.float32), img))
        record(f"{tag}.sampler.k100.np64", maxabs(O.sampler(sd, sh, x_T.numpy(), 100, np.float64), img))
        record(f"{tag}.sampler.k100.torchport", maxabs(OT.sampler(OT.to_torch_sd(sd), sh, x_T, 100), img))
        _, otr = O.denoise_loop(sd, sh, x_T.numpy(), 100, taps=True, teacher=[s["x_t"].numpy() for s in tr])
        record(f"{tag}.trace.x0.np32", max(maxabs(a["x0"], b["x0"]) for a, b in zip(otr, tr)))
        record(f"{tag}.trace.eps.np32", max(maxabs(a["eps"], b["eps"]) for a, b in zip(otr, tr)))
        _, otr64 = O.denoise_loop(sd, sh, x_T.numpy(), 100, dtype=np.float64, taps=True,
                                  teacher=[s["x_t"].numpy() for s in tr])
        record(f"{tag}.trace.x0.np64", max(maxabs(a["x0"], b["x0"]) for a, b in zip(otr64, tr)))
        record(f"{tag}.trace.eps.np64", max(maxabs(a["eps"], b["eps"]) for a, b in zip(otr64, tr)))
        fix["clamp_frac"] = np.array(np.mean([float((s["x0"].abs() >= 1).float().mean()) for s in tr]))
        record(f"{tag}.clamp_frac", float(fix["clamp_frac"]))
        np.savez_compressed(os.path.join(OUT, f"{tag}.npz"), **fix)


def gen_s3(ViT, D2D, record):
    # ------------------------------------------------------------------ S3: the declared 200x200 stand-in (L = 626)
    # Forward at three timesteps (one image each, different t per image -- forward() allows it, ViT.py:204) and a
    # short free-running sampler, k = 500 (4 steps), N = 1.
    sh = O.S3
    sd = O.synth_state_dict(sh, seed=2, rich=True)
    m = ref_model(ViT, sh, sd)
    fix = {}
    x = torch.from_numpy(O.synth_noise((3, 3, sh.img_h, sh.img_w), seed=41))
    t = torch.tensor([1999, 999, 19])
    with torch.no_grad():
        y = m.forward(x, t)
    fix["fwd_x"], fix["fwd_t"], fix["fwd_y"] = x.numpy(), t.numpy(), y.numpy()
    record("s3.forward.np32", maxabs(O.forward(sd, sh, x.numpy(), t.numpy(), np.float32), y))
    record("s3.forward.torchport", maxabs(OT.forward(OT.to_torch_sd(sd), sh, x, t), y))
    torch.manual_seed(77)
    x_T = torch.normal(0, 1, (1, 3, sh.img_h, sh.img_w))
    torch.manual_seed(77)
    img = quiet(m.sampler, "cpu", 500, 1)
    tr = ref_trace(m, x_T, 500)
    assert torch.equal((tr[-1]["x0"] + 1) / 2, img)
    fix["xT"], fix["samp500_img"] = x_T.numpy(), img.numpy()
    fix["tap_t"] = np.array([s["t"] for s in tr])
    for key in ("x_t", "x0", "eps"):
        fix[f"tap_{key}"] = np.stack([s[key].numpy() for s in tr])
    record("s3.sampler.k500.np32", maxabs(O.sampler(sd, sh, x_T.numpy(), 500, np.float32), img))
    record("s3.sampler.k500.torchport", maxabs(OT.sampler(OT.to_torch_sd(sd), sh, x_T, 500), img))
    np.savez_compressed(os.path.join(OUT, "s3.npz"), **fix)


def gen_fused(ViT, D2D, record):
    # ------------------------------------------------------------------ FUSED: 32x32 / p4 / C128 / depth 2 / 4 heads
    # L = 65 tokens, head_dim 32, embed_dim % 128 == 0: on the GPU this shape takes the fused block tail, the CTA-pair
    # GEMMs and the 65-token attention -- S1's code path at fixture size.  Every loop of the class API is recorded
    # from the reference here (TINY covers the same API on the separate-kernel path).
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    sd = O.synth_state_dict(sh, seed=3, rich=True, gain=2.0)
    m = ref_model(ViT, sh, sd)
    md = ref_model(D2D, sh, sd)
    fix = {}
    torch.manual_seed(15)
    x_T = torch.normal(0, 1, (3, 3, sh.img_h, sh.img_w))
    torch.manual_seed(15)
    with torch.no_grad():
        seq = quiet(m.diffusion_sequence, "cpu", 400, 3)
    fix["seq400_xT"], fix["seq400_imgs"] = x_T.numpy(), torch.stack(seq).numpy()
    record("fused.diffusion_sequence.k400.np32", maxabs(np.stack(O.diffusion_sequence(sd, sh, x_T.numpy(), 400)), fix["seq400_imgs"]))
    torch.manual_seed(15)
    img = quiet(m.sampler, "cpu", 100, 3)
    fix["samp100_img"] = img.numpy()
    record("fused.sampler.k100.np32", maxabs(O.sampler(sd, sh, x_T.numpy(), 100, np.float32), img))
    torch.manual_seed(19)
    z = torch.normal(0, 1, (4, 3))
    x_init = z.unsqueeze(2).unsqueeze(3).expand(-1, -1, sh.img_h, sh.img_w)
    torch.manual_seed(19)
    cimg = quiet(md.cold_sampler, "cpu", 10, 4)
    torch.manual_seed(19)
    cseq = quiet(md.cold_diffusion_sequence, "cpu", 4)
    fix["cold_z"], fix["cold_img"], fix["cold_seq"] = z.numpy(), cimg.numpy(), torch.stack(cseq).numpy()
    oc = O.cold_sequence(sd, sh, x_init.numpy())
    record("fused.cold_sampler.np32", maxabs(oc[-1], cimg))
    record("fused.cold_sequence.np32", maxabs(np.stack(oc), fix["cold_seq"]))
    draft = torch.from_numpy(O.synth_draft(sh))
    noise = torch.from_numpy(O.synth_noise((3, 3, sh.img_h, sh.img_w), seed=23))
    for t_start in (1599, 1999):
        alpha = 1 - math.sqrt(t_start / 2000)
        x = noise * math.sqrt(1 - alpha) + draft * math.sqrt(alpha)
        tr = ref_trace(md, x, 10, t_start=t_start)
        img = (tr[-1]["x0"] + 1) / 2
        fix[f"d2d{t_start}_img"] = img.numpy()
        record(f"fused.d2d.{t_start}.np32", maxabs(O.d2d(sd, sh, noise.numpy(), draft.numpy(), t_start), img))
    fix["d2d_noise_seed"] = np.array(23)
    np.savez_compressed(os.path.join(OUT, "fused.npz"), **fix)


def reference_evaluate():
    """The reference's `evaluate` (multi_gpu_trainer.py:32-45) executed from its own source text: the module itself
    cannot be imported here (it pulls in torchvision / tensorboard / a dataset path at import time)."""
    import torch.nn.functional as F
    with open(os.path.join(REF, "multi_gpu_trainer.py")) as f:
        src = f.read()
    a = src.index("def evaluate(")
    b = src.index("def main(", a)
    ns = {"torch": torch, "np": np, "F": F}
    exec(compile(src[a:b], "multi_gpu_trainer.py:evaluate", "exec"), ns)
    return ns["evaluate"]


def gen_evaluate(ViT, D2D, record):
    # ------------------------------------------------------------------ trainer-side evaluate() on the fused-path shape
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    sd = O.synth_state_dict(sh, seed=3, rich=True, gain=2.0)
    md = ref_model(D2D, sh, sd)                       # the trainer builds the ViT_draft2drawing class
    ev = reference_evaluate()
    batches = [tuple(torch.from_numpy(a) for a in b) for b in O.synth_eval_batches(sh)]
    mean = ev(md, batches, "cpu")
    per_batch = []
    import torch.nn.functional as F
    with torch.no_grad():
        for noisy, img, t in batches:
            per_batch.append(float(F.smooth_l1_loss(md(noisy, t), img)))
    assert abs(np.array(per_batch).mean() - mean) < 1e-12
    # the oracle's forward + a float64 smooth-L1 against the reference's numbers
    ol = []
    for noisy, img, t in O.synth_eval_batches(sh):
        d = np.abs(O.forward(sd, sh, noisy, t, np.float32).astype(np.float64) - img)
        ol.append(np.where(d < 1, 0.5 * d * d, d - 0.5).mean())
    record("evaluate.loss.np32", maxabs(np.array(ol), np.array(per_batch)))
    np.savez_compressed(os.path.join(OUT, "evaluate.npz"), mean=np.float64(mean), per_batch=np.array(per_batch, np.float64))


CASES = {"tiny": gen_tiny, "real": gen_real, "s3": gen_s3, "fused": gen_fused, "evaluate": gen_evaluate}


def main(argv=None):
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", choices=list(CASES), default=list(CASES),
                    help="regenerate only these fixture groups (PIN.json entries of the others are kept)")
    args = ap.parse_args(argv)
    torch.set_num_threads(8)
    ViT, D2D = import_reference()
    os.makedirs(OUT, exist_ok=True)
    pin_path = os.path.join(OUT, "PIN.json")
    pin = {"cases": {}}
    if os.path.exists(pin_path) and set(args.only) != set(CASES):
        with open(pin_path) as f:
            pin = json.load(f)
    pin["torch"], pin["numpy"] = torch.__version__, np.__version__

    def record(name, value):
        pin["cases"][name] = value
        print(f"{name}: {value}")

    for name in args.only:
        CASES[name](ViT, D2D, record)
    with open(pin_path, "w") as f:
        json.dump(pin, f, indent=1, sort_keys=True)
    worst = max(v for k, v in pin["cases"].items() if "clamp" not in k)
    print("worst oracle-vs-reference deviation:", worst)
    assert worst < 5e-4, worst


if __name__ == "__main__":
    main()
