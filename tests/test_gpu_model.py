"""GPU: the drop-in class (through the C ABI) against the CPU oracle and the fixtures recorded
from the unmodified reference.

Tolerances (north_star): teacher-forced per-step eps within 1e-2 max-abs with 16-bit operands; final
images within a PSNR bound.  The CUDA path feeds the tensor cores 16-bit operands (IEEE half by default,
bfloat16 on request) with fp32 accumulation, fp32 LayerNorm / softmax statistics and an fp32 residual
stream and image state.  The default format (fp16) is held to the 1e-2 bound; bfloat16 is an opt-in whose
measured miss of that bound (1.9e-2 at the lowest t) is asserted at 2.5e-2 and reported as a miss.
"""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

EPS_TOL = 1e-2          # north_star: per-step eps max-abs, teacher-forced; met with fp16 operands
X0_TOL = 1e-2
# bf16 operands cannot reach 1e-2 on eps at the lowest t (eps gain sqrt(a_t)/sqrt(1-a_t) = 3 at t=19):
# PyTorch's own bf16 autocast of the reference is 2.75e-2 off its fp32 self there.
# The bound asserted for bf16 is therefore "better than stock autocast", and is reported, not hidden.
EPS_TOL_BF16 = 2.5e-2
PSNR_MIN = 40.0         # final images in [0,1] vs the fp32 reference sampler, free-running


def psnr(a, b):
    mse = float(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2))
    return 99.0 if mse == 0 else 10 * math.log10(1.0 / mse)


def build(sh, sd, dev="cuda", fmt=None):
    from ddim_cold_b200 import DiffusionVisionTransformer
    m = DiffusionVisionTransformer(img_size=[sh.img_h, sh.img_w], patch_size=sh.patch, embed_dim=sh.embed,
                                   depth=sh.depth, num_heads=sh.heads, total_steps=sh.total_steps)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in sd.items()}, strict=True)
    if fmt is not None:
        m.operand_format = fmt
    assert m.operand_format in ("fp16", "bf16", "fp32")
    return m.eval().to(dev)


@pytest.fixture(scope="module")
def tiny(golden):
    from oracle import ddim_oracle as O
    sd = O.synth_state_dict(O.TINY, seed=11, rich=True, gain=4.0)
    return golden("tiny"), sd, build(O.TINY, sd)


def test_forward_tiny_vs_reference_fixture(tiny):
    g, sd, m = tiny
    y = m(torch.from_numpy(g["fwd_x"]).cuda(), torch.from_numpy(g["fwd_t"]).cuda()).cpu().numpy()
    assert np.abs(y - g["fwd_y"]).max() < 2e-2


def test_forward_rejects_bad_t(tiny):
    _, _, m = tiny
    with pytest.raises(IndexError):
        m(torch.zeros(1, 3, 16, 16).cuda(), torch.tensor([2000]).cuda())


@pytest.mark.parametrize("k", [100, 500, 2000])
def test_sampler_tiny_vs_reference_fixture(tiny, k, capsys):
    g, sd, m = tiny
    img = m.sampler("cuda", k, 3, noise=torch.from_numpy(g[f"samp{k}_xT"]))
    assert img.device.type == "cpu" and img.dtype == torch.float32 and tuple(img.shape) == (3, 3, 16, 16)
    assert float(img.min()) >= 0.0 and float(img.max()) <= 1.0
    assert psnr(img.numpy(), g[f"samp{k}_img"]) > PSNR_MIN
    assert "noise level" in capsys.readouterr().out


def test_sampler_seed_compat_with_reference(tiny):
    """Without noise= the sampler must consume the global CPU RNG like ViT.py:224."""
    g, sd, m = tiny
    torch.manual_seed(123)
    img = m.sampler("cuda", 100, 3)
    assert psnr(img.numpy(), g["samp100_img"]) > PSNR_MIN


def test_trace_teacher_forced_tiny(tiny):
    g, sd, m = tiny
    xs = torch.from_numpy(g["trace100_x_t"])          # [20, 3, 3, 16, 16]
    for j, t in enumerate(range(1999, 0, -100)):
        x0 = torch.clamp(m(xs[j].cuda(), torch.full((3,), t).cuda()), -1, 1).cpu().numpy()
        assert np.abs(x0 - g["trace100_x0"][j]).max() < X0_TOL, (j, t)


def test_bad_k_raises_before_gpu_work(tiny):
    _, _, m = tiny
    for k in (3, 7, 30, 4000):
        with pytest.raises(ValueError, match="math domain error"):
            m.sampler("cuda", k, 1)


def test_graph_equals_direct_launch(tiny):
    g, sd, m = tiny
    x = torch.from_numpy(g["samp100_xT"])
    a = m.denoise_trace("cuda", x, k=100, use_graph=True)
    b = m.denoise_trace("cuda", x, k=100, use_graph=False)
    for u, v in zip(a, b):
        assert torch.equal(u, v)


def test_deterministic_and_batch_invariant(tiny):
    """Same sample alone or inside a batch, run twice: bit-identical (sharding must not change
    any sample)."""
    g, sd, m = tiny
    x = torch.from_numpy(g["samp100_xT"])
    a = m.sampler("cuda", 100, 3, noise=x)
    b = m.sampler("cuda", 100, 3, noise=x)
    c = m.sampler("cuda", 100, 1, noise=x[1:2])
    assert torch.equal(a, b) and torch.equal(a[1:2], c)


def test_ragged_chunks_and_empty_batch(tiny):
    """Batches above `max_batch_per_call` run as chunks (here 3 + 3 + 1) overlapped on two streams: every sample must come
    out bit-identical to the single-call run, for the sampler and for the per-step sequence.  An empty batch is legal in
    the reference (ViT.py:224 draws a [0,3,H,W] tensor and every op accepts it) and returns an empty result."""
    g, sd, m = tiny
    x = torch.from_numpy(np.concatenate([g["samp100_xT"]] * 3)[:7].copy())
    whole = m.sampler("cuda", 100, 7, noise=x)
    seq_whole = m.diffusion_sequence("cuda", 500, 7, noise=x)
    from oracle import ddim_oracle as O
    m2 = build(O.TINY, sd)
    m2.max_batch_per_call = 3
    parts = m2.sampler("cuda", 100, 7, noise=x)
    seq_parts = m2.diffusion_sequence("cuda", 500, 7, noise=x)
    assert torch.equal(whole, parts)
    assert len(seq_whole) == len(seq_parts) and all(torch.equal(a, b) for a, b in zip(seq_whole, seq_parts))
    on_dev = m2.sampler("cuda", 100, 7, noise=x.cuda(), return_on_device=True)
    assert on_dev.is_cuda and torch.equal(on_dev.cpu(), whole)
    empty = m.sampler("cuda", 100, 0, noise=x[:0])
    assert tuple(empty.shape) == (0, 3) + tuple(x.shape[2:]) and empty.dtype == torch.float32
    seq = m.diffusion_sequence("cuda", 500, 0, noise=x[:0])
    assert len(seq) == len(seq_whole) and all(tuple(t.shape) == tuple(empty.shape) for t in seq)


def test_batch_invariant_on_the_fused_path():
    """The same check on a shape that takes the fused block tail, the tcgen05 GEMM pair kernels and the 65-token
    attention (embed 128, L = 65): one sample alone (a single, half-empty tile pair) == the same sample in a batch."""
    from oracle import ddim_oracle as O
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    m = build(sh, O.synth_state_dict(sh, seed=3, rich=True))
    x = torch.from_numpy(O.synth_noise((5, 3, 32, 32), seed=9))
    a = m.sampler("cuda", 200, 5, noise=x)
    c = m.sampler("cuda", 200, 1, noise=x[3:4])
    d = m.sampler("cuda", 200, 2, noise=x[3:5])
    assert torch.equal(a[3:4], c) and torch.equal(a[3:5], d)
    # 296 images = 74 four-image tiles on 74 CTA pairs: the fused qkv + attention kernel then takes whole tiles as work units,
    # where the small batches above cut every tile into two (per-launch cost model) -- a sample's bits must not depend on it
    big = torch.from_numpy(O.synth_noise((296, 3, 32, 32), seed=10))
    big[100:105] = x
    e = m.sampler("cuda", 200, 296, noise=big)
    assert torch.equal(e[100:105], a)


def test_sequence_cold_d2d_tiny(tiny):
    from oracle import ddim_oracle as O
    g, sd, m = tiny
    seq = m.diffusion_sequence("cuda", 400, 2, noise=torch.from_numpy(g["seq400_xT"]))
    assert len(seq) == 1 + 5 and np.array_equal(seq[0].numpy(), g["seq400_imgs"][0])
    assert min(psnr(s.numpy(), r) for s, r in zip(seq, g["seq400_imgs"])) > PSNR_MIN
    z = torch.from_numpy(g["cold_z"])
    assert psnr(m.cold_sampler("cuda", 10, 4, z=z).numpy(), g["cold_img"]) > PSNR_MIN
    cs = m.cold_diffusion_sequence("cuda", 4, z=z)
    assert len(cs) == 7 and min(psnr(s.numpy(), r) for s, r in zip(cs, g["cold_seq"])) > PSNR_MIN
    noise = torch.from_numpy(O.synth_noise((2, 3, 16, 16), seed=int(g["d2d_noise_seed"])))
    draft = torch.from_numpy(O.synth_draft(O.TINY))
    for t0 in (1599, 1999):
        img = m.draft2drawing("cuda", draft, t0, k=10, N=2, noise=noise)
        assert psnr(img.numpy(), g[f"d2d{t0}_img"]) > PSNR_MIN


@pytest.mark.parametrize("fmt", ["fp16", "bf16"])
@pytest.mark.parametrize("tag,seed", [("s1", 0), ("s2", 1)])
def test_real_shapes_teacher_forced_eps(golden, tag, seed, fmt):
    """S1 (yaml model) and S2 (OxfordFlower.pkl shape): per-step x0 and derived eps against the
    reference's own values at the recorded x_t, plus the free-running sampler image."""
    from oracle import ddim_oracle as O
    sh = {"s1": O.S1, "s2": O.S2}[tag]
    g = golden(tag)
    sd = O.synth_state_dict(sh, seed=seed, rich=True)
    m = build(sh, sd, fmt=fmt)
    worst_x0 = worst_eps = 0.0
    for j, t in enumerate(g["tap_t"]):
        x_t = g["tap_x_t"][j]
        x0 = torch.clamp(m(torch.from_numpy(x_t).cuda(), torch.tensor([int(t)]).cuda()), -1, 1).cpu().numpy()
        eps, _ = O.ddim_update(x_t, x0, int(t), 100)
        worst_x0 = max(worst_x0, float(np.abs(x0 - g["tap_x0"][j]).max()))
        worst_eps = max(worst_eps, float(np.abs(eps - g["tap_eps"][j]).max()))
    print(f"{tag} {fmt}: teacher-forced max|x0 err|={worst_x0:.3e} max|eps err|={worst_eps:.3e}")
    assert worst_x0 < X0_TOL and worst_eps < (EPS_TOL if fmt == "fp16" else EPS_TOL_BF16), (worst_x0, worst_eps)
    img = m.sampler("cuda", 100, 2, noise=torch.from_numpy(g["xT"])).numpy()
    p = psnr(img, g["samp100_img"])
    print(f"{tag} {fmt}: free-running sampler k=100 PSNR vs reference = {p:.1f} dB")
    assert p > PSNR_MIN


@pytest.mark.parametrize("fmt", ["fp16", "bf16"])
def test_s1_k20_eps_vs_oracle_all_steps(fmt):
    """Config #2 schedule (k=20, 100 steps) on S1, N=2: device-side eps taps of the free-running
    loop against the oracle evaluated at the device's own x_t (teacher-forced on the CUDA
    trajectory), every step."""
    from oracle import ddim_oracle as O
    from oracle import ddim_oracle_torch as OT
    sh = O.S1
    sd = O.synth_state_dict(sh, seed=0, rich=True)
    m = build(sh, sd, fmt=fmt)
    x_T = torch.from_numpy(O.synth_noise((2, 3, 64, 64), seed=123))
    x0_last, seq, eps = m.denoise_trace("cuda", x_T, k=20)
    # rebuild the device trajectory x_t from its own taps: x_{t-k} = a*x_t + b*x0
    sdt = OT.to_torch_sd(sd)
    x = x_T.clone()
    worst = 0.0
    for j, t in enumerate(range(1999, 0, -20)):
        x0_ref = torch.clamp(OT.forward(sdt, sh, x, torch.full((2,), t)), -1, 1)
        a_t = 1 - math.sqrt((t + 1) / 2000) + 1e-5
        eps_ref = (x - math.sqrt(a_t) * x0_ref) / math.sqrt(1 - a_t)
        worst = max(worst, float((eps[j] - eps_ref).abs().max()))
        a, b, _, _ = O.ddim_ab(t, 20)
        x = (a * x.double() + b * seq[j].double()).float()
    print(f"s1 k=20 {fmt}: worst teacher-forced eps error over 100 steps = {worst:.3e}")
    assert worst < (EPS_TOL if fmt == "fp16" else EPS_TOL_BF16)
    assert torch.equal(x0_last, seq[-1])


def test_default_operand_format_is_the_one_that_meets_the_bound():
    from ddim_cold_b200 import DiffusionVisionTransformer
    assert DiffusionVisionTransformer.operand_format == "fp16"


def test_fused_path_loops_vs_reference_fixture(golden):
    """diffusion_sequence / sampler / cold_sampler / cold_diffusion_sequence / draft->drawing on the shape that takes
    S1's kernels (fused block tail, CTA-pair GEMMs, 65-token attention: 32x32, patch 4, embed 128, 4 heads), against
    outputs recorded from the unmodified reference (ViT.py:220-256, ViT_draft2drawing.py:259-309,391-409)."""
    from oracle import ddim_oracle as O
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    g = golden("fused")
    m = build(sh, O.synth_state_dict(sh, seed=3, rich=True, gain=2.0))
    x_T = torch.from_numpy(g["seq400_xT"])
    seq = m.diffusion_sequence("cuda", 400, 3, noise=x_T)
    assert len(seq) == 1 + 5 and np.array_equal(seq[0].numpy(), g["seq400_imgs"][0])
    worst = max(float(np.abs(s.numpy() - r).max()) for s, r in zip(seq, g["seq400_imgs"]))
    print(f"fused-path diffusion_sequence: worst |image err| = {worst:.3e}")
    assert worst < X0_TOL / 2, worst                       # images are (x0 + 1) / 2
    assert psnr(m.sampler("cuda", 100, 3, noise=x_T).numpy(), g["samp100_img"]) > PSNR_MIN
    z = torch.from_numpy(g["cold_z"])
    cimg = m.cold_sampler("cuda", 10, 4, z=z).numpy()
    assert psnr(cimg, g["cold_img"]) > PSNR_MIN and np.abs(cimg - g["cold_img"]).max() < 2e-2
    cs = m.cold_diffusion_sequence("cuda", 4, z=z)
    assert len(cs) == 7 and min(psnr(s.numpy(), r) for s, r in zip(cs, g["cold_seq"])) > PSNR_MIN
    noise = torch.from_numpy(O.synth_noise((3, 3, 32, 32), seed=int(g["d2d_noise_seed"])))
    draft = torch.from_numpy(O.synth_draft(sh))
    for t0 in (1599, 1999):
        img = m.draft2drawing("cuda", draft, t0, k=10, N=3, noise=noise)
        p = psnr(img.numpy(), g[f"d2d{t0}_img"])
        print(f"fused-path d2d t_start={t0}: PSNR vs reference = {p:.1f} dB")
        assert p > PSNR_MIN


def test_s3_forward_and_sampler_vs_reference_fixture(golden):
    """S3 (declared 200x200 stand-in: 625 patches + cls = 626 tokens, head_dim 32): forward at three timesteps,
    teacher-forced x0 / eps on the recorded k = 500 trajectory, and the free-running sampler image."""
    from oracle import ddim_oracle as O
    g = golden("s3")
    sd = O.synth_state_dict(O.S3, seed=2, rich=True)
    m = build(O.S3, sd)
    y = m(torch.from_numpy(g["fwd_x"]).cuda(), torch.from_numpy(g["fwd_t"]).cuda()).cpu().numpy()
    err = float(np.abs(y - g["fwd_y"]).max())
    print(f"s3 forward: max|y err| = {err:.3e}")
    assert err < X0_TOL
    worst_x0 = worst_eps = 0.0
    for j, t in enumerate(g["tap_t"]):
        x_t = g["tap_x_t"][j]
        x0 = torch.clamp(m(torch.from_numpy(x_t).cuda(), torch.tensor([int(t)]).cuda()), -1, 1).cpu().numpy()
        eps, _ = O.ddim_update(x_t, x0, int(t), 500)
        worst_x0 = max(worst_x0, float(np.abs(x0 - g["tap_x0"][j]).max()))
        worst_eps = max(worst_eps, float(np.abs(eps - g["tap_eps"][j]).max()))
    print(f"s3: teacher-forced max|x0 err|={worst_x0:.3e} max|eps err|={worst_eps:.3e}")
    assert worst_x0 < X0_TOL and worst_eps < EPS_TOL
    img = m.sampler("cuda", 500, 1, noise=torch.from_numpy(g["xT"])).numpy()
    p = psnr(img, g["samp500_img"])
    print(f"s3: free-running sampler k=500 PSNR vs reference = {p:.1f} dB")
    assert p > PSNR_MIN


EPS_TOL_FP32 = 1e-4     # north_star: per-step eps max-abs in fp32 mode


@pytest.mark.parametrize("tag,seed", [("s1", 0), ("s2", 1)])
def test_fp32_mode_teacher_forced_eps(golden, tag, seed):
    """operand_format = "fp32" (every contraction in fp32 on the CUDA cores): per-step x0 and derived eps against the
    reference's own values at the recorded x_t within 1e-4, and the free-running sampler image."""
    from oracle import ddim_oracle as O
    sh = {"s1": O.S1, "s2": O.S2}[tag]
    g = golden(tag)
    sd = O.synth_state_dict(sh, seed=seed, rich=True)
    m = build(sh, sd, fmt="fp32")
    worst_x0 = worst_eps = 0.0
    for j, t in enumerate(g["tap_t"]):
        x_t = g["tap_x_t"][j]
        x0 = torch.clamp(m(torch.from_numpy(x_t).cuda(), torch.tensor([int(t)]).cuda()), -1, 1).cpu().numpy()
        eps, _ = O.ddim_update(x_t, x0, int(t), 100)
        worst_x0 = max(worst_x0, float(np.abs(x0 - g["tap_x0"][j]).max()))
        worst_eps = max(worst_eps, float(np.abs(eps - g["tap_eps"][j]).max()))
    print(f"{tag} fp32 mode: teacher-forced max|x0 err|={worst_x0:.3e} max|eps err|={worst_eps:.3e}")
    assert worst_x0 < EPS_TOL_FP32 and worst_eps < EPS_TOL_FP32, (worst_x0, worst_eps)
    img = m.sampler("cuda", 100, 2, noise=torch.from_numpy(g["xT"])).numpy()
    err = float(np.abs(img - g["samp100_img"]).max())
    print(f"{tag} fp32 mode: free-running sampler k=100 max|image err| = {err:.3e}, PSNR {psnr(img, g['samp100_img']):.1f} dB")
    assert err < 1e-3 and psnr(img, g["samp100_img"]) > 80.0


def test_fp32_mode_tiny_api_and_hot_weights(tiny):
    """fp32 mode on the fixture-sized model whose weights are scaled x4 (53 % of the pixels hit the clamp, softmax and GELU
    see large arguments): forward with per-image t, sampler, cold sampler, and switching the same model between formats."""
    g, sd, m16 = tiny
    from oracle import ddim_oracle as O
    m = build(O.TINY, sd, fmt="fp32")
    y = m(torch.from_numpy(g["fwd_x"]).cuda(), torch.from_numpy(g["fwd_t"]).cuda()).cpu().numpy()
    assert np.abs(y - g["fwd_y"]).max() < 2e-4 * max(1.0, np.abs(g["fwd_y"]).max())
    for k in (100, 2000):
        img = m.sampler("cuda", k, 3, noise=torch.from_numpy(g[f"samp{k}_xT"])).numpy()
        assert np.abs(img - g[f"samp{k}_img"]).max() < 2e-3, k          # free-running over up to 20 steps, hot weights
    z = torch.from_numpy(g["cold_z"])
    assert np.abs(m.cold_sampler("cuda", 10, 4, z=z).numpy() - g["cold_img"]).max() < 2e-3
    a = m.sampler("cuda", 500, 3, noise=torch.from_numpy(g["samp500_xT"]))
    m.operand_format = "fp16"
    b = m.sampler("cuda", 500, 3, noise=torch.from_numpy(g["samp500_xT"]))
    m.operand_format = "fp32"
    c = m.sampler("cuda", 500, 3, noise=torch.from_numpy(g["samp500_xT"]))
    assert torch.equal(a, c) and psnr(a.numpy(), b.numpy()) > PSNR_MIN and not torch.equal(a, b)


def test_fp32_mode_s3_forward(golden):
    from oracle import ddim_oracle as O
    g = golden("s3")
    m = build(O.S3, O.synth_state_dict(O.S3, seed=2, rich=True), fmt="fp32")
    y = m(torch.from_numpy(g["fwd_x"]).cuda(), torch.from_numpy(g["fwd_t"]).cuda()).cpu().numpy()
    err = float(np.abs(y - g["fwd_y"]).max())
    print(f"s3 fp32 mode forward: max|y err| = {err:.3e}")
    assert err < EPS_TOL_FP32


def test_s1_full_batch_properties_of_the_benchmarked_configuration():
    """BASELINE config #2 at full size (S1, 1024 images): properties that need no oracle.  (1) Every image's prediction
    equals the prediction of the same image in a batch of four (the fused qkv + attention kernel cuts the work
    differently for the two batch sizes); (2) two runs are bit-identical; (3) the plan reports the fused path for this
    shape in a 16-bit format and not in fp32 mode."""
    from ddim_cold_b200 import _lib
    from oracle import ddim_oracle as O
    sh = O.S1
    m = build(sh, O.synth_state_dict(sh, seed=0, rich=True))
    x = torch.from_numpy(O.synth_noise((1024, 3, sh.img_h, sh.img_w), seed=77)).cuda()
    t = (torch.arange(1024, device="cuda") * 37 + 11) % sh.total_steps
    y = m(x, t)
    assert torch.isfinite(y).all()
    assert torch.equal(y, m(x, t))
    for lo in (0, 508, 1020):
        assert torch.equal(y[lo:lo + 4], m(x[lo:lo + 4], t[lo:lo + 4])), lo
    lib = _lib.load()
    plan = m._plan(torch.device("cuda", torch.cuda.current_device()), 4)
    assert lib.ddc_plan_fused_qkv_attention(plan.handle) == 1
    m.operand_format = "fp32"
    plan = m._plan(torch.device("cuda", torch.cuda.current_device()), 4)
    assert lib.ddc_plan_fused_qkv_attention(plan.handle) == 0


def test_separate_qkv_attention_setting_agrees_with_the_fused_kernel():
    """`fused_qkv_attention = False` (the low-latency choice for small batches) runs the qkv GEMM and the attention kernel
    separately: same prediction within the 16-bit tolerance, and the step graph is rebuilt when the setting flips."""
    from oracle import ddim_oracle as O
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    sd = O.synth_state_dict(sh, seed=3, rich=True)
    m = build(sh, sd)
    x = torch.from_numpy(O.synth_noise((6, 3, 32, 32), seed=21))
    a = m.sampler("cuda", 200, 6, noise=x)
    m.fused_qkv_attention = False
    b = m.sampler("cuda", 200, 6, noise=x)
    m.fused_qkv_attention = True
    c = m.sampler("cuda", 200, 6, noise=x)
    assert torch.equal(a, c)
    assert not torch.equal(a, b) and (a - b).abs().max().item() < 5e-3
    ref = O.sampler(sd, sh, x.numpy(), 200, np.float32)
    assert np.abs(b.numpy() - ref).max() < 2e-2
