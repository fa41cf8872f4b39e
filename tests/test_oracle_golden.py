"""CPU: the oracle (numpy fp32/fp64 and the torch port) against the fixtures recorded from
the unmodified reference by oracle/make_golden.py.  Tolerances are the fp32 re-association
noise observed at generation time (tests/golden/PIN.json), with head-room."""
import json
import math
import os

import numpy as np
import pytest
import torch

from oracle import ddim_oracle as O
from oracle import ddim_oracle_torch as OT

TOL = 2e-5


def maxabs(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))))


@pytest.fixture(scope="module")
def tiny(golden):
    return golden("tiny"), O.synth_state_dict(O.TINY, seed=11, rich=True, gain=4.0)


def test_pin_file_records_small_deviation():
    with open(os.path.join(os.path.dirname(__file__), "golden", "PIN.json")) as f:
        pin = json.load(f)
    worst = max(v for k, v in pin["cases"].items() if "clamp" not in k)
    assert worst < 2e-5
    assert all(v == 0.0 for k, v in pin["cases"].items() if k.endswith("torchport"))


def test_state_dict_spec_counts():
    assert len(O.state_dict_spec(O.S1)) == 93 and len(O.state_dict_spec(O.S2)) == 81
    n1 = sum(int(np.prod(s)) for _, s in O.state_dict_spec(O.S1))
    n2 = sum(int(np.prod(s)) for _, s in O.state_dict_spec(O.S2))
    assert (n1, n2) == (7162176, 2978096)


def test_flops_formula():
    assert abs(O.flops_per_image_step(O.S1) / 1e6 - 869.56) < 0.01
    assert abs(O.flops_per_image_step(O.S2) / 1e6 - 1631.09) < 0.01
    assert abs(O.flops_per_image_step(O.S3) / 1e6 - 12151.74) < 0.01


def test_forward_tiny(tiny):
    g, sd = tiny
    for dt in (np.float32, np.float64):
        assert maxabs(O.forward(sd, O.TINY, g["fwd_x"], g["fwd_t"], dt), g["fwd_y"]) < TOL
    y = OT.forward(OT.to_torch_sd(sd), O.TINY, torch.from_numpy(g["fwd_x"]), torch.from_numpy(g["fwd_t"]))
    assert maxabs(y.numpy(), g["fwd_y"]) < TOL


@pytest.mark.parametrize("k", [100, 500, 2000])
def test_sampler_tiny(tiny, k):
    g, sd = tiny
    assert maxabs(O.sampler(sd, O.TINY, g[f"samp{k}_xT"], k), g[f"samp{k}_img"]) < TOL
    img = OT.sampler(OT.to_torch_sd(sd), O.TINY, torch.from_numpy(g[f"samp{k}_xT"]), k)
    assert maxabs(img.numpy(), g[f"samp{k}_img"]) < TOL


def test_trace_teacher_forced_tiny(tiny):
    g, sd = tiny
    xs = list(g["trace100_x_t"])
    _, tr = O.denoise_loop(sd, O.TINY, xs[0], 100, taps=True, teacher=xs)
    assert len(tr) == 20
    for j, s in enumerate(tr):
        assert maxabs(s["x0"], g["trace100_x0"][j]) < TOL
        assert maxabs(s["eps"], g["trace100_eps"][j]) < TOL
        assert maxabs(s["x_next"], g["trace100_x_next"][j]) < TOL
    # the last step of a k | T run lands on alpha_tk = 1  ->  x_next == x0 (A = 0, B = 1)
    assert maxabs(tr[-1]["x_next"], tr[-1]["x0"]) < 1e-5


def test_bad_k_raises_like_reference(tiny):
    g, sd = tiny
    for k in g["bad_k"]:
        with pytest.raises(ValueError, match="math domain error"):
            O.sampler(sd, O.TINY, np.zeros((1, 3, 16, 16), np.float32), int(k))
    for k, n in ((1, 1999), (10, 200), (20, 100), (50, 40)):
        assert len(O.steps_of(k)) == n


def test_sequence_cold_d2d_tiny(tiny):
    g, sd = tiny
    assert maxabs(np.stack(O.diffusion_sequence(sd, O.TINY, g["seq400_xT"], 400)), g["seq400_imgs"]) < TOL
    x_init = np.broadcast_to(g["cold_z"][:, :, None, None], (4, 3, 16, 16))
    seq = O.cold_sequence(sd, O.TINY, x_init)
    assert maxabs(np.stack(seq), g["cold_seq"]) < TOL and maxabs(seq[-1], g["cold_img"]) < TOL
    noise = O.synth_noise((2, 3, 16, 16), seed=int(g["d2d_noise_seed"]))
    draft = O.synth_draft(O.TINY)
    for t0 in (1599, 1999):
        assert maxabs(O.d2d(sd, O.TINY, noise, draft, t0), g[f"d2d{t0}_img"]) < TOL


def test_collapsed_coefficients_match_literal_update():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 3, 8, 8))
    x0 = np.clip(rng.standard_normal((2, 3, 8, 8)), -1, 1)
    for k in (1, 10, 20, 50):
        for t in O.steps_of(k)[:: max(1, len(O.steps_of(k)) // 7)] + [O.steps_of(k)[-1]]:
            eps, xn = O.ddim_update(x, x0, t, k)
            a, b, sa, inv = O.ddim_ab(t, k)
            assert maxabs(a * x + b * x0, xn) < 1e-9
            assert maxabs((x - sa * x0) * inv, eps) < 1e-9
    a, b, _, _ = O.ddim_ab(O.steps_of(20)[-1], 20)
    assert a == 0.0 and b == 1.0


@pytest.mark.parametrize("tag,sh,seed", [("s1", O.S1, 0), ("s2", O.S2, 1)])
def test_real_shapes_taps(golden, tag, sh, seed):
    g = golden(tag)
    sd = O.synth_state_dict(sh, seed=seed, rich=True)
    sdt = OT.to_torch_sd(sd)
    for j, t in enumerate(g["tap_t"]):
        x_t = g["tap_x_t"][j]
        x0 = np.clip(O.forward(sd, sh, x_t, np.array([t])), -1, 1)
        eps, xn = O.ddim_update(x_t, x0, int(t), 100)
        assert maxabs(x0, g["tap_x0"][j]) < TOL
        assert maxabs(eps, g["tap_eps"][j]) < TOL
        assert maxabs(xn, g["tap_x_next"][j]) < TOL
        y = torch.clamp(OT.forward(sdt, sh, torch.from_numpy(x_t), torch.tensor([int(t)])), -1, 1)
        assert maxabs(y.numpy(), g["tap_x0"][j]) < TOL


FUSED = O.Shape(32, 32, 4, 128, 2, 4)      # the shape whose GPU run takes the fused block tail + 65-token attention


def test_fused_shape_loops(golden):
    """Every loop of the class API on the fixture-sized twin of S1's code path (tests/golden/fused.npz)."""
    g = golden("fused")
    sd = O.synth_state_dict(FUSED, seed=3, rich=True, gain=2.0)
    assert maxabs(np.stack(O.diffusion_sequence(sd, FUSED, g["seq400_xT"], 400)), g["seq400_imgs"]) < TOL
    assert maxabs(O.sampler(sd, FUSED, g["seq400_xT"], 100), g["samp100_img"]) < TOL
    x_init = np.broadcast_to(g["cold_z"][:, :, None, None], (4, 3, 32, 32))
    seq = O.cold_sequence(sd, FUSED, x_init)
    assert maxabs(np.stack(seq), g["cold_seq"]) < TOL and maxabs(seq[-1], g["cold_img"]) < TOL
    noise = O.synth_noise((3, 3, 32, 32), seed=int(g["d2d_noise_seed"]))
    draft = O.synth_draft(FUSED)
    for t0 in (1599, 1999):
        assert maxabs(O.d2d(sd, FUSED, noise, draft, t0), g[f"d2d{t0}_img"]) < TOL


def test_s3_forward_and_taps(golden):
    """S3 (declared 200x200 stand-in, L = 626): the torch port against the reference's forward at three timesteps
    and against the first step of the recorded k = 500 trajectory (the numpy oracle on one image to keep it quick)."""
    g = golden("s3")
    sd = O.synth_state_dict(O.S3, seed=2, rich=True)
    sdt = OT.to_torch_sd(sd)
    y = OT.forward(sdt, O.S3, torch.from_numpy(g["fwd_x"]), torch.from_numpy(g["fwd_t"]))
    assert maxabs(y.numpy(), g["fwd_y"]) < TOL
    assert maxabs(O.forward(sd, O.S3, g["fwd_x"][:1], g["fwd_t"][:1]), g["fwd_y"][:1]) < TOL
    x0 = torch.clamp(OT.forward(sdt, O.S3, torch.from_numpy(g["tap_x_t"][0]), torch.tensor([int(g["tap_t"][0])])), -1, 1)
    assert maxabs(x0.numpy(), g["tap_x0"][0]) < TOL
    eps, _ = O.ddim_update(g["tap_x_t"][0], g["tap_x0"][0], int(g["tap_t"][0]), 500)
    assert maxabs(eps, g["tap_eps"][0]) < TOL


def test_evaluate_fixture_vs_oracle(golden):
    """Trainer-side evaluate() (multi_gpu_trainer.py:32-45): the per-batch smooth-L1 losses the reference's own function
    returned for the synthetic batches, against the oracle's forward + a float64 smooth-L1."""
    g = golden("evaluate")
    sh = O.Shape(32, 32, 4, 128, 2, 4)
    sd = O.synth_state_dict(sh, seed=3, rich=True, gain=2.0)
    losses = [O.smooth_l1_mean(O.forward(sd, sh, noisy, t, np.float32), img) for noisy, img, t in O.synth_eval_batches(sh)]
    assert np.abs(np.array(losses) - g["per_batch"]).max() < 1e-6
    assert abs(np.mean(losses) - float(g["mean"])) < 1e-6
