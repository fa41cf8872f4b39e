"""Benchmark of the DDIM sampling hot path (BASELINE.json metric: DDIM images/sec (N, k)).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--shape s1|s2|s3]

One "step" of this benchmark is one complete pass of the hot path over one batch: the whole
k-strided DDIM loop (k=20 -> 100 predictor evaluations + updates) for `--batch` images per GPU
of the workload BASELINE.json quotes its metric on (configs[1]: the yaml vit_tiny model "S1",
N=1024, k=20, one B200).  With N GPUs every rank runs its own 1024 images (weak scaling, no
collective on the step path; the finished images are gathered once with NCCL in the e2e leg).

Keys (one JSON line, rank 0): value = images/s with x_T already resident in HBM (CUDA events
around ddc_sample, max over ranks); e2e = the same through the public API
`DiffusionVisionTransformer.sampler(device, k, N, noise=<host tensor>)` with the host->device
copy of x_T and the device->host read of the result inside the timed region; roofline = the
dominant kernel class (LN+qkv GEMM) measured with CUDA events between launches by
ddc_profile_steps; cpu_baseline = the oracle's torch-CPU port of the reference sampler timed on
this box's host cores on config #1 (N=16, k=20).

--impl reference times that CPU port alone (the reference is pure PyTorch and is not present on
the GPU box; the port calls the same ATen CPU kernels in the same order).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

SHAPES = {
    "s1": dict(img=64, patch=8, embed=384, depth=7, heads=12, desc="S1 yaml vit_tiny_diffusion 64x64 p8 C384 d7 h12 (L=65, head_dim 32)"),
    "s2": dict(img=64, patch=4, embed=256, depth=6, heads=4, desc="S2 OxfordFlower.pkl shape 64x64 p4 C256 d6 h4 (L=257, head_dim 64)"),
    "s3": dict(img=200, patch=8, embed=384, depth=7, heads=12, desc="S3 declared 200x200 stand-in p8 C384 d7 h12 (L=626, head_dim 32)"),
}
CLASS_NAMES = ["patchify", "embed_gemm", "ln_qkv_gemm", "attention", "proj_gemm", "ln_fc1_gemm", "fc2_gemm",
               "ln_head_gemm", "head_ddim"]


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons in the background during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        sm, reasons, mx = [], set(), None
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s in sm if mx and s > 0.3 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_model(shape_key: str, dev):
    from ddim_cold_b200 import DiffusionVisionTransformer
    from oracle import ddim_oracle as O     # only for the deterministic synthetic weights
    s = SHAPES[shape_key]
    sh = O.Shape(s["img"], s["img"], s["patch"], s["embed"], s["depth"], s["heads"])
    sd = O.synth_state_dict(sh, seed=0, rich=True)
    m = DiffusionVisionTransformer(img_size=[s["img"], s["img"]], patch_size=s["patch"], embed_dim=s["embed"],
                                   depth=s["depth"], num_heads=s["heads"])
    m.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()}, strict=True)
    return m.eval().to(dev), sh, sd


def cpu_port_run(shape_key: str, n_img: int, k: int, threads: int, reps: int = 1, warm: bool = True):
    """Times the torch-CPU port of the reference sampler (oracle/ddim_oracle_torch.py)."""
    from oracle import ddim_oracle as O
    from oracle import ddim_oracle_torch as OT
    s = SHAPES[shape_key]
    sh = O.Shape(s["img"], s["img"], s["patch"], s["embed"], s["depth"], s["heads"])
    sd = OT.to_torch_sd(O.synth_state_dict(sh, seed=0, rich=True))
    torch.set_num_threads(threads)
    x = torch.from_numpy(O.synth_noise((n_img, 3, s["img"], s["img"]), seed=123))
    if warm:
        OT.sampler(sd, sh, x, 500)
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        OT.sampler(sd, sh, x, k)
        times.append(time.perf_counter() - t0)
    return times


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (port), rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_img, k = 16, args.k
    s = SHAPES[args.shape]
    times = cpu_port_run(args.shape, n_img, k, cores, reps=args.warmup + args.steps, warm=False)[args.warmup:]
    dt = sum(times) / len(times)
    val = n_img / dt
    unit = "images/s"
    line = {
        "impl": "reference", "metric": f"DDIM images/sec (N={args.batch}/GPU, k={k})", "value": val, "unit": unit,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{s['desc']}, N={args.batch}/GPU, k={k}; CPU sample: N={n_img}, k={k} per step"},
        "cpu_baseline": {"value": val, "unit": unit, "cores": cores, "kind": "port",
                         "sample": f"torch-CPU port of ViT.py:220-237 (same ATen ops), fp32, N={n_img}, k={k} "
                                   f"({len(range(1999, 0, -k))} denoise steps), {cores} threads"},
        "e2e": {"value": val, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--shape", default="s1", choices=list(SHAPES))
    ap.add_argument("--batch", type=int, default=None, help="images per GPU (default 1024; 256 total for s3)")
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.k is None:
        args.k = 10 if args.shape == "s3" else 20
    if args.batch is None:
        args.batch = 256 if args.shape == "s3" else 1024
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)
    if args.impl == "reference":
        return run_reference(args)

    import torch.distributed as dist
    from ddim_cold_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py (ours) needs a B200; there is no CPU path"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    model, sh, _ = make_model(args.shape, dev)
    B, k = args.batch, args.k
    H = SHAPES[args.shape]["img"]
    table, n_steps = _lib.ddim_table(k, 2000)
    from oracle import ddim_oracle as O     # synthetic inputs + FLOP formula only
    x_host = torch.from_numpy(O.synth_noise((B, 3, H, H), seed=123 + rank)).pin_memory()
    plan = model._plan(dev, B)
    stream = plan.stream
    launches_per_pass = lib.ddc_launches_per_step(plan.handle) * n_steps

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # ---------------- device-resident leg: x_T in HBM, CUDA events around the loop
    x_dev = x_host.to(dev)
    x_work = torch.empty_like(x_dev)
    x0 = torch.empty_like(x_dev)

    def one_pass():
        x_work.copy_(x_dev, non_blocking=True)           # restore x_T (device-to-device, ~0.05 ms)
        _lib.check(lib.ddc_sample(plan.handle, x_work.data_ptr(), B, table, n_steps, x0.data_ptr(), None, None, 1,
                                  stream.cuda_stream))

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            one_pass()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            one_pass()
        e1.record(stream)
    barrier()
    dev_ms = e0.elapsed_time(e1)
    clk = clocks.stop() if rank == 0 else None
    assert torch.isfinite(x0).all()

    # ---------------- end-to-end leg: public API, host buffers in, host tensor out
    import contextlib
    with contextlib.redirect_stdout(sys.stderr):       # the sampler prints a progress line like the reference
        for _ in range(2):
            model.sampler(dev, k, B, noise=x_host)
    e2e_out = torch.empty((B * world, 3, H, H), dtype=torch.float32).pin_memory() if (world > 1 and rank == 0) else None
    bufs = [torch.empty((B, 3, H, H), dtype=torch.float32, device=dev) for _ in range(world)] if (world > 1 and rank == 0) else None

    def e2e_pass():
        with contextlib.redirect_stdout(sys.stderr):
            img = model.sampler(dev, k, B, noise=x_host, return_on_device=world > 1)
        if world > 1:                                  # the one NCCL call: final gather of the images, then one D2H
            dist.gather(img, bufs, dst=0)
            if rank == 0:
                for r_, b_ in enumerate(bufs):
                    e2e_out[r_ * B:(r_ + 1) * B].copy_(b_, non_blocking=True)
            torch.cuda.synchronize(dev)

    if world > 1:
        e2e_pass()                                     # untimed: NCCL sets up its gather connections on first use
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_pass()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    print("", file=sys.stderr)

    # ---------------- per-class kernel timing (CUDA events between launches, direct launches)
    cls_ms = (C.c_float * 9)()
    cls_n = (C.c_int32 * 9)()
    prof_steps = min(n_steps, 10)
    x_work.copy_(x_dev)
    torch.cuda.synchronize(dev)
    _lib.check(lib.ddc_profile_steps(plan.handle, x_work.data_ptr(), B, table, prof_steps, cls_ms, cls_n,
                                     stream.cuda_stream))

    if world > 1:
        t = torch.tensor([dev_ms, e2e_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms = t.tolist()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    total_images = B * world
    value = total_images * args.steps / (dev_ms / 1e3)
    e2e_val = total_images * args.steps / (e2e_ms / 1e3)
    pk, pk_src = peaks()
    flops_img_step = O.flops_per_image_step(sh)
    Cdim, L = sh.embed, sh.tokens
    M = B * L
    # algorithmic FLOPs per launch of each tensor-core class (2*M*N*K; attention 4*h*L^2*d per image)
    heads = SHAPES[args.shape]["heads"]
    fused_tail = cls_n[5] == 0 and cls_n[4] > 0          # proj + fc1 + fc2 run as one block-tail kernel
    P = sh.patch_dim
    cls_flops = {1: 2.0 * M * Cdim * P, 2: 2.0 * M * 3 * Cdim * Cdim, 3: 4.0 * B * L * L * Cdim,
                 4: (6.0 if fused_tail else 2.0) * M * Cdim * Cdim, 5: 2.0 * M * Cdim * Cdim, 6: 2.0 * M * Cdim * Cdim,
                 7: 2.0 * M * Cdim * P}
    names = list(CLASS_NAMES)
    if fused_tail:
        names[4] = "block_tail(proj+ln+fc1+gelu+fc2+ln)"
    peak_tf = pk.get("bf16_tflops_sustained", pk["bf16_tflops"])
    kernels = {}
    step_ms_sum = sum(cls_ms[i] for i in range(9)) / prof_steps
    for i, name in enumerate(names):
        if cls_n[i]:
            ms = cls_ms[i] / cls_n[i]
            kernels[name] = {"launches_per_step": cls_n[i] // prof_steps, "ms_per_launch": ms,
                             "share": (cls_ms[i] / prof_steps) / step_ms_sum}
            if i in cls_flops:
                kernels[name]["tflops"] = cls_flops[i] / (ms * 1e-3) / 1e12
    # roofline object: the kernel class with the largest share of the step (tensor-bound: dense contractions)
    dom = max((i for i in (2, 4) if cls_n[i]), key=lambda i: cls_ms[i])
    dom_ms = cls_ms[dom] / cls_n[dom]
    dom_flops = cls_flops[dom]
    achieved = dom_flops / (dom_ms * 1e-3) / 1e12
    dom_name = {2: "gemm2_kernel<S_BF16> (qkv GEMM, tcgen05 cta_group::2)",
                4: "block_tail_kernel (proj + LN2 + fc1 + GELU + fc2 + LN, tcgen05 cta_group::2)" if fused_tail
                   else "gemm2_kernel<R_RESID> (proj GEMM + residual + LN)"}[dom]
    whole_tf = flops_img_step * B * n_steps * args.steps * world / (dev_ms / 1e3) / 1e12 / world
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this workload
    # (S1, 1024 images per GPU); other workloads were not captured -> null
    ncu_traffic = {4: (250.5e6, "profiles/r01d_ncu_full.txt"), 2: (163.7e6, "profiles/r01d_ncu_full.txt")}
    traffic, traffic_src = (ncu_traffic[dom] if (args.shape == "s1" and B == 1024 and (dom != 4 or fused_tail))
                            else (None, None))
    line = {
        "metric": f"DDIM images/sec (N={B}/GPU, k={k})", "value": value, "unit": "images/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"{SHAPES[args.shape]['desc']}, N={B}/GPU, k={k} ({n_steps} denoise steps per pass), "
                               "random-init weights (no checkpoint ships)",
                   "l2": "per-pass working set (fp32 residual + bf16 qkv/attn/mlp activations, "
                         f"{plan.work.numel() / 1e6:.0f} MB) exceeds the 126 MB L2; no explicit flush",
                   "parallelism": f"batch-sharded x{world}, no step-path collective"},
        "ms_per_denoise_step": dev_ms / args.steps / n_steps,
        "tensor_tflops_whole_loop_per_gpu": whole_tf,
        "tensor_frac_whole_loop": whole_tf / peak_tf,
        "e2e": {"value": e2e_val, "unit": "images/s", "h2d_bytes_per_step": int(x_host.numel() * 4) * world,
                "d2h_bytes_per_step": int(x_host.numel() * 4) * world, "ms_per_step": e2e_ms / args.steps,
                "note": "N>1: each rank uploads its own x_T; one NCCL gather to rank 0, which reads all images back"},
        "gpu_launches": launches_per_pass * args.steps,
        "clocks": clk,
        "roofline": {"bound": "tensor", "kernel": dom_name,
                     "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": f"{pk_src} bf16 sustained (MEASURED_PEAKS.json)",
                     "flops_per_launch": dom_flops, "ms_per_launch": dom_ms,
                     "share_of_step": (cls_ms[dom] / prof_steps) / step_ms_sum},
        "kernels": kernels,
    }
    if not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        tms = cpu_port_run("s1" if args.shape == "s1" else args.shape, 16, 20, cores, reps=1)
        line["cpu_baseline"] = {"value": 16 / tms[0], "unit": "images/s", "cores": cores, "kind": "port",
                                "sample": f"config #1: torch-CPU port of the reference sampler (ViT.py:220-237), fp32, "
                                          f"N=16, k=20 (100 denoise steps), {cores} threads, {tms[0]:.2f} s"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
