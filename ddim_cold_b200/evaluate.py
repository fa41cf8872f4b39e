"""Validation pass of the trainer on the B200 forward (reference multi_gpu_trainer.py:32-45).

The reference's `evaluate(model, dataloader, device)` runs `model(noisy_img, t)` under `no_grad` for every batch of the
test loader, takes `F.smooth_l1_loss(denoised_img, img)` and returns the mean of the per-batch losses.  Training itself
(backward, optimiser, DDP) stays stock PyTorch; this module only lets the trainer's evaluation reuse the inference path:
the prediction comes from `DiffusionVisionTransformer.forward` (the CUDA chain behind `ddc_forward`) and the loss is
reduced on the device by `ddc_k_smooth_l1`, one float per batch coming back to the host as in the reference
(`loss.item()`).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

_scratch: dict = {}


def smooth_l1_loss(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """F.smooth_l1_loss(pred, target) (beta = 1, mean reduction) for fp32 CUDA tensors: a 0-dim tensor on the device."""
    if pred.shape != target.shape:
        raise RuntimeError(f"The size of tensor a {tuple(pred.shape)} must match the size of tensor b {tuple(target.shape)}")
    if pred.device.type != "cuda" or target.device != pred.device:
        raise RuntimeError("ddim_cold_b200 computes on CUDA (sm_100a) only; there is no CPU path")
    a = pred.detach().to(torch.float32).contiguous()
    b = target.detach().to(torch.float32).contiguous()
    dev = a.device
    scr = _scratch.get(dev)
    if scr is None:
        scr = _scratch[dev] = torch.zeros(1024, dtype=torch.float32, device=dev)
    out = torch.empty((), dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(_lib.load().ddc_k_smooth_l1(a.data_ptr(), b.data_ptr(), a.numel(), scr.data_ptr(), out.data_ptr(), st))
    return out


def evaluate(model, dataloader, device):
    """Mean smooth-L1 loss of the x0 prediction over `dataloader`, whose batches are (noisy_img, img, t)
    (reference multi_gpu_trainer.py:32-45: same iteration, same per-batch `.item()`, same numpy mean)."""
    model.eval()
    datagenerator = iter(dataloader)
    loss_arr = []
    for _ in range(len(dataloader)):
        noisy_img, img, t = next(datagenerator)
        noisy_img = noisy_img.to(device)
        img = img.to(device)
        t = t.to(device)
        with torch.no_grad():
            denoised_img = model(noisy_img, t)
            loss = smooth_l1_loss(denoised_img, img)
        loss_arr.append(loss.item())
    return np.array(loss_arr).mean()
