"""Headless command line of the sampling path: the reference's ``main()`` (ViT.py:258-305) without a display.

Same options (``--sample_n``, ``--acc_k``, same defaults), same two products -- a denoising-sequence sheet
(``model.diffusion_sequence(device, 100, N=6)``, one row per sample, one column per step, ViT.py:283-293) and a sample
sheet (``model.sampler(device, acc_k, sample_n)``, 16 x 16 grid, ViT.py:295-305) -- written as PNG files by PIL instead
of matplotlib figures (matplotlib is presentation only and not installed here).

Deliberate differences:
- ``get_next_path`` (ViT.py:307-313) never increments ``i`` in the reference and spins forever once the file exists;
  here it counts up.
- the checkpoint path, model shape and output directory are options instead of literals; without ``--ckpt`` the weights
  are the constructor's seeded random init (there is no checkpoint in the reference tree), which the program says.
- tiles are laid out with a fixed pixel padding instead of matplotlib's inch-based ``ImageGrid``.

Run:  python -m ddim_cold_b200.cli --sample_n 256 --acc_k 20 --out_dir out/
"""
from __future__ import annotations

import os

import click
import numpy as np
import torch

from .checkpoint import load_checkpoint, model_from_yaml
from .vit import DiffusionVisionTransformer

# the two literal model shapes of the reference's main() (ViT.py:274 commented-out miniImageNet line, :277 OxfordFlower)
_SHAPES = {
    "oxfordflower": dict(img_size=[64, 64], total_steps=2000, patch_size=4, embed_dim=256, depth=6, num_heads=4),
    "miniimagenet": dict(img_size=[64, 64], total_steps=2000, patch_size=8, embed_dim=384, depth=7, num_heads=12),
}


def get_next_path(pth: str) -> str:
    """First of pth, prefix_1.ext, prefix_2.ext, ... that does not exist yet (ViT.py:307-313, with the loop fixed)."""
    prefix, ext = os.path.splitext(pth)
    i, candidate = 1, pth
    while os.path.isfile(candidate):
        candidate = f"{prefix}_{i}{ext}"
        i += 1
    return candidate


def image_grid(imgs: torch.Tensor, ncols: int, pad: int = 2, fill: float = 1.0) -> np.ndarray:
    """[N,3,H,W] floats in [0,1] -> one uint8 [rows*(H+pad)+pad, ncols*(W+pad)+pad, 3] sheet, row-major like
    ``zip(grid, img)`` over an ImageGrid (ViT.py:290-292).  Conversion is ``to_pil_image``'s: x*255, truncated."""
    if imgs.dim() != 4 or imgs.shape[1] != 3:
        raise ValueError(f"expected [N,3,H,W], got {tuple(imgs.shape)}")
    n, _, h, w = imgs.shape
    if n == 0 or ncols <= 0:
        raise ValueError("empty grid")
    nrows = (n + ncols - 1) // ncols
    sheet = np.full((nrows * (h + pad) + pad, ncols * (w + pad) + pad, 3), int(fill * 255), dtype=np.uint8)
    u8 = imgs.detach().to("cpu", torch.float32).clamp(0, 1).mul(255).to(torch.uint8).permute(0, 2, 3, 1).numpy()
    for i in range(n):
        r, c = divmod(i, ncols)
        y0, x0 = pad + r * (h + pad), pad + c * (w + pad)
        sheet[y0:y0 + h, x0:x0 + w] = u8[i]
    return sheet


def save_png(path: str, sheet: np.ndarray) -> str:
    from PIL import Image

    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    path = get_next_path(path)
    Image.fromarray(sheet, mode="RGB").save(path)
    return path


def sequence_sheet(seq) -> np.ndarray:
    """List of per-step [N,3,H,W] tensors -> N rows x len(seq) columns (ViT.py:289: stack, transpose(0,1), flatten)."""
    imgs = torch.stack(list(seq), dim=0).transpose(0, 1).flatten(0, 1)
    return image_grid(imgs, ncols=len(seq))


def build_model(ckpt, config, shape, heads, seed) -> DiffusionVisionTransformer:
    if ckpt:
        if heads is None:
            heads = _SHAPES[shape]["num_heads"]
        return load_checkpoint(ckpt, num_heads=heads)
    if seed is not None:
        torch.manual_seed(seed)
    if config:
        return model_from_yaml(config)
    return DiffusionVisionTransformer(**_SHAPES[shape])


@click.command()
@click.option("--sample_n", default=256, help="Number of samples you'll get.")
@click.option("--acc_k", default=1, help="Number of step jumped during sampling.")
@click.option("--ckpt", default=None, type=click.Path(exists=True, dir_okay=False),
              help="state_dict pickle (plain, or nested under 'state_dict' with a 'module.' prefix).")
@click.option("--config", default=None, type=click.Path(exists=True, dir_okay=False),
              help="yaml with image_size / patch_size / embed_dim / depth / head (used when no --ckpt).")
@click.option("--shape", default="oxfordflower", type=click.Choice(sorted(_SHAPES)), show_default=True,
              help="Literal model shape of the reference's main() (used when neither --ckpt nor --config).")
@click.option("--heads", default=None, type=int, help="num_heads for --ckpt (not recoverable from tensor shapes).")
@click.option("--out_dir", default="Saved_Models", show_default=True, help="Where the PNG sheets go.")
@click.option("--seq_n", default=6, show_default=True, help="Rows of the denoising-sequence sheet (0 = skip it).")
@click.option("--seq_k", default=100, show_default=True, help="Step jump of the denoising-sequence sheet.")
@click.option("--seed", default=None, type=int, help="torch.manual_seed before random init and before sampling.")
@click.option("--operand_format", default="fp16", type=click.Choice(["fp16", "bf16"]), show_default=True)
def main(sample_n, acc_k, ckpt, config, shape, heads, out_dir, seq_n, seq_k, seed, operand_format):
    """Sample with the B200 path and write denoise_sequence.png and samples.png (headless ViT.py main())."""
    if not torch.cuda.is_available():
        raise click.ClickException("no CUDA device: ddim_cold_b200 has no CPU path")
    device = torch.device("cuda")
    model = build_model(ckpt, config, shape, heads, seed)
    if not ckpt:
        click.echo("no --ckpt given: weights are the constructor's random init" +
                   (f" (seed {seed})" if seed is not None else ""))
    model.operand_format = operand_format
    model.eval()
    model.to(device)
    if seed is not None:
        torch.manual_seed(seed)
    if seq_n > 0:
        seq = model.diffusion_sequence(device, seq_k, N=seq_n)
        click.echo("\nwrote " + save_png(os.path.join(out_dir, "denoise_sequence.png"), sequence_sheet(seq)))
    img = model.sampler(device, acc_k, sample_n)
    click.echo("\nwrote " + save_png(os.path.join(out_dir, "samples.png"), image_grid(img[:256], ncols=16)))


if __name__ == "__main__":
    main()
