"""Checkpoint / config ingestion for the drop-in class (host logic, no GPU needed).

Formats follow the reference: ``bestloss.pkl`` / ``OxfordFlower*.pkl`` are bare pickled
state dicts (multi_gpu_trainer.py:152-154, ViT.py:278-280); ``lastepoch.pkl`` nests the DDP
state dict under ``['state_dict']`` with a ``module.`` prefix (multi_gpu_trainer.py:155-163);
the yaml carries ``image_size, patch_size, embed_dim, depth, head`` (20220822.yaml:10-15,
multi_gpu_trainer.py:205-210) and never ``total_steps`` (always 2000, ViT.py:162).
"""
from __future__ import annotations

import math

import torch

from .vit import DiffusionVisionTransformer


def normalize_state_dict(obj):
    """Accept a bare state dict, a `lastepoch.pkl` dict, or DDP-prefixed keys."""
    if isinstance(obj, dict) and "state_dict" in obj and "cls_token" not in obj:
        obj = obj["state_dict"]
    return {(k[len("module."):] if k.startswith("module.") else k): v for k, v in obj.items()}


def infer_shape(state_dict, num_heads: int, img_size=None) -> dict:
    """Constructor kwargs from tensor shapes.  ``num_heads`` cannot be inferred (qkv is one
    fused [3C, C] matrix -- SURVEY.md F9) and must be given; a non-square image needs
    ``img_size`` because only the patch count is recorded."""
    sd = normalize_state_dict(state_dict)
    C, in_chans, p, p2 = sd["patch_embed.proj.weight"].shape
    if p != p2:
        raise ValueError("non-square patches are not part of the reference")
    L = sd["pos_embed"].shape[1]
    depth = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("blocks."))
    if sd["blocks.0.mlp.fc1.weight"].shape[0] != C:
        raise NotImplementedError("mlp_ratio != 1.0")
    if C % num_heads != 0:
        raise ValueError(f"embed_dim {C} not divisible by num_heads {num_heads}")
    if img_size is None:
        side = math.isqrt(L - 1)
        if side * side != L - 1:
            raise ValueError(f"{L - 1} patches is not a square grid: pass img_size")
        img_size = [side * p, side * p]
    if (img_size[0] // p) * (img_size[1] // p) != L - 1:
        raise ValueError(f"img_size {img_size} / patch {p} does not give {L - 1} patches")
    return dict(img_size=list(img_size), patch_size=int(p), in_chans=int(in_chans), embed_dim=int(C), depth=depth,
                num_heads=int(num_heads), total_steps=int(sd["time_embed.weight"].shape[0]))


def load_checkpoint(path_or_obj, num_heads: int, img_size=None, map_location="cpu") -> DiffusionVisionTransformer:
    """Build the model a checkpoint was saved from and load it ``strict=True``."""
    obj = torch.load(path_or_obj, map_location=map_location) if isinstance(path_or_obj, (str, bytes)) else path_or_obj
    sd = normalize_state_dict(obj)
    model = DiffusionVisionTransformer(**infer_shape(sd, num_heads, img_size))
    model.load_state_dict(sd, strict=True)
    return model.eval()


def model_from_yaml(path_or_dict, total_steps: int = 2000) -> DiffusionVisionTransformer:
    """Model from the trainer's yaml shape keys (multi_gpu_trainer.py:69,205-210)."""
    cfg = path_or_dict
    if not isinstance(cfg, dict):
        import yaml
        with open(path_or_dict) as f:
            cfg = yaml.safe_load(f)
    return DiffusionVisionTransformer(img_size=list(cfg["image_size"]), patch_size=cfg["patch_size"],
                                      embed_dim=cfg["embed_dim"], depth=cfg["depth"], num_heads=cfg["head"],
                                      total_steps=total_steps).eval()
