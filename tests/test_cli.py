"""Headless writer + CLI: the reference's main() (ViT.py:258-313) without matplotlib."""
import os

import numpy as np
import pytest
import torch
from click.testing import CliRunner

from ddim_cold_b200 import cli


def test_get_next_path_counts_up(tmp_path):
    # the reference's loop never increments i and would spin forever on the second call (ViT.py:307-313)
    p = str(tmp_path / "samples.png")
    assert cli.get_next_path(p) == p
    open(p, "w").close()
    p1 = cli.get_next_path(p)
    assert p1 == str(tmp_path / "samples_1.png")
    open(p1, "w").close()
    assert cli.get_next_path(p) == str(tmp_path / "samples_2.png")


def test_image_grid_layout_and_conversion():
    imgs = torch.zeros(5, 3, 4, 6)
    for i in range(5):
        imgs[i, 0] = (i + 1) / 10.0           # red channel encodes the tile index
    imgs[4, 1] = 0.999                         # 0.999*255 = 254.7 -> 254 (to_pil_image truncates)
    sheet = cli.image_grid(imgs, ncols=3, pad=1)
    assert sheet.shape == (2 * 5 + 1, 3 * 7 + 1, 3) and sheet.dtype == np.uint8
    for i in range(5):
        r, c = divmod(i, 3)
        tile = sheet[1 + r * 5:1 + r * 5 + 4, 1 + c * 7:1 + c * 7 + 6]
        assert (tile[..., 0] == int((i + 1) / 10.0 * 255)).all()
    assert (sheet[6:10, 8:14, 1] == 254).all()
    assert (sheet[0] == 255).all() and (sheet[6:10, 15:21] == 255).all()    # padding and the empty sixth cell
    with pytest.raises(ValueError):
        cli.image_grid(torch.zeros(2, 1, 4, 4), ncols=2)


def test_sequence_sheet_is_one_row_per_sample():
    steps, n = 4, 3
    seq = [torch.full((n, 3, 2, 2), 0.0) for _ in range(steps)]
    for s in range(steps):
        for i in range(n):
            seq[s][i, 0] = (s * n + i + 1) / 20.0
    sheet = cli.sequence_sheet(seq)
    assert sheet.shape == (n * 4 + 2, steps * 4 + 2, 3)
    for i in range(n):
        for s in range(steps):       # row = sample, column = step (ViT.py:289)
            assert sheet[2 + i * 4, 2 + s * 4, 0] == int((s * n + i + 1) / 20.0 * 255)


def test_cli_options_match_the_reference():
    names = {p.name: p for p in cli.main.params}
    assert names["sample_n"].default == 256 and names["acc_k"].default == 1     # ViT.py:261-262
    assert "--sample_n" in names["sample_n"].opts and "--acc_k" in names["acc_k"].opts
    res = CliRunner().invoke(cli.main, ["--help"])
    assert res.exit_code == 0 and "--sample_n" in res.output and "--acc_k" in res.output


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_cli_refuses_to_run_without_cuda(tmp_path):
    res = CliRunner().invoke(cli.main, ["--sample_n", "2", "--out_dir", str(tmp_path)])
    assert res.exit_code != 0 and "no CPU path" in res.output
    assert not os.listdir(tmp_path)


@pytest.mark.gpu
def test_cli_writes_both_sheets(tmp_path):
    from PIL import Image
    import yaml

    cfg = tmp_path / "tiny.yaml"
    cfg.write_text(yaml.safe_dump({"image_size": [16, 16], "patch_size": 4, "embed_dim": 64, "depth": 2, "head": 2}))
    args = ["--sample_n", "20", "--acc_k", "500", "--config", str(cfg), "--out_dir", str(tmp_path / "out"),
            "--seq_n", "3", "--seq_k", "500", "--seed", "7"]
    res = CliRunner().invoke(cli.main, args)
    assert res.exit_code == 0, res.output
    seq = np.asarray(Image.open(tmp_path / "out" / "denoise_sequence.png"))
    smp = np.asarray(Image.open(tmp_path / "out" / "samples.png"))
    n_steps = len(range(1999, 0, -500))
    assert seq.shape == (3 * 18 + 2, (1 + n_steps) * 18 + 2, 3)            # 3 rows, 1 + steps columns
    assert smp.shape == (2 * 18 + 2, 16 * 18 + 2, 3)                       # 20 samples in rows of 16
    assert smp[2:18, 2:18].std() > 0                                       # an image, not a blank tile
    # a second run must not overwrite (and must terminate, unlike the reference's get_next_path)
    res = CliRunner().invoke(cli.main, args)
    assert res.exit_code == 0, res.output
    assert (tmp_path / "out" / "samples_1.png").exists() and (tmp_path / "out" / "denoise_sequence_1.png").exists()
    # same seed, same weights, same noise -> same sheet
    assert (np.asarray(Image.open(tmp_path / "out" / "samples_1.png")) == smp).all()
