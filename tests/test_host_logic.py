"""CPU: host-side logic of the drop-in -- C-ABI symbols, schedule tables, state_dict layout,
checkpoint / yaml ingestion, sharding -- no GPU compute."""
import ctypes as C
import math
import os
import re

import numpy as np
import pytest
import torch

from oracle import ddim_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from ddim_cold_b200 import _lib
    lib = _lib.load()
    with open(os.path.join(ROOT, "include", "ddimcold.h")) as f:
        hdr = f.read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ddc_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in ddimcold.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert C.sizeof(_lib.Step) == 32 and C.sizeof(_lib.Shape) == 32


@pytest.mark.parametrize("k", [1, 10, 20, 50, 100, 1999, 2000])
def test_ddim_table_matches_oracle(k):
    from ddim_cold_b200 import _lib
    tab, n = _lib.ddim_table(k, 2000)
    ts = O.steps_of(k)
    assert n == len(ts)
    for j in (0, n // 3, n - 1):
        a, b, sa, inv = O.ddim_ab(ts[j], k)
        e = tab[j]
        assert e.t == ts[j]
        assert (e.a, e.b, e.sqrt_at, e.inv_sqrt_1mat) == tuple(np.float32(v) for v in (a, b, sa, inv))
        assert e.flags == (1 if j == n - 1 else 0)


def test_ddim_table_errors_match_reference():
    from ddim_cold_b200 import _lib
    for k in (3, 7, 30, 4000):
        with pytest.raises(ValueError, match="math domain error"):
            _lib.ddim_table(k, 2000)
    tab, n = _lib.ddim_table(10, 2000, 1599)          # draft->drawing schedule
    assert n == 160 and tab[0].t == 1599 and tab[n - 1].t == 9
    tab, n = _lib.cold_table()
    assert [tab[i].t for i in range(n)] == [6, 5, 4, 3, 2, 1] and tab[5].flags == 1


def test_arena_bytes_and_shape_validation():
    from ddim_cold_b200 import _lib
    lib = _lib.load()
    wb, kb = C.c_size_t(), C.c_size_t()
    s1 = _lib.Shape(64, 64, 8, 384, 7, 12, 3, 2000)
    assert lib.ddc_arena_bytes(C.byref(s1), 1024, C.byref(wb), C.byref(kb)) == 0
    assert 15e6 < wb.value < 25e6          # 12.8 MB 16-bit weights + 6.2 MB head-pair copy of the qkv weights + 3 MB fp32 time table + biases
    assert kb.value < 600e6
    bad = _lib.Shape(64, 64, 8, 192, 7, 3, 3, 2000)     # DINO vit_tiny: 192 is not a multiple of 64 ... head_dim 64
    assert lib.ddc_arena_bytes(C.byref(bad), 8, C.byref(wb), C.byref(kb)) == 0
    bad = _lib.Shape(64, 64, 8, 384, 7, 5, 3, 2000)
    assert lib.ddc_arena_bytes(C.byref(bad), 8, C.byref(wb), C.byref(kb)) == _lib.DDC_E_INVALID
    assert b"head_dim" in lib.ddc_last_error()


def test_state_dict_layout_and_strict_load():
    from ddim_cold_b200 import DiffusionVisionTransformer
    for sh in (O.S1, O.S2, O.TINY):
        m = DiffusionVisionTransformer(img_size=[sh.img_h, sh.img_w], patch_size=sh.patch, embed_dim=sh.embed,
                                       depth=sh.depth, num_heads=sh.heads)
        spec = O.state_dict_spec(sh)
        assert [(k, tuple(v.shape)) for k, v in m.state_dict().items()] == spec
        assert all(v.dtype == torch.float32 for v in m.state_dict().values())
        sd = O.synth_state_dict(sh, seed=1)
        m.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()}, strict=True)
        bad = dict(sd)
        bad.pop("head.bias")
        with pytest.raises(RuntimeError):
            m.load_state_dict({k: torch.from_numpy(v) for k, v in bad.items()}, strict=True)
    assert (m.embed_dim, m.num_features, m.patch_size, m.in_chans, m.img_size, m.total_steps) == \
        (64, 64, 4, 3, [16, 16], 2000)


def test_unsupported_ctor_args_fail_loudly():
    from ddim_cold_b200 import DiffusionVisionTransformer
    with pytest.raises(NotImplementedError):
        DiffusionVisionTransformer(mlp_ratio=4.0)
    with pytest.raises(NotImplementedError):
        DiffusionVisionTransformer(qkv_bias=False)


def test_no_cpu_fallback():
    from ddim_cold_b200 import DiffusionVisionTransformer
    m = DiffusionVisionTransformer(img_size=[16, 16], patch_size=4, embed_dim=64, depth=2, num_heads=2).eval()
    with pytest.raises(RuntimeError, match="no CPU path"):
        m.sampler("cpu", 100, 1)
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.zeros(1, 3, 16, 16), torch.zeros(1, dtype=torch.long))


@pytest.mark.skipif(not os.path.exists("/root/reference/ViT.py"), reason="reference not mounted")
def test_random_init_is_seed_identical_to_reference():
    import sys
    sys.dont_write_bytecode = True
    sys.path.insert(0, "/root/reference")
    import ViT
    from ddim_cold_b200 import DiffusionVisionTransformer
    kw = dict(img_size=[16, 16], patch_size=4, embed_dim=64, depth=2, num_heads=2, total_steps=50)
    torch.manual_seed(7)
    ref = ViT.DiffusionVisionTransformer(**kw).state_dict()
    torch.manual_seed(7)
    mine = DiffusionVisionTransformer(**kw).state_dict()
    assert list(ref.keys()) == list(mine.keys())
    for k in ref:
        assert torch.equal(ref[k], mine[k]), k


def test_checkpoint_ingestion(tmp_path):
    from ddim_cold_b200 import infer_shape, load_checkpoint, model_from_yaml
    sd = {k: torch.from_numpy(v) for k, v in O.synth_state_dict(O.TINY, seed=2).items()}
    kw = infer_shape(sd, num_heads=2)
    assert kw == dict(img_size=[16, 16], patch_size=4, in_chans=3, embed_dim=64, depth=2, num_heads=2,
                      total_steps=2000)
    p = tmp_path / "bestloss.pkl"
    torch.save(sd, p)
    m = load_checkpoint(str(p), num_heads=2)
    assert torch.equal(m.state_dict()["head.weight"], sd["head.weight"])
    # lastepoch.pkl: nested + DDP prefix (multi_gpu_trainer.py:155-163)
    p2 = tmp_path / "lastepoch.pkl"
    torch.save({"epoch": 3, "state_dict": {"module." + k: v for k, v in sd.items()}}, p2)
    m2 = load_checkpoint(str(p2), num_heads=2)
    assert torch.equal(m2.state_dict()["cls_token"], sd["cls_token"])
    with pytest.raises(ValueError):
        infer_shape(sd, num_heads=5)
    y = model_from_yaml(dict(image_size=[64, 64], patch_size=8, embed_dim=384, depth=7, head=12))
    assert len(y.state_dict()) == 93 and y.total_steps == 2000


def test_shard_bounds_cover_everything():
    from ddim_cold_b200 import shard_bounds
    for n in (1, 5, 8, 1024, 1000, 3):
        for w in (1, 2, 4, 8):
            spans = [shard_bounds(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
