"""ctypes binding of the C ABI declared in include/ddimcold.h.

There is no fallback: if ``libddimcold.so`` is missing or a call fails, an exception is
raised.  The library is built in-tree by ``__graft_entry__.build()`` (or ``make -C
ddim_cold_b200/csrc``) for sm_100a only.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libddimcold.so")

DDC_OK, DDC_E_INVALID, DDC_E_CUDA, DDC_E_DOMAIN, DDC_E_STATE = 0, -1, -2, -3, -4


class Shape(C.Structure):
    """ddc_shape"""
    _fields_ = [(n, C.c_int32) for n in
                ("img_h", "img_w", "patch", "embed", "depth", "heads", "in_chans", "total_steps")]


class Step(C.Structure):
    """ddc_step"""
    _fields_ = [("t", C.c_int32), ("flags", C.c_int32), ("a", C.c_float), ("b", C.c_float),
                ("sqrt_at", C.c_float), ("inv_sqrt_1mat", C.c_float), ("pad0", C.c_float), ("pad1", C.c_float)]


class DdcError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"ddimcold error {code}: {msg}")
        self.code = code
        self.msg = msg


# name -> (restype, argtypes); every entry must be declared in include/ddimcold.h
_vp, _i32, _sz = C.c_void_p, C.c_int32, C.c_size_t
SIGNATURES = {
    "ddc_last_error": (C.c_char_p, []),
    "ddc_version": (C.c_int, []),
    "ddc_ddim_table": (C.c_int, [_i32, _i32, _i32, C.POINTER(Step), _i32, C.POINTER(_i32)]),
    "ddc_cold_table": (C.c_int, [C.POINTER(Step), _i32, C.POINTER(_i32)]),
    "ddc_arena_bytes": (C.c_int, [C.POINTER(Shape), _i32, C.POINTER(_sz), C.POINTER(_sz)]),
    "ddc_plan_create": (C.c_int, [C.POINTER(Shape), _i32, _i32, _vp, _sz, _vp, _sz, C.POINTER(_vp)]),
    "ddc_plan_destroy": (C.c_int, [_vp]),
    "ddc_pack_weights": (C.c_int, [_vp, C.POINTER(_vp), _i32, _vp]),
    "ddc_forward": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp]),
    "ddc_forward_fixed_scratch": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp]),
    "ddc_sample": (C.c_int, [_vp, _vp, _i32, C.POINTER(Step), _i32, _vp, _vp, _vp, _i32, _vp]),
    "ddc_launches_per_step": (C.c_int, [_vp]),
    "ddc_plan_max_batch": (C.c_int, [_vp]),
    "ddc_plan_fused_qkv_attention": (C.c_int, [_vp]),
    "ddc_plan_set_fused_qkv_attention": (C.c_int, [_vp, _i32]),
    "ddc_profile_steps": (C.c_int, [_vp, _vp, _i32, C.POINTER(Step), _i32, C.POINTER(C.c_float), C.POINTER(_i32), _vp]),
    "ddc_plan_set_operand_format": (C.c_int, [_vp, _i32]),
    "ddc_k_gemm": (C.c_int, [_i32, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _i32, _vp]),
    "ddc_k_gemm_resid_ln": (C.c_int, [_i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ddc_k_embed": (C.c_int, [_i32] * 6 + [_vp] * 10),
    "ddc_k_block_tail": (C.c_int, [_i32, _i32, _i32] + [_vp] * 10),
    "ddc_k_attention": (C.c_int, [_i32, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "ddc_k_qkv_attn": (C.c_int, [_i32, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "ddc_k_patchify": (C.c_int, [_i32, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "ddc_k_affine": (C.c_int, [_vp, _vp, _sz, C.c_float, C.c_float, _vp]),
    "ddc_k_smooth_l1": (C.c_int, [_vp, _vp, _sz, _vp, _vp, _vp]),
    "ddc_k_mix": (C.c_int, [_vp, _vp, _vp, _sz, _sz, C.c_float, C.c_float, _vp]),
    "ddc_k_head_ddim": (C.c_int, [_vp, _i32, _vp, _i32, _i32, _i32, _i32, C.POINTER(Step), _vp, _vp, _vp]),
}

_lib = None


def load() -> C.CDLL:
    """Load the CUDA library; raises if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  ddim_cold_b200 has no CPU or PyTorch fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)      # AttributeError if the .so is stale
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(code: int) -> None:
    if code == DDC_OK:
        return
    msg = load().ddc_last_error().decode("utf-8", "replace")
    if code == DDC_E_DOMAIN:
        # the reference fails mid-loop with these exact Python exceptions (ViT.py:231,234)
        if "division by zero" in msg:
            raise ZeroDivisionError(msg)
        raise ValueError(msg)
    raise DdcError(code, msg)


def ddim_table(k: int, total_steps: int, t_start: int | None = None):
    """Host schedule for `for t in range(t_start, 0, -k)` -> (ctypes array, n)."""
    lib = load()
    t0 = total_steps - 1 if t_start is None else int(t_start)
    cap = max(1, total_steps)
    tab = (Step * cap)()
    n = _i32(0)
    check(lib.ddc_ddim_table(int(k), int(total_steps), t0, tab, cap, C.byref(n)))
    return tab, n.value


def cold_table():
    lib = load()
    tab = (Step * 6)()
    n = _i32(0)
    check(lib.ddc_cold_table(tab, 6, C.byref(n)))
    return tab, n.value
