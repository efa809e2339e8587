"""Operator-level access to the CUDA kernels (hfg_op_tapped_conv) for parity tests.

torch is used only to own the device buffers."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from . import _lib

_ACT = {"fp32": 0, "fp16": 1, "bf16": 2}


def tapped_conv(x, w: np.ndarray, bias: np.ndarray, tap_off: Sequence[int], *, engine: str = "tc",
                out_stride: int = 1, out_off: int = 0, in_slope: float = 1.0, res=None, res_inv: float = 1.0,
                res2=None, out_scale: float = 1.0, out_slope: float = 1.0, acc=None,
                mrf_mode: int = 0, mrf_div: float = 3.0, out=None):
    """x: torch CUDA tensor [B, rows, c_in] (float32 / float16 / bfloat16); w: fp32 [ntaps, c_in, c_out].
    engine "ffma" = direct kernel, "tc" = persistent tcgen05 kernel (see include/hifigan_b200.h)."""
    import torch
    lib = _lib.load()
    act = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}[x.dtype]
    B, rows, c_in = (int(v) for v in x.shape)
    ntaps, c_in_w, c_out = (int(v) for v in w.shape)
    if c_in_w != c_in or len(tap_off) != ntaps:
        raise ValueError("weight / tap shapes do not match the input")
    x = x.contiguous()
    w = np.ascontiguousarray(w, dtype=np.float32)
    bias = np.ascontiguousarray(bias, dtype=np.float32)
    if out is None:
        out = torch.zeros((B, rows * out_stride, c_out), dtype=x.dtype, device=x.device)
    taps = (C.c_int * ntaps)(*[int(t) for t in tap_off])
    ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
    rc = lib.hfg_op_tapped_conv(act, {"ffma": 0, "tc": 1, "tf32": 2}[engine], ptr(x), B, rows, c_in,
                                w.ctypes.data_as(C.c_void_p), bias.ctypes.data_as(C.c_void_p), c_out, ntaps, taps,
                                out_stride, out_off, float(in_slope), ptr(res), float(res_inv), ptr(res2),
                                float(out_scale), float(out_slope), ptr(acc), int(mrf_mode), float(mrf_div), ptr(out),
                                C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream))
    _lib.check(rc, None, "hfg_op_tapped_conv")
    return out


def resblock_fused(x, w1: np.ndarray, b1: np.ndarray, off1: Sequence[int], w2: np.ndarray, b2: np.ndarray, off2: Sequence[int], *,
                   mode: int = 1, mid_slope: float = 0.1, res_inv: float = 10.0, res2=None, out_scale: float = 1.0,
                   out_slope: float = 0.1, out=None):
    """One fused ResBlock1 iteration (mode 1) or whole ResBlock2 (mode 2) (hfg_op_resblock_fused): x torch CUDA
    [B, rows, 64] float16 / bfloat16 stored activated; w1 / w2 fp32 [ntaps, 64, 64]."""
    import torch
    lib = _lib.load()
    act = {torch.float16: 1, torch.bfloat16: 2}[x.dtype]
    B, rows, c = (int(v) for v in x.shape)
    if c != 64 or tuple(w1.shape[1:]) != (64, 64) or tuple(w2.shape[1:]) != (64, 64):
        raise ValueError("the fused kernel works on 64 kernel-view channels")
    x = x.contiguous()
    w1 = np.ascontiguousarray(w1, dtype=np.float32); w2 = np.ascontiguousarray(w2, dtype=np.float32)
    b1 = np.ascontiguousarray(b1, dtype=np.float32); b2 = np.ascontiguousarray(b2, dtype=np.float32)
    if out is None:
        out = torch.zeros_like(x)
    o1 = (C.c_int * len(off1))(*[int(t) for t in off1]); o2 = (C.c_int * len(off2))(*[int(t) for t in off2])
    ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
    rc = lib.hfg_op_resblock_fused(act, int(mode), ptr(x), B, rows, w1.ctypes.data_as(C.c_void_p), b1.ctypes.data_as(C.c_void_p), len(off1), o1,
                                   w2.ctypes.data_as(C.c_void_p), b2.ctypes.data_as(C.c_void_p), len(off2), o2,
                                   float(mid_slope), float(res_inv), ptr(res2), float(out_scale), float(out_slope), ptr(out),
                                   C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream))
    _lib.check(rc, None, "hfg_op_resblock_fused")
    return out
