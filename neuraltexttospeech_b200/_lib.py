"""ctypes binding of libhifigan_b200.so (include/hifigan_b200.h).

There is deliberately no fallback: if the shared library is missing or fails to load the
import error is raised to the caller -- nothing in this package computes on the CPU.
"""
from __future__ import annotations

import ctypes as C
import os

from .build import LIB_PATH

HFG_MAX_UPSAMPLES = 8
HFG_MAX_KERNELS = 8
HFG_MAX_DILATIONS = 4

HFG_ERR_INVALID, HFG_ERR_STATE, HFG_ERR_CUDA, HFG_ERR_NOMEM, HFG_ERR_UNSUPPORTED = -1, -2, -3, -4, -5
PRECISIONS = {"fp32": 0, "fp16": 1, "bf16": 2, "tf32": 3}


class HfgConfig(C.Structure):
    _fields_ = [
        ("num_mels", C.c_int32),
        ("upsample_initial_channel", C.c_int32),
        ("resblock", C.c_int32),
        ("num_upsamples", C.c_int32),
        ("upsample_rates", C.c_int32 * HFG_MAX_UPSAMPLES),
        ("upsample_kernel_sizes", C.c_int32 * HFG_MAX_UPSAMPLES),
        ("num_kernels", C.c_int32),
        ("resblock_kernel_sizes", C.c_int32 * HFG_MAX_KERNELS),
        ("num_dilations", C.c_int32 * HFG_MAX_KERNELS),
        ("resblock_dilation_sizes", (C.c_int32 * HFG_MAX_DILATIONS) * HFG_MAX_KERNELS),
        ("compat", C.c_int32),
    ]


class HfgMelConfig(C.Structure):
    _fields_ = [
        ("fft_length", C.c_int32),
        ("frame_length", C.c_int32),
        ("frame_step", C.c_int32),
        ("n_mel_channels", C.c_int32),
        ("sampling_rate", C.c_int32),
        ("mel_fmin", C.c_float),
        ("mel_fmax", C.c_float),
    ]


# every symbol include/hifigan_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "hfg_version": (C.c_int, []),
    "hfg_create": (C.c_int, [C.POINTER(HfgConfig), C.c_int, C.POINTER(C.c_void_p)]),
    "hfg_set_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int]),
    "hfg_finalize_weights": (C.c_int, [C.c_void_p, C.c_int]),
    "hfg_total_upsample": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "hfg_workspace_bytes": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "hfg_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "hfg_forward_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "hfg_forward_pcm16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "hfg_forward_host_pcm16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "hfg_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_longlong]),
    "hfg_layer_engines": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "hfg_auto_micro_batch": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "hfg_layer_folds": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "hfg_last_launch_count": (C.c_int, [C.c_void_p, C.POINTER(C.c_longlong)]),
    "hfg_stage_times_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "hfg_stage_launches": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "hfg_op_tapped_conv": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                     C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_float,
                                     C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_int, C.c_float, C.c_void_p, C.c_void_p]),
    "hfg_op_tapped_conv_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                             C.POINTER(C.c_int), C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_float, C.c_void_p,
                                             C.c_void_p, C.c_char_p, C.c_size_t]),
    "hfg_op_resblock_fused": (C.c_int, [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int),
                                        C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.c_float, C.c_float, C.c_void_p,
                                        C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "hfg_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "hfg_host_free": (C.c_int, [C.c_void_p]),
    "hfg_last_error": (C.c_char_p, [C.c_void_p]),
    "hfg_destroy": (None, [C.c_void_p]),
    "hfg_mel_create": (C.c_int, [C.POINTER(HfgMelConfig), C.c_int, C.POINTER(C.c_void_p)]),
    "hfg_mel_num_frames": (C.c_int, [C.c_void_p, C.c_longlong, C.POINTER(C.c_int)]),
    "hfg_mel_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p]),
    "hfg_mel_forward_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_void_p]),
    "hfg_mel_last_error": (C.c_char_p, [C.c_void_p]),
    "hfg_mel_destroy": (None, [C.c_void_p]),
    "hfg_disc_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    "hfg_disc_set_weight": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.POINTER(C.c_int64), C.c_int]),
    "hfg_disc_finalize_weights": (C.c_int, [C.c_void_p]),
    "hfg_disc_num_maps": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "hfg_disc_layout": (C.c_int, [C.c_void_p, C.c_int, C.c_longlong, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong),
                                  C.POINTER(C.c_longlong)]),
    "hfg_disc_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_void_p, C.c_void_p]),
    "hfg_disc_forward_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.c_void_p]),
    "hfg_disc_pair_losses": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_longlong, C.POINTER(C.c_float), C.c_void_p]),
    "hfg_disc_last_error": (C.c_char_p, [C.c_void_p]),
    "hfg_disc_destroy": (None, [C.c_void_p]),
}

_lib = None


def load():
    """Load the shared library (once) and bind the signatures.  Raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    # The .so is a git-ignored artefact.  When nvcc and the sources are here, (re)build it whenever the source digest
    # differs from the stamp of the binary (a stale library would be used silently otherwise); one process builds under a
    # file lock, the other ranks of a torchrun launch wait for it.  Without nvcc (a box that only received the prebuilt
    # library) the existing binary is used as shipped.  There is never a fallback to another implementation.
    from . import build as _build
    alt = os.environ.get("HFG_LIB")          # A/B experiments: a library built elsewhere (tools/build_variant.py); same ABI, no digest check
    if alt:
        lib = C.CDLL(alt)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib
    stale = not os.path.exists(LIB_PATH)
    if not stale and _build.have_nvcc():
        stale = not _build.is_current()
    if stale:
        try:
            _build.build_locked()
        except Exception as e:
            if not os.path.exists(LIB_PATH):
                raise ImportError(
                    f"{LIB_PATH} is missing and could not be built ({e}): run `python -m neuraltexttospeech_b200.build` "
                    "(or __graft_entry__.build()). There is no CPU / PyTorch fallback for the generator.") from e
            raise
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error(handle=None) -> str:
    msg = load().hfg_last_error(handle)
    return msg.decode() if msg else ""


def check(rc: int, handle=None, what: str = "", mel: bool = False, disc: bool = False):
    if rc == 0:
        return
    if mel or disc:
        m = (load().hfg_mel_last_error if mel else load().hfg_disc_last_error)(handle)
        msg = (m.decode() if m else "") or f"error {rc}"
    else:
        msg = last_error(handle) or f"error {rc}"
    if what:
        msg = f"{what}: {msg}"
    if rc in (HFG_ERR_INVALID, HFG_ERR_UNSUPPORTED):
        raise ValueError(msg)
    if rc == HFG_ERR_NOMEM:
        raise MemoryError(msg)
    raise RuntimeError(msg)
