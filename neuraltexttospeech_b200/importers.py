"""Weight importers for the drop-in `Generator`.

The reference ships no trained HiFi-GAN weights (its own save path, HiFiGAN_TF/gan.py:199-203, is "Not
Implemented" per HiFiGAN_TF/README.md), so real audio needs one of:

  (i)  `export_keras_generator(gen)` -- walks a live reference `Generator` (HiFiGAN_TF/model.py:116-152:
       `conv_pre`, `ups[i]`, `resblocks[n].convs1/convs2/convs[m]`, `conv_post`, each a
       tfa.layers.WeightNormalization with variables `g`, `v` and `layer.bias`) into the name -> array
       mapping `Generator.load_weights` takes.  Duck-typed: anything with those attributes works, so it
       runs (and is tested) without TensorFlow.
  (ii) `from_upstream_state_dict(sd, hparams)` -- the upstream jik876/hifi-gan PyTorch checkpoint format the
       reference's vocoder loader reads (FastSpeech2_TF/FastSpeech2/utils/model.py:57-69:
       `ckpt["generator"]`, `weight_g` / `weight_v` / `bias`; also torch >= 2.1 parametrizations and
       checkpoints saved after remove_weight_norm()).  Use with `Generator(..., compat="upstream")`: upstream
       weights were trained with symmetric padding and final slope 0.01, not the reference's quirks.

Layout conversions (PyTorch -> Keras, what `load_weights` expects):
    Conv1d           weight_v [Cout, Cin, k] -> v [k, Cin, Cout];   weight_g [Cout,1,1] -> g [Cout]
    ConvTranspose1d  weight_v [Cin, Cout, k] -> v [k, Cout, Cin];   weight_g [Cin,1,1]  -> g [Cin]
torch.nn.utils.weight_norm (dim=0) normalises over every axis but the first, tfa over every axis but the
last: after the transpose both are "per Cout" for Conv1d and "per Cin" for ConvTranspose1d -- the same fold.
"""
from __future__ import annotations

from typing import Dict, Mapping

import numpy as np

from .hparams import coerce_hparams
from .topology import enumerate_layers


def _np(t) -> np.ndarray:
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    elif hasattr(t, "numpy") and not isinstance(t, np.ndarray):
        t = t.numpy()
    return np.ascontiguousarray(np.asarray(t, dtype=np.float32))


def _upstream_key(sd: Mapping[str, object], name: str, num_kernels: int) -> str:
    """Our layer name -> the module prefix used by the checkpoint (NVIDIA DLE groups resblocks per stage)."""
    if any(k.startswith(name + ".") for k in sd):
        return name
    parts = name.split(".")
    if parts[0] == "resblocks":
        n = int(parts[1])
        alt = ".".join(["resblocks", str(n // num_kernels), str(n % num_kernels)] + parts[2:])
        if any(k.startswith(alt + ".") for k in sd):
            return alt
    raise KeyError(f"checkpoint has no parameters for layer {name!r}")


def from_upstream_state_dict(state_dict: Mapping[str, object], hparams) -> Dict[str, np.ndarray]:
    h = coerce_hparams(hparams)
    sd = state_dict["generator"] if "generator" in state_dict and not hasattr(state_dict["generator"], "shape") else state_dict
    nk = len(h.resblock_kernel_sizes)
    out: Dict[str, np.ndarray] = {}
    for L in enumerate_layers(h):
        p = _upstream_key(sd, L.name, nk)
        if p + ".weight_v" in sd:
            v, g = _np(sd[p + ".weight_v"]), _np(sd[p + ".weight_g"]).reshape(-1)
        elif p + ".parametrizations.weight.original1" in sd:
            v, g = _np(sd[p + ".parametrizations.weight.original1"]), _np(sd[p + ".parametrizations.weight.original0"]).reshape(-1)
        elif p + ".weight" in sd:               # saved after remove_weight_norm(): g = ||v|| reproduces the kernel
            v = _np(sd[p + ".weight"])
            g = np.sqrt((v.astype(np.float64) ** 2).sum(axis=(1, 2))).astype(np.float32)
        else:
            raise KeyError(f"{p}: none of weight_v / parametrizations.weight.original1 / weight found")
        if v.ndim != 3:
            raise ValueError(f"{p}: expected a 3-D conv kernel, got shape {v.shape}")
        v = np.ascontiguousarray(np.transpose(v, (2, 1, 0)))       # both conv kinds: reverse the axes
        b = _np(sd[p + ".bias"]) if p + ".bias" in sd else np.zeros(L.c_out, np.float32)
        for suffix, arr, shape in (("v", v, L.v_shape), ("g", g, L.g_shape), ("b", b, L.b_shape)):
            if tuple(arr.shape) != tuple(shape):
                raise ValueError(f"{p}: {suffix} has shape {tuple(arr.shape)}, config expects {tuple(shape)}")
            out[f"{L.name}.{suffix}"] = arr
    return out


def load_upstream_checkpoint(path: str, hparams) -> Dict[str, np.ndarray]:
    """`generator_*.pth.tar` of upstream hifi-gan (read like FastSpeech2_TF/FastSpeech2/utils/model.py:62-66)."""
    import torch
    ckpt = torch.load(path, map_location="cpu", weights_only=True)
    return from_upstream_state_dict(ckpt, hparams)


def export_keras_generator(gen, hparams=None) -> Dict[str, np.ndarray]:
    """Live reference `Generator` (built, i.e. called once) -> mapping for `load_weights` / `np.savez`."""
    h = coerce_hparams(hparams if hparams is not None else gen.hparams)
    out: Dict[str, np.ndarray] = {}

    def take(name: str, wrapper) -> None:
        v = getattr(wrapper, "v", None)
        if v is None:
            v = wrapper.layer.kernel
        out[name + ".v"] = _np(v)
        out[name + ".g"] = _np(wrapper.g).reshape(-1)
        out[name + ".b"] = _np(wrapper.layer.bias)

    take("conv_pre", gen.conv_pre)
    for i, up in enumerate(gen.ups):
        take(f"ups.{i}", up)
    for n, rb in enumerate(gen.resblocks):
        if h.resblock == "1":
            for m, (c1, c2) in enumerate(zip(rb.convs1, rb.convs2)):
                take(f"resblocks.{n}.convs1.{m}", c1)
                take(f"resblocks.{n}.convs2.{m}", c2)
        else:
            for m, c in enumerate(rb.convs):
                take(f"resblocks.{n}.convs.{m}", c)
    take("conv_post", gen.conv_post)
    for L in enumerate_layers(h):
        for suffix, shape in (("v", L.v_shape), ("g", L.g_shape), ("b", L.b_shape)):
            key = f"{L.name}.{suffix}"
            if key not in out:
                raise KeyError(f"generator has no variable for {key}")
            if tuple(out[key].shape) != tuple(shape):
                raise ValueError(f"{key}: shape {tuple(out[key].shape)} != expected {tuple(shape)}")
    return out
