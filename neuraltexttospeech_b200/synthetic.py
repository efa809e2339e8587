"""Deterministic synthetic mels and weights.

There is no trained HiFi-GAN checkpoint anywhere in the reference tree
(README.md:53-56), so parity and benchmarks run on random-init weights drawn the
way Keras would (Glorot-uniform kernels, HiFiGAN_TF/model.py:26-29 defaults) and on
mels in the range of log-compressed spectrograms
(HiFiGAN_TF/common/audio_processing_tf.py:27-32).
"""
from __future__ import annotations

import math
from typing import Dict

import numpy as np

from .topology import enumerate_layers

MEL_SEED = 1234          # = "seed" of config_v*.json
MEL_FLOOR = math.log(1e-5)


def synthetic_mel(batch: int, frames: int, num_mels: int = 80, seed: int = MEL_SEED) -> np.ndarray:
    rng = np.random.default_rng(seed)
    mel = rng.standard_normal((batch, frames, num_mels), dtype=np.float32) * 2.0 - 5.0
    return np.clip(mel, MEL_FLOOR, 2.0).astype(np.float32)


def _norm_over_leading(v: np.ndarray) -> np.ndarray:
    return np.sqrt(np.sum(np.square(v.astype(np.float64)), axis=(0, 1)))


def synthetic_weights(hparams, variant: str = "A", seed: int = MEL_SEED) -> Dict[str, np.ndarray]:
    """variant 'A': g = ||v||, b = 0 (TF random-init after the first call).
    variant 'B': g = ||v||*U(0.5,1.5), b ~ U(-0.05,0.05) (exercises fold + bias)."""
    if variant not in ("A", "B"):
        raise ValueError("variant must be 'A' or 'B'")
    out: Dict[str, np.ndarray] = {}
    for idx, L in enumerate(enumerate_layers(hparams)):
        rng = np.random.default_rng(seed + idx)
        lim = math.sqrt(6.0 / (L.kernel * L.c_in + L.kernel * L.c_out))
        v = rng.uniform(-lim, lim, size=L.v_shape).astype(np.float32)
        g = _norm_over_leading(v)
        b = np.zeros(L.b_shape, dtype=np.float32)
        if variant == "B":
            g = g * rng.uniform(0.5, 1.5, size=g.shape)
            b = rng.uniform(-0.05, 0.05, size=L.b_shape).astype(np.float32)
        out[f"{L.name}.v"] = v
        out[f"{L.name}.g"] = g.astype(np.float32)
        out[f"{L.name}.b"] = b
    return out


def synthetic_disc_weights(kind: str, variant: str = "A", seed: int = 4321) -> Dict[str, np.ndarray]:
    """Seeded Glorot-uniform weights for the discriminators (kind "mpd" / "msd"; discriminators.py), drawn layer by layer
    like `synthetic_weights`.  The first DiscriminatorS is SpectralNormalization-wrapped in the reference
    (HiFiGAN_TF/model.py:353) and gets a plain kernel `w`; every other layer gets WeightNormalization's (v, g, b)."""
    from .discriminators import MPD_PERIODS, _P_LAYERS, _S_LAYERS
    if kind not in ("mpd", "msd") or variant not in ("A", "B"):
        raise ValueError("kind must be 'mpd' or 'msd', variant 'A' or 'B'")
    n_sub, layers = (len(MPD_PERIODS), _P_LAYERS) if kind == "mpd" else (3, _S_LAYERS)
    out: Dict[str, np.ndarray] = {}
    n = 0
    for i in range(n_sub):
        for l, (cin, cout, k, groups) in enumerate(layers):
            name = f"discriminators.{i}.convs.{l}" if l + 1 < len(layers) else f"discriminators.{i}.conv_post"
            rng = np.random.default_rng(seed + n)
            n += 1
            lim = math.sqrt(6.0 / (k * (cin // groups) + k * cout))
            v = rng.uniform(-lim, lim, size=(k, cin // groups, cout)).astype(np.float32)
            nrm = _norm_over_leading(v)
            g = nrm.copy()
            b = np.zeros(cout, dtype=np.float32)
            if variant == "B":
                g = g * rng.uniform(0.5, 1.5, size=g.shape)
                b = rng.uniform(-0.05, 0.05, size=cout).astype(np.float32)
            if kind == "msd" and i == 0:
                out[name + ".w"] = (v * (g / nrm)).astype(np.float32)
            else:
                out[name + ".v"] = v
                out[name + ".g"] = g.astype(np.float32)
            out[name + ".b"] = b
    return out
