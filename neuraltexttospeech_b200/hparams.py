"""Hyper-parameter carrier for the HiFi-GAN generator drop-in.

Mirrors the reference's config object (reference: HiFiGAN_TF/hparams.py:9-41 — a
dataclass with 25 required fields built as ``HParams(**json.load(f))``,
HiFiGAN_TF/train.py:44-47).  The generator path only reads six of them
(HiFiGAN_TF/model.py:119-143) plus ``num_mels``; the rest are carried so that the
reference's construction idiom keeps working verbatim.
"""
from __future__ import annotations

import dataclasses
import json
import os
from typing import Any, Mapping, Optional

_CONFIG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "configs")

# Fields the generator forward pass actually consumes.
GENERATOR_FIELDS = (
    "resblock",
    "upsample_rates",
    "upsample_kernel_sizes",
    "upsample_initial_channel",
    "resblock_kernel_sizes",
    "resblock_dilation_sizes",
    "num_mels",
)


@dataclasses.dataclass
class HParams:
    # -- training knobs (unused by the forward pass, kept for construction parity)
    resblock: str
    num_gpus: int
    batch_size: int
    learning_rate: float
    adam_b1: float
    adam_b2: float
    lr_decay: float
    seed: int
    # -- generator topology
    upsample_rates: list
    upsample_kernel_sizes: list
    upsample_initial_channel: int
    resblock_kernel_sizes: list
    resblock_dilation_sizes: list
    # -- audio front-end
    segment_size: int
    num_mels: int
    num_freq: int
    n_fft: int
    hop_size: int
    win_size: int
    sampling_rate: int
    fmin: int
    fmax: int
    fmax_for_loss: Optional[int]
    num_workers: int
    dist_config: dict

    @classmethod
    def from_json(cls, path: str) -> "HParams":
        with open(path, "r") as f:
            return cls(**json.load(f))

    @classmethod
    def builtin(cls, version: str) -> "HParams":
        """Load one of the three stock configs ('v1', 'v2', 'v3')."""
        version = str(version).lower().lstrip("v")
        if version not in ("1", "2", "3"):
            raise ValueError(f"unknown HiFi-GAN config version {version!r}")
        return cls.from_json(os.path.join(_CONFIG_DIR, f"config_v{version}.json"))


def coerce_hparams(obj: Any) -> Any:
    """Accept an HParams, any attribute object, a mapping, or a json path.

    Returns an object exposing the GENERATOR_FIELDS as attributes.  ``num_mels``
    defaults to 80 when the carrier lacks it (the reference lets Keras infer the
    input width at first call, HiFiGAN_TF/model.py:121-125).
    """
    if isinstance(obj, (str, os.PathLike)):
        obj = HParams.from_json(os.fspath(obj))
    if isinstance(obj, Mapping):
        missing = [k for k in GENERATOR_FIELDS if k not in obj and k != "num_mels"]
        if missing:
            raise ValueError(f"hparams mapping lacks generator fields: {missing}")
        ns = _Namespace(**{k: obj[k] for k in obj})
        obj = ns
    for k in GENERATOR_FIELDS:
        if k == "num_mels":
            continue
        if not hasattr(obj, k):
            raise ValueError(f"hparams object lacks generator field {k!r}")
    if not hasattr(obj, "num_mels"):
        obj = _Overlay(obj, num_mels=80)
    return obj


class _Namespace:
    def __init__(self, **kw):
        self.__dict__.update(kw)


class _Overlay:
    def __init__(self, base, **kw):
        self._base = base
        self.__dict__.update(kw)

    def __getattr__(self, name):
        return getattr(self._base, name)
