"""Layer enumeration of the HiFi-GAN generator (what the reference's
``Generator.__init__`` builds, HiFiGAN_TF/model.py:116-152), as plain data.

Every weight-normalised layer is described by a :class:`LayerSpec`; the order is
the reference's construction/call order and defines the synthetic-weight seed
index: conv_pre, ups[0], resblocks[0..2] (convs1[0], convs2[0],
convs1[1], ...), ups[1], ..., conv_post.
"""
from __future__ import annotations

import dataclasses
from typing import List, Tuple

from .utils import get_padding


@dataclasses.dataclass(frozen=True)
class LayerSpec:
    name: str            # e.g. "resblocks.4.convs1.2"
    kind: str            # "conv" | "convT"
    kernel: int
    c_in: int
    c_out: int
    dilation: int = 1
    stride: int = 1      # upsample rate for convT
    padding: str = "same"

    @property
    def v_shape(self) -> Tuple[int, int, int]:
        """Keras kernel layout: Conv1D [k, Cin, Cout]; Conv1DTranspose [k, Cout, Cin]."""
        if self.kind == "conv":
            return (self.kernel, self.c_in, self.c_out)
        return (self.kernel, self.c_out, self.c_in)

    @property
    def g_shape(self) -> Tuple[int]:
        """tfa WeightNormalization: g spans the kernel's LAST axis."""
        return (self.v_shape[-1],)

    @property
    def b_shape(self) -> Tuple[int]:
        return (self.c_out,)


def stage_channels(h) -> List[int]:
    return [h.upsample_initial_channel // (2 ** (i + 1)) for i in range(len(h.upsample_rates))]


def total_upsample(h) -> int:
    u = 1
    for r in h.upsample_rates:
        u *= int(r)
    return u


def is_resblock1(h) -> bool:
    # reference: `ResBlock1 if hparams.resblock == '1' else ResBlock2` (model.py:126)
    return h.resblock == "1"


def used_dilations(h) -> List[List[int]]:
    """The dilation entries the reference's resblocks consume: ResBlock1 reads dilation[0..2] (reference
    HiFiGAN_TF/model.py:21-66), ResBlock2 dilation[0..1] (model.py:84-101).  Longer lists are truncated exactly like
    the C engine does (generator._config_from_hparams); shorter ones are an error there."""
    need = 3 if is_resblock1(h) else 2
    return [[int(d) for d in list(ds)[:need]] for ds in h.resblock_dilation_sizes]


def enumerate_layers(h) -> List[LayerSpec]:
    layers: List[LayerSpec] = []
    c0 = int(h.upsample_initial_channel)
    layers.append(LayerSpec("conv_pre", "conv", 7, int(h.num_mels), c0, 1, 1, "same"))
    chans = stage_channels(h)
    nk = len(h.resblock_kernel_sizes)
    c_prev = c0
    for i, (u, ku) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
        c = chans[i]
        layers.append(LayerSpec(f"ups.{i}", "convT", int(ku), c_prev, c, 1, int(u), "same"))
        for j, (k, dils) in enumerate(zip(h.resblock_kernel_sizes, used_dilations(h))):
            n = i * nk + j
            if is_resblock1(h):
                for m, d in enumerate(dils):
                    layers.append(LayerSpec(f"resblocks.{n}.convs1.{m}", "conv", int(k), c, c,
                                            int(d), 1, get_padding(int(k), int(d))))
                    layers.append(LayerSpec(f"resblocks.{n}.convs2.{m}", "conv", int(k), c, c,
                                            1, 1, get_padding(int(k), 1)))
            else:
                for m, d in enumerate(dils):
                    layers.append(LayerSpec(f"resblocks.{n}.convs.{m}", "conv", int(k), c, c,
                                            int(d), 1, get_padding(int(k), int(d))))
        c_prev = c
    layers.append(LayerSpec("conv_post", "conv", 7, c_prev, 1, 1, 1, "same"))
    return layers


def count_macs_per_frame(h) -> int:
    """Dense MACs per input mel frame (SURVEY.md §8a invariants)."""
    macs = 0
    rate = 1
    for L in enumerate_layers(h):
        if L.kind == "convT":
            macs += rate * L.kernel * L.c_in * L.c_out
            rate *= L.stride
        else:
            macs += rate * L.kernel * L.c_in * L.c_out
    return macs


def count_params(h) -> int:
    return sum(L.kernel * L.c_in * L.c_out + L.c_out for L in enumerate_layers(h))


def boundary_bytes_per_frame(h, act_bytes: int = 2) -> int:
    """Layer-boundary activation bytes per mel frame."""
    total = 0
    rate = 1
    for L in enumerate_layers(h):
        if L.kind == "convT":
            total += act_bytes * (rate * L.c_in + rate * L.stride * L.c_out)
            rate *= L.stride
        else:
            res = L.c_out if (".convs2." in L.name or ".convs." in L.name) else 0
            total += act_bytes * rate * (L.c_in + L.c_out + res)
    return total


def per_stage_work(h, act_bytes: int = 2):
    """Per mel frame, per measurement stage of hfg_stage_times_ms: [conv_pre, stage 0.., conv_post] ->
    list of dicts {label, macs, bytes, launches} (FLOP = 2*MAC, layer-boundary bytes).
    `launches` counts weight-bearing layers (one kernel launch each on the tcgen05 path)."""
    out = []
    rate = 1
    cur = None
    for L in enumerate_layers(h):
        if L.name == "conv_pre" or L.name == "conv_post" or L.kind == "convT":
            label = L.name if L.kind == "conv" else f"stage{L.name.split('.')[1]}"
            cur = {"label": label, "macs": 0, "bytes": 0, "launches": 0}
            out.append(cur)
        if L.kind == "convT":
            cur["macs"] += rate * L.kernel * L.c_in * L.c_out
            cur["bytes"] += act_bytes * (rate * L.c_in + rate * L.stride * L.c_out)
            rate *= L.stride
        else:
            res = L.c_out if (".convs2." in L.name or ".convs." in L.name) else 0
            cur["macs"] += rate * L.kernel * L.c_in * L.c_out
            cur["bytes"] += act_bytes * rate * (L.c_in + L.c_out + res)
        cur["launches"] += 1
    return out
