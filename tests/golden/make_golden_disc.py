"""Regenerates tests/golden/disc_upstream_{mpd,msd}.npz -- run ONLY in the build container (needs /root/reference):
    python tests/golden/make_golden_disc.py

The fixtures are outputs of the torch discriminators the reference vendors (FastPitch_TF/DeepLearningExamples/PyTorch/
SpeechSynthesis/HiFi-GAN/hifigan/models.py:273-409: DiscriminatorP / MultiPeriodDiscriminator / DiscriminatorS /
MultiScaleDiscriminator), imported from the reference tree, unmodified, and run on the CPU on the seeded synthetic
weights of oracle/disc_oracle.synthetic_weights (variant "B"; weights are NOT stored, they are re-derived from the seed)
and a seeded waveform.  oracle/disc_oracle.py in compat="upstream" must reproduce them (tests/test_disc_oracle.py):
that pins the layer list, grouped / strided conv placement, the period reshape and reflect padding, the weight-norm
fold axes and the pooling windows.  The Keras-only semantics of compat="reference" (TensorFlow SAME padding with a
stride, the "valid" first DiscriminatorS conv, average pooling that ignores padding) are pinned by nothing runnable here.

Stored per discriminator: the score (full) and, per feature map, its shape, sum, sum of |.| and 97 samples at fixed
strided positions of the flattened channels-last map (whole maps would be megabytes).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import disc_oracle as D       # noqa: E402

REF = "/root/reference"
B, T = 2, 1031                            # odd and prime: every period reflect-pads, every stride sees a remainder


def waveform(b=B, t=T, seed=99):
    rng = np.random.default_rng(seed)
    n = np.arange(t)[None, :]
    f = rng.uniform(0.01, 0.2, size=(b, 1))
    return (0.6 * np.sin(2 * np.pi * f * n) + 0.2 * rng.standard_normal((b, t))).astype(np.float32)


def sample_positions(n):
    return (np.arange(97) * max(n // 97, 1)) % n


def summarise(fmap):
    flat = np.asarray(fmap, np.float64).reshape(-1)
    return np.array(fmap.shape, np.int64), np.array([flat.sum(), np.abs(flat).sum()]), flat[sample_positions(flat.size)]


def _vendored_module():
    base = os.path.join(REF, "FastPitch_TF/DeepLearningExamples/PyTorch/SpeechSynthesis/HiFi-GAN")
    common = types.ModuleType("common")                       # the module imports common.*; none of it is arithmetic used here
    common.monkey_patch = types.ModuleType("common.monkey_patch")
    stft = types.ModuleType("common.stft"); stft.STFT = object
    utils = types.ModuleType("common.utils")
    utils.AttrDict = dict
    utils.init_weights = lambda m, mean=0.0, std=0.01: None
    utils.get_padding = lambda k, d=1: int((k * d - d) / 2)
    sys.modules.update({"common": common, "common.monkey_patch": common.monkey_patch, "common.stft": stft, "common.utils": utils})
    import importlib.util
    spec = importlib.util.spec_from_file_location("dle_hifigan_models", os.path.join(base, "hifigan/models.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _load(conv, w, b, two_d):
    """Folded kernel [k, cin_g, cout] -> the torch module's effective weight (weight norm / spectral norm removed first)."""
    import torch
    wt = np.transpose(w, (2, 1, 0))                            # [cout, cin_g, k]
    if two_d:
        wt = wt[..., None]                                     # Conv2d((k, 1)): [cout, cin, k, 1]
    assert tuple(conv.weight.shape) == tuple(wt.shape), (conv.weight.shape, wt.shape)
    conv.weight.data.copy_(torch.from_numpy(np.ascontiguousarray(wt)).float())
    conv.bias.data.copy_(torch.from_numpy(b).float())


def _strip_norm(module):
    import torch
    from torch.nn.utils import remove_spectral_norm, remove_weight_norm
    for m in module.modules():
        if isinstance(m, (torch.nn.Conv1d, torch.nn.Conv2d)):
            try:
                remove_weight_norm(m)
            except ValueError:
                remove_spectral_norm(m)


def run(kind):
    import torch
    mod = _vendored_module()
    model = mod.MultiPeriodDiscriminator(list(D.MPD_PERIODS)) if kind == "mpd" else mod.MultiScaleDiscriminator()
    _strip_norm(model)
    weights = D.synthetic_weights(kind, "B")
    for i, d in enumerate(model.discriminators):
        for l, conv in enumerate(d.convs):
            w, b = D.fold(weights, f"discriminators.{i}.convs.{l}")
            _load(conv, w, b.astype(np.float32), kind == "mpd")
        w, b = D.fold(weights, f"discriminators.{i}.conv_post")
        _load(d.conv_post, w, b.astype(np.float32), kind == "mpd")
    model = model.double().eval()                              # fp64 so the fixture is not limited by torch's fp32 conv order
    y = torch.from_numpy(waveform()).double()[:, None, :]
    with torch.no_grad():
        scores, _, fmaps, _ = model(y, y)
    out = {"B": B, "T": T}
    for i, (s, fm) in enumerate(zip(scores, fmaps)):
        out[f"score_{i}"] = s.numpy().astype(np.float64)
        for l, f in enumerate(fm):
            f = f.numpy()
            f = np.transpose(f, (0, 2, 3, 1)) if kind == "mpd" else np.transpose(f, (0, 2, 1))   # channels-last like the reference
            out[f"shape_{i}_{l}"], out[f"sums_{i}_{l}"], out[f"samples_{i}_{l}"] = summarise(f)
    path = os.path.join(HERE, f"disc_upstream_{kind}.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    for kind in (sys.argv[1:] or ["mpd", "msd"]):
        run(kind)
