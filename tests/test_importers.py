"""Weight importers, checked on the CPU against the oracle's own fold: a torch-format
state dict (weight_g / weight_v, the upstream checkpoint format the reference's vocoder loader reads) and a
duck-typed Keras generator must both map to exactly the tensors `Generator.load_weights` takes."""
import types

import numpy as np
import pytest

from neuraltexttospeech_b200 import HParams
from neuraltexttospeech_b200.importers import export_keras_generator, from_upstream_state_dict
from neuraltexttospeech_b200.synthetic import synthetic_weights
from neuraltexttospeech_b200.topology import enumerate_layers
from oracle import hifigan_oracle as O


def _torch_style(h, w, style):
    """Our Keras-layout weights -> a fake upstream state dict in one of the three on-disk styles."""
    sd = {}
    nk = len(h.resblock_kernel_sizes)
    folded = O.fold_all(h, w, np.float64)
    for L in enumerate_layers(h):
        name = L.name
        if style == "dle" and name.startswith("resblocks."):
            p = name.split(".")
            name = ".".join(["resblocks", str(int(p[1]) // nk), str(int(p[1]) % nk)] + p[2:])
        v = np.transpose(w[L.name + ".v"], (2, 1, 0)).copy()
        g = w[L.name + ".g"].reshape(-1, 1, 1).copy()
        if style == "weight_norm" or style == "dle":
            sd[name + ".weight_v"], sd[name + ".weight_g"] = v, g
        elif style == "parametrize":
            sd[name + ".parametrizations.weight.original1"], sd[name + ".parametrizations.weight.original0"] = v, g
        else:   # "removed": plain folded weight, torch layout
            kern, _ = folded[L.name]
            sd[name + ".weight"] = np.transpose(kern, (2, 1, 0)).astype(np.float32).copy()
        sd[name + ".bias"] = w[L.name + ".b"].copy()
    return {"generator": sd}


@pytest.mark.parametrize("v", "123")
@pytest.mark.parametrize("style", ["weight_norm", "parametrize", "dle"])
def test_upstream_state_dict_round_trips(v, style):
    h = HParams.builtin(v)
    w = synthetic_weights(h, "B")
    got = from_upstream_state_dict(_torch_style(h, w, style), h)
    assert set(got) == set(w)
    for k in w:
        assert got[k].shape == w[k].shape and (got[k] == w[k]).all(), k


def test_checkpoint_saved_after_remove_weight_norm_gives_the_same_folded_kernels():
    h = HParams.builtin("2")
    w = synthetic_weights(h, "B")
    got = from_upstream_state_dict(_torch_style(h, w, "removed"), h)
    a, b = O.fold_all(h, w, np.float64), O.fold_all(h, got, np.float64)
    for name in a:
        assert np.abs(a[name][0] - b[name][0]).max() < 1e-6
    with pytest.raises(KeyError):
        from_upstream_state_dict({"conv_pre.weight_v": np.zeros((1, 1, 1))}, h)
    bad = _torch_style(h, w, "weight_norm")
    bad["generator"]["conv_pre.weight_v"] = np.zeros((3, 3, 3), np.float32)
    with pytest.raises(ValueError):
        from_upstream_state_dict(bad, h)


def test_importer_agrees_with_the_real_vendored_torch_module_names():
    """Names and shapes come from torch itself: build torch Conv1d / ConvTranspose1d with weight_norm like
    upstream's Generator does and import its state_dict."""
    import torch
    from torch.nn.utils import weight_norm
    h = HParams.builtin("3")

    class G(torch.nn.Module):
        def __init__(self):
            super().__init__()
            c0 = h.upsample_initial_channel
            self.conv_pre = weight_norm(torch.nn.Conv1d(80, c0, 7, padding=3))
            self.ups = torch.nn.ModuleList()
            self.resblocks = torch.nn.ModuleList()
            for i, (u, k) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
                self.ups.append(weight_norm(torch.nn.ConvTranspose1d(c0 >> i, c0 >> (i + 1), k, u, padding=(k - u) // 2)))
                for kk, ds in zip(h.resblock_kernel_sizes, h.resblock_dilation_sizes):
                    rb = torch.nn.Module()
                    rb.convs = torch.nn.ModuleList([weight_norm(torch.nn.Conv1d(c0 >> (i + 1), c0 >> (i + 1), kk, dilation=d)) for d in ds[:2]])
                    self.resblocks.append(rb)
            self.conv_post = weight_norm(torch.nn.Conv1d(c0 >> len(h.upsample_rates), 1, 7, padding=3))

    torch.manual_seed(0)
    g = G()
    w = from_upstream_state_dict(g.state_dict(), h)
    # the oracle's fold of the imported (v, g) must equal torch's own folded weight
    folded = O.fold_all(h, w, np.float64)
    for name, mod in (("conv_pre", g.conv_pre), ("ups.1", g.ups[1]), ("resblocks.4.convs.1", g.resblocks[4].convs[1]), ("conv_post", g.conv_post)):
        want = mod.weight.detach().numpy()                         # torch recomputes g * v / ||v||
        assert np.abs(np.transpose(folded[name][0], (2, 1, 0)) - want).max() < 1e-6, name


def test_export_keras_generator_duck_typed():
    h = HParams.builtin("1")
    w = synthetic_weights(h, "B")

    def wrap(name):
        return types.SimpleNamespace(v=w[name + ".v"], g=w[name + ".g"], layer=types.SimpleNamespace(bias=w[name + ".b"], kernel=None))

    nk = len(h.resblock_kernel_sizes)
    gen = types.SimpleNamespace(
        hparams=h, conv_pre=wrap("conv_pre"), conv_post=wrap("conv_post"),
        ups=[wrap(f"ups.{i}") for i in range(len(h.upsample_rates))],
        resblocks=[types.SimpleNamespace(convs1=[wrap(f"resblocks.{n}.convs1.{m}") for m in range(3)],
                                         convs2=[wrap(f"resblocks.{n}.convs2.{m}") for m in range(3)])
                   for n in range(nk * len(h.upsample_rates))])
    got = export_keras_generator(gen)
    assert set(got) == set(w) and all((got[k] == w[k]).all() for k in w)
    gen.ups = gen.ups[:-1]
    with pytest.raises(KeyError):
        export_keras_generator(gen)


@pytest.mark.parametrize("v", "13")
def test_real_upstream_generator_state_dict_imports_and_reproduces_its_forward(v):
    """Container-only (needs the vendored upstream torch generator under /root/reference; skipped elsewhere):
    random-init upstream Generator WITH weight norm -> state_dict -> importer -> oracle(compat=upstream) must
    reproduce the upstream module's own forward."""
    import os
    import sys
    if not os.path.isdir("/root/reference/FastSpeech2_TF/FastSpeech2/hifigan"):
        pytest.skip("reference tree not present")
    import torch
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden
    h = HParams.builtin(v)
    torch.manual_seed(int(v))
    gen = make_golden._upstream_generator(h, keep_weight_norm=True)
    with torch.no_grad():                                          # upstream inits N(0, 0.01): perturb g, b so the fold matters
        for n, p in gen.named_parameters():
            if n.endswith("weight_g") or "original0" in n:
                p.mul_(1.0 + 0.3 * torch.rand_like(p))
            if n.endswith("bias"):
                p.add_(0.02 * torch.randn_like(p))
    w = from_upstream_state_dict({"generator": gen.state_dict()}, h)
    rng = np.random.default_rng(5)
    mel = rng.standard_normal((1, 6, 80)).astype(np.float32)
    with torch.no_grad():
        y = gen(torch.from_numpy(mel).permute(0, 2, 1).contiguous()).permute(0, 2, 1).numpy()
    mine = O.generator_forward(h, w, mel, "numpy64", "upstream")
    assert np.abs(mine - y).max() < 2e-5
