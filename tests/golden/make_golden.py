"""Regenerates the committed golden fixtures in this directory.

Run ONLY in the build container (needs /root/reference for the upstream cross
check):   python tests/golden/make_golden.py

Fixtures
  oracle_v{1,2,3}_{A,B}.npz  mel + fp64-oracle audio (stored fp32) + a few interior taps'
                              checksums, for synthetic weight variants A/B at B=2, T=9.
                              Weights are NOT stored: they are re-derived from the seed by
                              neuraltexttospeech_b200.synthetic.synthetic_weights.
  upstream_v{1,2,3}.npz       audio produced by the *importable upstream PyTorch generator*
                              that the reference vendors (FastSpeech2_TF/FastSpeech2/hifigan/
                              models.py:112-165 for ResBlock1 configs; NVIDIA DLE
                              PyTorch/SpeechSynthesis/HiFi-GAN/hifigan/models.py for the
                              ResBlock2 config) on the same folded weights.  The oracle in
                              compat="upstream" mode must reproduce it (tests/test_oracle.py).
                              This pins conv / transposed-conv placement, the weight-norm
                              fold axes, the MRF order and the residual wiring.  The two
                              reference quirks (causal k=3 convs, slope 0.3) are NOT covered
                              by any runnable reference code here: "parity unpinned" for them.
  quirks_v{1,2,3}.npz         a SECOND, independently written derivation of the two reference quirks, run on
                              the vendored torch modules themselves (not on the oracle's conv code):
                              Q1 -- every ResBlock Conv1d for which the reference's own get_padding
                              (HiFiGAN_TF/utils.py:29-37) answers "causal" is wrapped so that the vendored,
                              symmetrically padded torch conv sees its input delayed by p = d(k-1)/2 samples
                              (p zeros prepended, last p outputs dropped): y_sym[t - p] = sum_j x[t + j*d - 2p] W[j],
                              which is Keras' causal Conv1D (left pad d(k-1), right pad 0);
                              Q2 -- the module-level `F` of the vendored file is replaced by a shim whose
                              leaky_relu default slope is Keras' LeakyReLU() default 0.3 (HiFiGAN_TF/model.py:152),
                              so the vendored forward() runs unmodified.
                              The oracle in compat="reference" must reproduce it (tests/test_oracle.py) and so
                              must the CUDA fp32 mode (tests/test_generator_gpu.py).  TensorFlow itself is still
                              not in the loop: the TF boundary stays "parity unpinned".
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from neuraltexttospeech_b200.hparams import HParams            # noqa: E402
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights  # noqa: E402
from oracle import hifigan_oracle as O                          # noqa: E402

REF = "/root/reference"


def _upstream_generator(h, keep_weight_norm=False):
    """Build the vendored upstream torch generator (weight norm removed unless keep_weight_norm)."""
    import torch
    if h.resblock == "1":
        sys.path.insert(0, os.path.join(REF, "FastSpeech2_TF/FastSpeech2"))
        # the package __init__ pulls the whole FastSpeech2 tree; load the module file directly
        import importlib.util
        spec = importlib.util.spec_from_file_location(
            "upstream_hifigan_models", os.path.join(REF, "FastSpeech2_TF/FastSpeech2/hifigan/models.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        gen = mod.Generator(h)
    else:
        base = os.path.join(REF, "FastPitch_TF/DeepLearningExamples/PyTorch/SpeechSynthesis/HiFi-GAN")
        # DLE's module imports common.{monkey_patch,stft,utils}; stub what is not arithmetic
        common = types.ModuleType("common")
        common.monkey_patch = types.ModuleType("common.monkey_patch")
        stft = types.ModuleType("common.stft"); stft.STFT = object
        utils = types.ModuleType("common.utils")

        class AttrDict(dict):
            def __init__(self, *a, **k):
                super().__init__(*a, **k); self.__dict__ = self

        utils.AttrDict = AttrDict
        utils.init_weights = lambda m, mean=0.0, std=0.01: None
        utils.get_padding = lambda k, d=1: int((k * d - d) / 2)
        sys.modules.update({"common": common, "common.monkey_patch": common.monkey_patch,
                            "common.stft": stft, "common.utils": utils})
        import importlib.util
        spec = importlib.util.spec_from_file_location("dle_hifigan_models", os.path.join(base, "hifigan/models.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        gen = mod.Generator(dict(vars(h)))
        for group in gen.resblocks:          # DLE's ResBlock2 lists lrelu_slope in __constants__
            for rb in group:                 # but never assigns it; give it the module constant
                if not hasattr(rb, "lrelu_slope"):
                    rb.lrelu_slope = mod.LRELU_SLOPE
    if not keep_weight_norm:
        gen.remove_weight_norm()
    gen._vendored_module = mod
    return gen.eval()


def _load_folded(gen, h, weights):
    import torch
    folded = O.fold_all(h, weights, np.float64)

    def put(module, name, kind):
        w, b = folded[name]
        wt = torch.from_numpy(np.ascontiguousarray(np.transpose(w, (2, 1, 0)))).float()
        assert tuple(module.weight.shape) == tuple(wt.shape), (name, module.weight.shape, wt.shape)
        module.weight.data.copy_(wt)
        module.bias.data.copy_(torch.from_numpy(b).float())

    put(gen.conv_pre, "conv_pre", "conv")
    put(gen.conv_post, "conv_post", "conv")
    for i, up in enumerate(gen.ups):
        put(up, f"ups.{i}", "convT")
    blocks = list(gen.resblocks)
    if blocks and isinstance(blocks[0], torch.nn.Sequential):      # DLE groups them per stage
        blocks = [rb for group in blocks for rb in group]
    for n, rb in enumerate(blocks):
        if h.resblock == "1":
            for m in range(3):
                put(rb.convs1[m], f"resblocks.{n}.convs1.{m}", "conv")
                put(rb.convs2[m], f"resblocks.{n}.convs2.{m}", "conv")
        else:
            for m in range(2):
                put(rb.convs[m], f"resblocks.{n}.convs.{m}", "conv")


def _reference_padding_string(k, d):
    """reference HiFiGAN_TF/utils.py:29-37, restated here on purpose (this script must not lean on the oracle for Q1)."""
    pad = int((k * d - d) / 2)
    return "valid" if pad == 0 else ("causal" if pad == 1 else "same")


def _apply_reference_quirks(gen, mod):
    """Turn the vendored upstream torch generator into the reference's arithmetic without touching its forward()."""
    import torch
    import torch.nn.functional as TF

    class DelayedInput(torch.nn.Module):          # symmetric conv on a delayed input == Keras causal conv
        def __init__(self, conv):
            super().__init__()
            self.conv = conv
            k, d = conv.kernel_size[0], conv.dilation[0]
            self.p = d * (k - 1) // 2
            assert conv.padding[0] == self.p and conv.stride[0] == 1

        def forward(self, x):
            n = x.shape[-1]
            return self.conv(TF.pad(x, (self.p, 0)))[..., :n]

    blocks = list(gen.resblocks)
    if blocks and isinstance(blocks[0], torch.nn.Sequential):
        blocks = [rb for group in blocks for rb in group]
    n_wrapped = 0
    for rb in blocks:
        for lst in [getattr(rb, a) for a in ("convs1", "convs2", "convs") if hasattr(rb, a)]:
            for m in range(len(lst)):
                c = lst[m]
                if _reference_padding_string(c.kernel_size[0], c.dilation[0]) == "causal":
                    lst[m] = DelayedInput(c)
                    n_wrapped += 1

    class KerasDefaultsF:                          # `F` as the vendored file sees it
        def __getattr__(self, name):
            return getattr(TF, name)

        @staticmethod
        def leaky_relu(x, negative_slope=0.3):     # keras.layers.LeakyReLU() default alpha (model.py:152)
            return TF.leaky_relu(x, negative_slope)

    mod.F = KerasDefaultsF()
    return n_wrapped


def make_adjacent():
    """adjacent_frontend.npz: seeded waveform + the fp64 oracle's log-mel (stored fp32) and its 16-bit PCM."""
    from oracle import adjacent_oracle as A
    rng = np.random.default_rng(4321)
    n = 1024 + 256 * 9 + 17
    t = np.arange(n)[None, :] / 22050.0
    f0 = np.array([[110.0], [196.0]])
    wav = sum((0.3 / h_) * np.sin(2 * np.pi * f0 * h_ * t + h_) for h_ in range(1, 24)) + 0.005 * rng.standard_normal((2, n))
    wav = (wav / np.abs(wav).max() * 0.9).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "adjacent_frontend.npz"), wav=wav,
                        mel=A.mel_spectrogram(wav.astype(np.float64)).astype(np.float32), pcm=A.pcm16(wav))


def main():
    make_adjacent()
    import torch
    torch.manual_seed(0)
    B, T = 2, 9
    for v in "123":
        h = HParams.builtin(v)
        mel = synthetic_mel(B, T, h.num_mels, seed=1234 + int(v))
        for variant in "AB":
            w = synthetic_weights(h, variant)
            taps = {}
            audio = O.generator_forward(h, w, mel, "numpy64", "reference", taps=taps)
            np.savez_compressed(
                os.path.join(HERE, f"oracle_v{v}_{variant}.npz"),
                mel=mel, audio=audio.astype(np.float32),
                conv_pre_sum=np.float64(taps["conv_pre"].sum()),
                stage0_sum=np.float64(taps["stage.0"].sum()),
                stage_last_abs_sum=np.float64(np.abs(taps[f"stage.{len(h.upsample_rates) - 1}"]).sum()))
        # upstream cross-check fixture (variant B: non-trivial g and b)
        w = synthetic_weights(h, "B")
        if os.path.isdir(REF):
            gen = _upstream_generator(h)
            _load_folded(gen, h, w)
            with torch.no_grad():
                y = gen(torch.from_numpy(mel).permute(0, 2, 1).contiguous())   # NCT in, [B,1,L] out
            y = y.permute(0, 2, 1).contiguous().numpy()
            mine = O.generator_forward(h, w, mel, "numpy64", "upstream")
            err = float(np.abs(mine - y).max())
            print(f"v{v}: upstream torch vs oracle(compat=upstream) max-abs {err:.3e}; amp {np.abs(y).max():.3f}")
            assert err < 2e-5, err
            np.savez_compressed(os.path.join(HERE, f"upstream_v{v}.npz"), mel=mel, audio=y.astype(np.float32))
            # second derivation of the reference quirks Q1/Q2 on the vendored modules (see the header)
            gen = _upstream_generator(h)
            _load_folded(gen, h, w)
            nw = _apply_reference_quirks(gen, gen._vendored_module)
            mel_q = synthetic_mel(B, 13, h.num_mels, seed=4321 + int(v))
            with torch.no_grad():
                yq = gen(torch.from_numpy(mel_q).permute(0, 2, 1).contiguous()).permute(0, 2, 1).contiguous().numpy()
            mine = O.generator_forward(h, w, mel_q, "numpy64", "reference")
            up = O.generator_forward(h, w, mel_q, "numpy64", "upstream")
            err = float(np.abs(mine - yq).max())
            print(f"v{v}: vendored torch + quirks ({nw} causal convs) vs oracle(compat=reference) max-abs {err:.3e}; "
                  f"distance to compat=upstream {np.abs(up - yq).max():.3e}")
            assert err < 2e-5, err
            assert np.abs(up - yq).max() > 1e-3            # the quirks matter: the fixture tells the two modes apart
            np.savez_compressed(os.path.join(HERE, f"quirks_v{v}.npz"), mel=mel_q, audio=yq.astype(np.float32), causal_convs=np.int64(nw))


if __name__ == "__main__":
    main()
