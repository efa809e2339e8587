#!/usr/bin/env python
"""Tiny end-to-end run for compute-sanitizer: wav -> mel front-end -> generator (fp16 tcgen05 path and fp32 path) -> PCM."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from neuraltexttospeech_b200 import Generator, HParams, STFT
from neuraltexttospeech_b200.synthetic import synthetic_weights
rng = np.random.default_rng(0)
wav = (0.3 * rng.standard_normal((2, 1024 + 256 * 5))).astype(np.float32)
fe = STFT()
mel = fe.mel_spectrogram(wav)
for v in "123":
    h = HParams.builtin(v)
    for prec in ("fp16", "fp32"):
        g = Generator(h, precision=prec).load_weights(synthetic_weights(h, "B"))
        a = g(mel); p = g.synthesize_pcm16(mel)
        for _ in range(2): g(mel)          # capture + replay
        print(v, prec, a.shape, float(np.abs(a).max()), int(np.abs(p).max()))
        g.close()
fe.close()
