#!/usr/bin/env python
"""Tiny end-to-end run for compute-sanitizer: wav -> mel front-end -> generator -> PCM, for all three configs and
every precision (fp16: tcgen05 conv_tc + fused resblock_tc kernels incl. CTA pairs; tf32: conv_tf32; fp32: direct
kernels), host path (plain, graph capture, graph replay) and device path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from neuraltexttospeech_b200 import Generator, HParams, STFT
from neuraltexttospeech_b200.synthetic import synthetic_weights
rng = np.random.default_rng(0)
wav = (0.3 * rng.standard_normal((2, 1024 + 256 * 5))).astype(np.float32)
fe = STFT()
mel = fe.mel_spectrogram(wav)
for v in "123":
    h = HParams.builtin(v)
    for prec in (os.environ.get("SAN_PRECISIONS", "fp16,tf32,fp32").split(",")):
        g = Generator(h, precision=prec).load_weights(synthetic_weights(h, "B"))
        a = g(mel); p = g.synthesize_pcm16(mel)
        for _ in range(2): g(mel)          # capture + replay
        d = g(torch.from_numpy(mel).cuda()).cpu().numpy()
        print(v, prec, a.shape, float(np.abs(a).max()), int(np.abs(p).max()), bool(np.array_equal(a, d)), g.last_launch_count())
        g.close()
fe.close()
