#!/usr/bin/env python
"""Host-call latency of small batches (Generator.call on numpy): ms per call, V1."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
v = sys.argv[1] if len(sys.argv) > 1 else "1"
h = HParams.builtin(v); gen = Generator(h).load_weights(synthetic_weights(h, "A"))
for B, T in ((1, 64), (1, 256), (4, 256), (1, 862), (8, 862)):
    mel = synthetic_mel(B, T, 80)
    for _ in range(5): gen(mel)
    t0 = time.perf_counter()
    for _ in range(50): gen(mel)
    ms = (time.perf_counter() - t0) / 50 * 1e3
    print(f"v{v} B {B} T {T} ms/call {ms:.3f}  audio-s/s {B * T * 256 / 22050 / ms * 1e3:.0f}  (HFG_CHAIN_FRAMES={os.environ.get('HFG_CHAIN_FRAMES', 'default')})")
