#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python - <<'PY'
import numpy as np, sys
sys.path.insert(0, '.')
from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
from oracle import hifigan_oracle as O
for v in "123":
    h = HParams.builtin(v); w = synthetic_weights(h, "B"); mel = synthetic_mel(2, 37, h.num_mels)
    ref = O.generator_forward(h, w, mel, "torch32")
    g = Generator(h, precision="fp16", device=0).load_weights(w)
    out = g(mel)
    err = float(np.abs(out - ref).max()); snr = 10 * np.log10((ref ** 2).sum() / ((out - ref) ** 2).sum())
    eng = g.layer_engines()
    print("V%s fp16 max-abs %.2e snr %.1f dB engines %s" % (v, err, snr, {e: list(eng.values()).count(e) for e in set(eng.values())}))
PY
EXTRA=--no-e2e bash tools/other_configs.sh 2>&1 | grep -v "^    {'label': 'conv_pre'"
