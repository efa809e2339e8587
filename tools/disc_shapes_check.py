#!/usr/bin/env python
"""Discriminator forward against the CPU oracle on awkward shapes (very short waveforms where sequences are shorter than a
conv's stride, a single waveform, a long one): prints the worst map-relative difference per case."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from neuraltexttospeech_b200.discriminators import MultiPeriodDiscriminator, MultiScaleDiscriminator
from neuraltexttospeech_b200.synthetic import synthetic_disc_weights
from oracle import disc_oracle as D

rng = np.random.default_rng(3)
cases = {"mpd": [(1, 12), (1, 37), (5, 100), (3, 131), (1, 40000)], "msd": [(1, 300), (4, 333), (2, 1000), (1, 40000)]}
worst_all = 0.0
for kind, cls in (("mpd", MultiPeriodDiscriminator), ("msd", MultiScaleDiscriminator)):
    w = synthetic_disc_weights(kind, "B")
    for compat in ("reference", "upstream"):
        m = cls(compat=compat).load_weights(w)
        for B, T in cases[kind]:
            y = (rng.standard_normal((B, T)) * 0.4).astype(np.float32)
            try:
                s, f = m.forward(y)
            except ValueError as e:
                print(kind, compat, B, T, "rejected:", str(e)[:90]); continue
            os_, of = (D.mpd_forward if kind == "mpd" else D.msd_forward)(w, y, compat=compat)
            worst = 0.0
            for fm, ofm in zip(f, of):
                for a, b in zip(fm, ofm):
                    assert a.shape == b.shape, (a.shape, b.shape)
                    worst = max(worst, float(np.abs(np.asarray(a) - b).max() / max(np.abs(b).max(), 1e-30)))
            worst_all = max(worst_all, worst)
            print(kind, compat, B, T, "worst %.2e" % worst, "OK" if worst <= 2e-4 else "FAIL")
print("overall worst %.2e" % worst_all)
sys.exit(0 if worst_all <= 2e-4 else 1)
