#!/usr/bin/env python
"""Discriminator forward (hfg_disc_*) timing at the training-step shape of the reference (config_v1.json: batch_size 16,
segment_size 8192; gan.py runs every discriminator on y and y_hat -> 32 waveforms per step).  Device-resident input,
CUDA events on torch's current stream (the stream the forward is launched on), 3 warm-ups, 10 timed forwards."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--samples", type=int, default=8192)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--once", action="store_true", help="one warm-up + one forward per model (for ncu)")
args = ap.parse_args()

import numpy as np
import torch
from neuraltexttospeech_b200.discriminators import MultiPeriodDiscriminator, MultiScaleDiscriminator, disc_flops
from neuraltexttospeech_b200.synthetic import synthetic_disc_weights

rng = np.random.default_rng(0)
y = torch.from_numpy(rng.standard_normal((args.batch, args.samples)).astype(np.float32) * 0.3).cuda()
rows = {}
for kind, cls in (("mpd", MultiPeriodDiscriminator), ("msd", MultiScaleDiscriminator)):
    m = cls().load_weights(synthetic_disc_weights(kind, "A"))
    for _ in range(1 if args.once else 3):
        m.forward(y)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    n = 1 if args.once else args.iters
    e0.record()
    for _ in range(n):
        m.forward(y)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    fl, by = disc_flops(m, args.batch, args.samples)
    rows[kind] = {"ms": round(ms, 4), "gflop": round(fl / 1e9, 2), "tflops": round(fl / ms / 1e9, 2), "map_mbytes": round(by / 1e6, 1),
                  "waveforms": args.batch, "samples": args.samples}
print(json.dumps(rows))
