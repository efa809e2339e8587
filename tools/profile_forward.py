#!/usr/bin/env python
"""One warm-up forward + one profiled forward of the generator on device-resident mels (for ncu).

Launch order of one forward (V1): cast_pad, conv_pre, then per stage: ups[i] + 18 ResBlock1 convs
(k=3: 6, k=7: 6, k=11: 6), then conv_post.  conv_tc launches per forward: 1 + 4*19 = 77.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ap = argparse.ArgumentParser()
ap.add_argument("--version", default="1")
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--frames", type=int, default=862)
ap.add_argument("--precision", default="fp16")
ap.add_argument("--forwards", type=int, default=2)
args = ap.parse_args()

import torch
from neuraltexttospeech_b200 import Generator, HParams
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights

h = HParams.builtin(args.version)
gen = Generator(h, precision=args.precision, device=0).load_weights(synthetic_weights(h, "A"))
mel = torch.from_numpy(synthetic_mel(args.batch, args.frames, h.num_mels)).cuda()
out = None
for _ in range(args.forwards):
    out = gen._call_device(mel, out)
torch.cuda.synchronize()
print("ok", tuple(out.shape), float(out.abs().max()), gen.last_launch_count())
