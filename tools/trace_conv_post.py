#!/usr/bin/env python
"""Role timeline of CTA 0 of the LAST kernel of a generator forward (conv_post on the tcgen05 path); needs a library built
with HFG_BUILD_TRACE=1.  Every conv_tc launch stamps the same buffer, so only the last launch's tiles are meaningful."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuraltexttospeech_b200 import Generator, HParams, _lib
from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights

version = sys.argv[1] if len(sys.argv) > 1 else "1"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
h = HParams.builtin(version)
gen = Generator(h, precision="fp16", device=0).load_weights(synthetic_weights(h, "A"))
mel = torch.from_numpy(synthetic_mel(B, 862, h.num_mels)).cuda()
lib = _lib.load()
trace = torch.zeros(7 * 64 * 8, dtype=torch.int64, device="cuda")
lib.hfgdbg_set_trace.argtypes = [C.c_void_p]
lib.hfgdbg_set_trace(C.c_void_p(trace.data_ptr()))
out = None
for _ in range(2):
    trace.zero_()
    out = gen._call_device(mel, out)
torch.cuda.synchronize()
t = trace.cpu().numpy()[:6 * 64 * 8].reshape(6, 64, 8)
names = {0: "slab", 1: "stg", 2: "mma", 3: "store", 4: "epi0", 5: "epi1"}
ev = {0: ["issue"], 1: ["grant"], 2: ["arrive", "acc_free", "slab_in", "issued"], 3: ["first", "last"],
      4: ["arrive", "acc_full", "stg_rdy", "done", "ld", "staged", "fenced", "handed"], 5: ["arrive", "acc_full", "stg_rdy", "done", "ld", "staged", "fenced", "handed"]}
t0 = t[2, 0, 0]
print(f"V{version} B={B}: conv_post CTA 0, clocks relative to the MMA warp's arrival at tile 0")
for tile in range(0, 11):
    parts = []
    for r in (0, 2, 4, 5, 3):
        parts.append(names[r] + "[" + " ".join(f"{n}={int(t[r, tile, i] - t0)}" for i, n in enumerate(ev[r]) if t[r, tile, i] > 0) + "]")
    print(f"tile {tile}: " + " ".join(parts))
print("clocks per tile (MMA issued, tiles 2..9):", (t[2, 9, 3] - t[2, 2, 3]) / 7.0)
