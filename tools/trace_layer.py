#!/usr/bin/env python
"""Per-role timeline of conv_tc_kernel's CTA 0 for one layer shape (needs a library built with HFG_BUILD_TRACE=1).
usage: trace_layer.py C K DIL [RES] [B] [ROWS]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuraltexttospeech_b200 import _lib
from neuraltexttospeech_b200.ops import tapped_conv

c, k, dil = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
use_res = len(sys.argv) > 4 and sys.argv[4] == "1"
B = int(sys.argv[5]) if len(sys.argv) > 5 else 16
rows = int(sys.argv[6]) if len(sys.argv) > 6 else 220672 * 32 // c
lib = _lib.load()
rng = np.random.default_rng(0)
x = torch.from_numpy(rng.standard_normal((B, rows, c)).astype(np.float32)).cuda().half()
w = (rng.standard_normal((k, c, c)) * 0.05).astype(np.float32)
bias = np.zeros(c, np.float32)
taps = [j * dil - dil * (k - 1) // 2 for j in range(k)]
res = x.clone() if use_res else None
out = torch.empty_like(x)
trace = torch.zeros(6 * 64 * 8, dtype=torch.int64, device="cuda")
lib.hfgdbg_set_trace.argtypes = [C.c_void_p]
lib.hfgdbg_set_trace(C.c_void_p(trace.data_ptr()))
for _ in range(2):
    trace.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tapped_conv(x, w, bias, taps, engine="tc", res=res, res_inv=10.0, out_slope=0.1, out=out)
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(6, 64, 8)
t0 = t[t > 0].min()
names = {0: "slab", 1: "stg", 2: "mma", 3: "store", 4: "epi0", 5: "epi1"}
ev = {0: ["issue"], 1: ["grant"], 2: ["arrive", "acc_free", "slab_in", "issued"], 3: ["first", "last"], 4: ["arrive", "acc_full", "stg_rdy", "done", "ld", "staged", "fenced", "handed"], 5: ["arrive", "acc_full", "stg_rdy", "done", "ld", "staged", "fenced", "handed"]}
print(f"C={c} k={k} d={dil} res={use_res} B={B} rows={rows}; clocks relative to the first stamp")
for tile in range(4, 14):
    parts = []
    for r in (0, 1, 2, 4, 5, 3):
        parts.append(names[r] + "[" + " ".join(f"{n}={int(t[r, tile, i] - t0)}" for i, n in enumerate(ev[r]) if t[r, tile, i] > 0) + "]")
    print(f"tile {tile}: " + " ".join(parts))
per_tile = (t[2, 40, 3] - t[2, 8, 3]) / 32.0
print(f"steady-state clocks per tile (MMA issue to MMA issue, tiles 8..40): {per_tile:.0f}")
