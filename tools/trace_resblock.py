#!/usr/bin/env python
"""Per-role timeline of resblock_tc_kernel's CTA 0 (needs a library built with HFG_BUILD_TRACE=1).
usage: trace_resblock.py K DIL [FOLD] [B] [ROWS]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from neuraltexttospeech_b200 import _lib
from neuraltexttospeech_b200.ops import resblock_fused

k, dil = int(sys.argv[1]), int(sys.argv[2])
f = int(sys.argv[3]) if len(sys.argv) > 3 else 1
B = int(sys.argv[4]) if len(sys.argv) > 4 else 16
rows = int(sys.argv[5]) if len(sys.argv) > 5 else 110336
lib = _lib.load()
rng = np.random.default_rng(0)


def taps(kk, dd):
    pl = dd * (kk - 1) if (kk == 3 and dd == 1) else dd * (kk - 1) // 2
    return sorted({(s + j * dd - pl) // f for s in range(f) for j in range(kk)})


o1, o2 = taps(k, dil), taps(k, 1)
x = torch.from_numpy(rng.standard_normal((B, rows, 64)).astype(np.float32)).cuda().half()
w1 = (rng.standard_normal((len(o1), 64, 64)) * 0.03).astype(np.float32)
w2 = (rng.standard_normal((len(o2), 64, 64)) * 0.03).astype(np.float32)
b = np.zeros(64, np.float32)
out = torch.empty_like(x)
trace = torch.zeros(7 * 64 * 8, dtype=torch.int64, device="cuda")
lib.hfgdbg_set_trace.argtypes = [C.c_void_p]
lib.hfgdbg_set_trace(C.c_void_p(trace.data_ptr()))
for _ in range(2):
    trace.zero_()
    resblock_fused(x, w1, b, o1, w2, b, o2, out=out)
torch.cuda.synchronize()
t = trace.cpu().numpy().reshape(7, 64, 8)
t0 = t[t > 0].min()
names = ["slab", "mma", "store", "mid0", "mid1", "fin0", "fin1"]
ev = {0: ["req"], 1: ["c1_turn", "slab_in", "c1_issued", "c2_turn", "mid_done", "c2_issued"], 2: ["blk0", "blk1", "read"],
      3: ["arrive", "acc1", "free", "done"], 4: ["arrive", "acc1", "free", "done"], 5: ["arrive", "acc2", "", "done"], 6: ["arrive", "acc2", "", "done"]}
print(f"k={k} d={dil} fold={f} taps {len(o1)}+{len(o2)} B={B} rows={rows}; clocks relative to the first stamp")
for tile in range(6, 14):
    parts = []
    for r in range(7):
        grp = [g for g in range(3) if t[r, tile, 4 + g] > 0] if r >= 3 else []
        parts.append(names[r] + (f"(g{grp[0]})" if grp else "") + "[" + " ".join(f"{n}={int(t[r, tile, i] - t0)}" for i, n in enumerate(ev[r]) if n and t[r, tile, i] > 0) + "]")
    print(f"tile {tile}:\n   " + "\n   ".join(parts))
per_tile = (t[1, 40, 5] - t[1, 8, 5]) / 32.0
print(f"steady-state clocks per tile (conv2 issue to conv2 issue, tiles 8..40): {per_tile:.0f}")
for r in (3, 4, 5, 6):
    d = [(t[r, i, 3] - t[r, i, 1]) for i in range(8, 40) if t[r, i, 3] > 0 and t[r, i, 1] > 0]
    w = [(t[r, i, 1] - t[r, i, 0]) for i in range(8, 40) if t[r, i, 1] > 0 and t[r, i, 0] > 0]
    print(f"{names[r]}: pass (accumulator seen -> done) mean {np.mean(d):.0f} clk; wait for the accumulator mean {np.mean(w):.0f} clk")
d = [(t[1, i, 1] - t[0, i, 0]) for i in range(8, 40)]
print(f"slab: request -> landed (seen by the MMA warp) mean {np.mean(d):.0f} clk")
d = [(t[1, i, 4] - t[1, i, 3]) for i in range(8, 40)]
print(f"mma: wait for mid_done mean {np.mean(d):.0f} clk;  wait for slab mean {np.mean([(t[1, i, 1] - t[1, i, 0]) for i in range(8, 40)]):.0f} clk")
