#!/bin/bash
mkdir -p gpurun_out
for mb in 16 32 64; do
  HFG_TC_SKIP=0x60 timeout 300 python bench.py --steps 5 --no-cpu-baseline --micro-batch $mb > gpurun_out/mb_$mb.json 2> gpurun_out/mb_$mb.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/mb_$mb.json"))
    print("mb $mb value %.0f e2e %.0f ms %.2f clk %s"%(d["value"], d["e2e"]["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("mb $mb failed", e, open("gpurun_out/mb_$mb.err").read()[-600:])
PY
done
