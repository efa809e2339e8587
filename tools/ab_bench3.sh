#!/bin/bash
mkdir -p gpurun_out
export HFG_TC_SKIP=0x60
HFG_TC_DIRECT=3 timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for cfg in "0 64" "1 64" "2 64" "3 64" "3 128" "1 256"; do
  set -- $cfg
  HFG_TC_DIRECT=$1 HFG_TC_DIRECT_MAXC=$2 timeout 300 python bench.py --steps 5 --no-cpu-baseline --no-e2e --micro-batch 32 > gpurun_out/d_$1_$2.json 2> gpurun_out/d_$1_$2.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/d_$1_$2.json"))
    print("direct $1 maxc $2 value %.0f ms %.2f clk %s"%(d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("direct $1 $2 failed", e, open("gpurun_out/d_$1_$2.err").read()[-600:])
PY
done
