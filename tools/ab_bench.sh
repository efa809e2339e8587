#!/bin/bash
# A/B of conv_tc table rows on one box: HFG_TC_SKIP masks (bit i = skip row i of HFG_TC_TABLE)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for mask in ${MASKS:-0 0x70 0x40 0x10}; do
  HFG_TC_SKIP=$mask timeout 300 python bench.py --steps 5 --no-cpu-baseline --no-e2e ${BENCH_ARGS:-} > gpurun_out/ab_$mask.json 2> gpurun_out/ab_$mask.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_$mask.json"))
    print("mask $mask value %.0f ms %.2f clk %s"%(d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("mask $mask failed", e, open("gpurun_out/ab_$mask.err").read()[-600:])
PY
done
