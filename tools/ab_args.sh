#!/bin/bash
# A/B of bench.py arguments on one box: tools/ab_args.sh "label:--config v2_b256x10s --micro-batch 64" ...
mkdir -p gpurun_out
for spec in "$@"; do
  label=${spec%%:*}; args=${spec#*:}
  timeout 300 python bench.py $args --steps ${STEPS:-5} --no-cpu-baseline --no-e2e --no-adjacent > gpurun_out/ab_${label}.json 2> gpurun_out/ab_${label}.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_${label}.json"))
    print("$label value %.0f ms %.2f clk %s frac %.3f"%(d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"], d["roofline_step"]["frac_of_per_stage_roofline"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("$label failed", e, open("gpurun_out/ab_${label}.err").read()[-800:])
PY
done
