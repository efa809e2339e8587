#!/bin/bash
mkdir -p gpurun_out
for cfg in v3_b64x10s v2_b256x10s v1_b1x256; do
  timeout 600 python bench.py --config $cfg --steps 3 --no-cpu-baseline ${EXTRA:-} > gpurun_out/cfg_$cfg.json 2> gpurun_out/cfg_$cfg.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/cfg_$cfg.json"))
    print("$cfg value %.0f e2e %.0f ms %.2f frac_stage_roof %.3f"%(d["value"], (d["e2e"] or {"value":0})["value"], d["ms_per_step"], d["roofline_step"]["frac_of_per_stage_roofline"]))
    for r in d["stages"]["rows"]: print("   ", r)
except Exception as e:
    print("$cfg failed", e, open("gpurun_out/cfg_$cfg.err").read()[-800:])
PY
done
