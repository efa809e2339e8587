#!/bin/bash
# A/B of environment knobs on one box: tools/ab_env.sh "label:K=V,K=V" ... ; CFGS="v1_b64x10s v3_b64x10s" selects configs
mkdir -p gpurun_out
if [ "${RUN_TESTS:-1}" = "1" ]; then timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
for spec in "$@"; do
  label=${spec%%:*}; envs=${spec#*:}
  for cfg in ${CFGS:-v1_b64x10s}; do
    ( IFS=','; for kv in $envs; do [ -n "$kv" ] && export "$kv"; done
      timeout 300 python bench.py --config $cfg --steps ${STEPS:-5} --no-cpu-baseline --no-e2e > gpurun_out/ab_${label}_$cfg.json 2> gpurun_out/ab_${label}_$cfg.err )
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_${label}_$cfg.json"))
    print("$label $cfg value %.0f ms %.2f clk %s frac %.3f"%(d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"], d["roofline_step"]["frac_of_per_stage_roofline"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("$label $cfg failed", e, open("gpurun_out/ab_${label}_$cfg.err").read()[-800:])
PY
  done
done
