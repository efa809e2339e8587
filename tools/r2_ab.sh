#!/bin/bash
# tests + A/B bench lines (env knob per run) + the other configs
mkdir -p gpurun_out
if [ "${SKIP_TESTS:-0}" != "1" ]; then
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
fi
i=0
for kv in ${AB:-HFG_RB_PREF=0 HFG_RB_PREF=1}; do
  env $kv timeout 300 python bench.py --steps 5 --no-cpu-baseline --no-adjacent ${BENCH_ARGS:-} > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_$i.json"))
    print("$kv value %.0f e2e %.0f ms %.2f clk %s"%(d["value"], (d["e2e"] or {"value":0})["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("$kv failed", e, open("gpurun_out/ab_$i.err").read()[-800:])
PY
  i=$((i+1))
done
if [ "${SKIP_CFG:-0}" != "1" ]; then bash tools/other_configs.sh; fi
