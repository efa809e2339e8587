#!/bin/bash
# A/B of library builds on one box: tools/ab_lib.sh build/lib_a.so build/lib_b.so ... ; each is copied over the in-tree
# library in turn (interleaved ROUNDS times); CFGS selects bench configs; the LAST one stays installed and is tested.
mkdir -p gpurun_out
LIB=neuraltexttospeech_b200/libhifigan_b200.so
for round in $(seq 1 ${ROUNDS:-2}); do
for so in "$@"; do
  cp "$so" $LIB; label=$(basename $so .so)
  for cfg in ${CFGS:-v1_b64x10s}; do
    timeout 300 python bench.py --config $cfg --steps ${STEPS:-5} --no-cpu-baseline --no-e2e --no-adjacent > gpurun_out/ab_${label}_$cfg.json 2> gpurun_out/ab_${label}_$cfg.err
    python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_${label}_$cfg.json"))
    print("$label r$round $cfg value %.0f ms %.2f clk %s frac %.3f"%(d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"], d["roofline_step"]["frac_of_per_stage_roofline"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("$label $cfg failed", e, open("gpurun_out/ab_${label}_$cfg.err").read()[-800:])
PY
  done
done
done
if [ "${RUN_TESTS:-1}" = "1" ]; then timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; fi
