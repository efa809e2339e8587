#!/bin/bash
# Same-box A/B of bench.py runs (box-to-box variance is +-5 %, so only runs of ONE gpurun call compare).  Each spec is
#     label:ITEM,ITEM,...      ITEM = KEY=VALUE            environment knob (HFG_RB_MB=1, HFG_FUSE=0, ...)
#                                   | lib=PATH            library variant (tools/build_variant.py), loaded through HFG_LIB
#                                   | --flag / --flag=V   extra bench.py argument (--micro-batch=32)
# e.g.  tools/ab.sh base: nofuse:HFG_FUSE=0 deep:lib=build/variants/deep/libhifigan_b200.so mb32:--micro-batch=32
# ROUNDS (2) interleaves the specs; CFGS ("v1_b64x10s") selects bench configs; STEPS (5); E2E=1 keeps the end-to-end leg;
# RUN_TESTS=1 runs the GPU tests first.  One summary line per run goes to stdout and gpurun_out/ab_summary.txt.
mkdir -p gpurun_out
: > gpurun_out/ab_summary.txt
if [ "${RUN_TESTS:-0}" = "1" ]; then timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log; fi
extra="--no-cpu-baseline --no-adjacent"; [ "${E2E:-0}" = "1" ] || extra="$extra --no-e2e"
for round in $(seq 1 ${ROUNDS:-2}); do
for spec in "$@"; do
  label=${spec%%:*}; items=${spec#*:}; envs=(); args=()
  IFS=',' read -ra parts <<< "$items"
  for it in "${parts[@]}"; do
    case "$it" in
      "") ;;
      lib=*) envs+=("HFG_LIB=$(realpath "${it#lib=}")") ;;
      --*=*) args+=("${it%%=*}" "${it#*=}") ;;
      --*) args+=("$it") ;;
      *=*) envs+=("$it") ;;
    esac
  done
  for cfg in ${CFGS:-v1_b64x10s}; do
    out=gpurun_out/ab_${label}_${cfg}_r$round
    env "${envs[@]}" timeout 300 python bench.py --config $cfg --steps ${STEPS:-5} $extra "${args[@]}" > $out.json 2> $out.err
    python - "$label" "$round" "$cfg" "$out" <<'PY' | tee -a gpurun_out/ab_summary.txt
import json, sys
label, rnd, cfg, out = sys.argv[1:]
try:
    d = json.load(open(out + ".json"))
    e2e = (d.get("e2e") or {}).get("value") or 0
    print("%s r%s %s value %.0f e2e %.0f ms %.2f clk %s frac %.3f" % (label, rnd, cfg, d["value"], e2e, d["ms_per_step"], d["clocks"]["sm_mhz"],
          d["roofline_step"]["frac_of_per_stage_roofline"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print(label, "r" + rnd, cfg, "FAILED", e, open(out + ".err").read()[-500:].replace("\n", " | "))
PY
  done
done
done
