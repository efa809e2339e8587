#!/bin/bash
# A/B of library builds on one box, then ship-test the winner: tools/ab_auto.sh build/ab/lib_base.so build/ab/lib_x.so ...
# Every library is copied over the in-tree one in turn (interleaved ROUNDS times) and benched on CFGS; the library with the
# best mean of the FIRST config (a candidate must beat the first-listed library by MARGIN, default 2 %) stays installed and
# gets the GPU tests, smoke() and the full default bench line.
mkdir -p gpurun_out
LIB=neuraltexttospeech_b200/libhifigan_b200.so
: > gpurun_out/ab_auto.txt
for round in $(seq 1 ${ROUNDS:-2}); do
for so in "$@"; do
  cp "$so" $LIB; label=$(basename $so .so)
  for cfg in ${CFGS:-v1_b64x10s v3_b64x10s}; do
    timeout 240 python bench.py --config $cfg --steps ${STEPS:-5} --no-cpu-baseline --no-e2e --no-adjacent > gpurun_out/ab_${label}_${cfg}_r$round.json 2> gpurun_out/ab_${label}_${cfg}_r$round.err
    python - <<PY | tee -a gpurun_out/ab_auto.txt
import json
try:
    d=json.load(open("gpurun_out/ab_${label}_${cfg}_r$round.json"))
    print("$label r$round $cfg value %.0f ms %.2f clk %s frac %.3f"%(d["value"], d["ms_per_step"], d["clocks"]["sm_mhz"], d["roofline_step"]["frac_of_per_stage_roofline"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("$label r$round $cfg FAILED", e, open("gpurun_out/ab_${label}_${cfg}_r$round.err").read()[-500:].replace("\n", " | "))
PY
  done
done
done
WIN=$(python - "$@" <<'PY'
import re, sys, os, collections
first_cfg = (os.environ.get("CFGS") or "v1_b64x10s v3_b64x10s").split()[0]
vals, failed = collections.defaultdict(list), set()
for line in open("gpurun_out/ab_auto.txt"):
    p = line.split()
    if len(p) > 3 and p[3] == "FAILED": failed.add(p[0])
    elif len(p) > 4 and p[2] == first_cfg: vals[p[0]].append(float(p[4]))
libs = sys.argv[1:]
label = lambda so: os.path.basename(so)[:-3]
mean = lambda so: sum(vals[label(so)]) / max(len(vals[label(so)]), 1)
best = libs[0]
for so in libs[1:]:
    if label(so) not in failed and mean(so) > mean(best) * (1 + float(os.environ.get("MARGIN", "0.02")) * (best == libs[0])): best = so
print(best)
PY
)
echo "winner: $WIN" | tee -a gpurun_out/ab_auto.txt
cp "$WIN" $LIB
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 240 python bench.py > gpurun_out/bench_v1.json 2> gpurun_out/bench_v1.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_v1.json
