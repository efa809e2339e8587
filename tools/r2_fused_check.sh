#!/bin/bash
# Fused ResBlock kernel bring-up: op-level parity first (isolates failures), then the generator suite, then A/B bench.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -m gpu -x -q -k "resblock_fused" > gpurun_out/ops_fused.log 2>&1; echo "ops fused rc=$?"; tail -25 gpurun_out/ops_fused.log
if [ "${ONLY_OPS:-0}" = "1" ]; then exit 0; fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
for fuse in 1 0; do
  HFG_FUSE=$fuse timeout 300 python bench.py --steps 5 --no-cpu-baseline --no-adjacent > gpurun_out/bench_fuse$fuse.json 2> gpurun_out/bench_fuse$fuse.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_fuse$fuse.json"))
    print("fuse=$fuse value %.0f e2e %.0f ms %.2f clk %s"%(d["value"], (d["e2e"] or {"value":0})["value"], d["ms_per_step"], d["clocks"]["sm_mhz"]), [r["ms"] for r in d["stages"]["rows"]])
except Exception as e:
    print("fuse=$fuse failed", e, open("gpurun_out/bench_fuse$fuse.err").read()[-800:])
PY
done
