#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --precision tf32 --steps 5 --no-cpu-baseline --no-adjacent > gpurun_out/bench_v1_b64x10s_tf32.json 2> gpurun_out/bench_tf32.err; echo "tf32 rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_v1_b64x10s_tf32.json'))
print("tf32 value %.0f e2e %.0f ms %.2f"%(d['value'], d['e2e']['value'], d['ms_per_step']), [ (r['label'], r['ms'], r['frac']) for r in d['stages']['rows']])
PY
