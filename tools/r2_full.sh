#!/bin/bash
# final-build visit: GPU tests, smoke, bench lines of every BASELINE config, reference arm, CPU port batch effect
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_v1_b64x10s.json 2> gpurun_out/bench_v1.err; echo "bench rc=$?"
for cfg in v1_b1x256 v2_b256x10s v3_b64x10s v1_b512x10s; do
  timeout 900 python bench.py --config $cfg --steps ${STEPS:-5} --no-cpu-baseline > gpurun_out/bench_$cfg.json 2> gpurun_out/bench_$cfg.err; echo "$cfg rc=$?"
done
timeout 600 python bench.py --precision tf32 --steps 5 --no-cpu-baseline --no-adjacent > gpurun_out/bench_v1_b64x10s_tf32.json 2> gpurun_out/bench_tf32.err; echo "tf32 rc=$?"
timeout 600 python bench.py --precision bf16 --steps 5 --no-cpu-baseline --no-adjacent > gpurun_out/bench_v1_b64x10s_bf16.json 2> gpurun_out/bench_bf16.err; echo "bf16 rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_*.json')):
    try:
        d=json.load(open(f))
        e=(d.get('e2e') or {}).get('value')
        rs=d.get('roofline_step',{})
        print(f.split('/')[-1], "value %.0f e2e %s ms %.3f frac_stage %s clk %s"%(d['value'], ("%.0f"%e) if e else None, d['ms_per_step'], rs.get('frac_of_per_stage_roofline'), (d.get('clocks') or {}).get('sm_mhz')))
        for r in d.get('stages',{}).get('rows',[]): print("    ", r['label'], r['ms'], r['launches'], r['bound'], r['frac'])
    except Exception as ex:
        print(f, "failed", ex)
PY
