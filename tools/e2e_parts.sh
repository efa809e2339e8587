#!/bin/bash
# e2e (host-buffer) throughput against the micro-batch size of hfg_forward_host: tools/e2e_parts.sh "v3_b64x10s:0,16,8,4" ...
mkdir -p gpurun_out
for spec in "$@"; do
  cfg=${spec%%:*}
  for mb in $(echo ${spec#*:} | tr ',' ' '); do
    timeout 300 python bench.py --config $cfg --steps ${STEPS:-10} --micro-batch $mb --no-cpu-baseline --no-adjacent > gpurun_out/e2e_${cfg}_mb$mb.json 2> gpurun_out/e2e_${cfg}_mb$mb.err
    python - <<PY | tee -a gpurun_out/e2e_parts.txt
import json
try:
    d=json.load(open("gpurun_out/e2e_${cfg}_mb$mb.json"))
    print("$cfg micro_batch $mb (host path %s): device %.0f e2e %.0f audio-s/s (%.1f %%)"%(d["e2e"]["micro_batch"], d["value"], d["e2e"]["value"], 100*d["e2e"]["value"]/d["value"]))
except Exception as e:
    print("$cfg mb $mb failed", e)
PY
  done
done
