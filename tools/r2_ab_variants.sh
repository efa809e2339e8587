#!/bin/bash
# same-box A/B of library variants (build/variants/<name>/libhifigan_b200.so): launch list of the fused kernels, twice each, interleaved
mkdir -p gpurun_out
for round in 1 2; do
for v in ${VARIANTS:-rr deep}; do
  envs="HFG_LIB=$PWD/build/variants/${v%%:*}/libhifigan_b200.so"
  case "$v" in *:*) envs="$envs ${v#*:}";; esac
  env $envs python tools/profile_forward.py > gpurun_out/plain_$v.log 2>&1 && \
  env $envs ncu --metrics gpu__time_duration.sum --clock-control none -k regex:resblock_tc_kernel -s 18 -c 18 --csv --log-file gpurun_out/ll_${v}_$round.csv python tools/profile_forward.py > gpurun_out/ncu_$v.log 2>&1
  python - <<PY
import csv
rows=list(csv.reader(open('gpurun_out/ll_${v}_$round.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
hdr=rows[hi]; mv=hdr.index('Metric Value'); mu=hdr.index('Metric Unit')
v=[float(r[mv].replace(',',''))/(1000 if r[mu]=='ns' else 1) for r in rows[hi+1:] if len(r)>mv]
print("$v round $round: stage2 %.0f stage3 %.0f us |"%(sum(v[:9]),sum(v[9:18])), ' '.join('%.0f'%x for x in v))
PY
done
done
