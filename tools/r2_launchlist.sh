#!/bin/bash
mkdir -p gpurun_out
python tools/profile_forward.py ${PF_ARGS:-} > gpurun_out/plain_profile.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed.avg.per_cycle_active --clock-control none -s ${SKIP:-61} -c ${COUNT:-61} --csv --log-file gpurun_out/launches_fwd_b16.csv python tools/profile_forward.py ${PF_ARGS:-} > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/launches_fwd_b16.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
hdr=rows[hi]; kn=hdr.index('Kernel Name'); mn=hdr.index('Metric Name'); mv=hdr.index('Metric Value'); mu=hdr.index('Metric Unit')
d={}
for r in rows[hi+1:]:
    if len(r)<=mv: continue
    v=float(r[mv].replace(',','')); 
    if r[mn].startswith('gpu__time') and r[mu]=='ns': v/=1000
    d.setdefault(int(r[0]),{'name':r[kn]})[r[mn].split('.')[0]]=v
for i,e in sorted(d.items()):
    n=e['name']; inst=n[n.find('<')+1:n.find('>(')].replace('(int)','') if '<' in n else ''
    tag='RB' if 'resblock' in n else ('TC' if 'conv_tc' in n else n[:12])
    print(i,tag,inst,' '.join(f"{k[-22:]}={v:.1f}" for k,v in e.items() if k!='name'))
PY
