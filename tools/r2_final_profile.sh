#!/bin/bash
# final-build evidence: tests, bench lines (r2_full.sh), launch list of one forward and ncu --set full of the 18 fused launches
if [ "${SKIP_FULL:-0}" != "1" ]; then bash tools/r2_full.sh; fi
python tools/profile_forward.py > gpurun_out/plain_profile.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 61 -c 61 --csv --log-file gpurun_out/launches_fwd_b16.csv python tools/profile_forward.py > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
python tools/profile_forward.py > gpurun_out/plain_profile2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:resblock_tc_kernel -s 18 -c 18 -f -o gpurun_out/prof_fused python tools/profile_forward.py > gpurun_out/ncu_fused.log 2>&1
echo "ncu fused rc=$?"
ncu -i gpurun_out/prof_fused.ncu-rep --page raw --csv > gpurun_out/prof_fused_raw.csv 2>/dev/null
rm -f gpurun_out/prof_fused.ncu-rep
python tools/profile_forward.py > gpurun_out/plain_profile3.log 2>&1 && \
ncu --set full --clock-control none -k regex:conv_tc_kernel -s 60 -c 18 -f -o gpurun_out/prof_s1 python tools/profile_forward.py > gpurun_out/ncu_s1.log 2>&1
echo "ncu stage1 rc=$?"
ncu -i gpurun_out/prof_s1.ncu-rep --page raw --csv > gpurun_out/prof_stage1_raw.csv 2>/dev/null
rm -f gpurun_out/prof_s1.ncu-rep
du -sh gpurun_out
