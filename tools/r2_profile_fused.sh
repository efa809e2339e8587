#!/bin/bash
# tests, then launch list + ncu --set full of the 18 fused ResBlock launches of one V1 forward (16 utterances)
# gpurun_out/ must stay below 64 MiB: reports are converted to CSV on the box and removed, source pages gzipped
mkdir -p gpurun_out
if [ "${SKIP_TESTS:-0}" != "1" ]; then
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
fi
python tools/profile_forward.py > gpurun_out/plain_profile.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 61 -c 61 --csv --log-file gpurun_out/launches_fwd_b16.csv python tools/profile_forward.py > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
python tools/profile_forward.py > gpurun_out/plain_profile2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:resblock_tc_kernel -s 18 -c 18 -f -o gpurun_out/prof_fused python tools/profile_forward.py > gpurun_out/ncu_fused.log 2>&1
echo "ncu fused rc=$?"
ncu -i gpurun_out/prof_fused.ncu-rep --page raw --csv > gpurun_out/prof_fused_raw.csv 2>/dev/null
rm -f gpurun_out/prof_fused.ncu-rep
for spec in s2k7:22 s2k11:25 s3k3:27; do
  name=${spec%%:*}; skip=${spec#*:}
  ncu --set full --clock-control none --import-source on -k regex:resblock_tc_kernel -s $skip -c 1 -f -o gpurun_out/prof_$name python tools/profile_forward.py > gpurun_out/ncu_$name.log 2>&1
  ncu -i gpurun_out/prof_$name.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/prof_${name}_source.csv.gz
  rm -f gpurun_out/prof_$name.ncu-rep
done
du -sh gpurun_out; ls -la gpurun_out
