#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch list, ncu --set full of selected conv_tc launches.
# ncu exports are converted to CSV on the box; gpurun_out/ must stay below 64 MiB.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
fi
if [ "${SKIP_BENCH:-0}" != "1" ]; then
python bench.py ${BENCH_ARGS:-} > gpurun_out/bench_v1.json 2> gpurun_out/bench_v1.err; echo "bench rc=$?"; cat gpurun_out/bench_v1.json
fi
if [ "${SKIP_REF:-1}" != "1" ]; then
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
fi
if [ "${SKIP_NCU:-0}" != "1" ]; then
[ "${SKIP_LAUNCHLIST:-0}" != "1" ] && python tools/profile_forward.py > gpurun_out/plain_profile.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 79 -c 79 --csv --log-file gpurun_out/launches_fwd_b16.csv python tools/profile_forward.py > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
# conv_tc launch index within one forward: 0 conv_pre; stage i base = 1+19i: +0 ups[i], +1..6 k=3, +7..12 k=7, +13..18 k=11
# (c1 m0, c2 m0, c1 m1, c2 m1, c1 m2, c2 m2).  77 launches of the warm-up forward are skipped first.
DEFAULT_SPECS="s0k11:93:2 s1k11:112:2 s2k11:131:2 s3k11:150:2 ups:97:1"
for spec in ${NCU_SPECS:-$DEFAULT_SPECS}; do
  name=${spec%%:*}; rest=${spec#*:}; skip=${rest%%:*}; cnt=${rest#*:}
  python tools/profile_forward.py > gpurun_out/plain_$name.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s $skip -c $cnt -f -o gpurun_out/prof_$name python tools/profile_forward.py > gpurun_out/ncu_$name.log 2>&1
  echo "ncu $name rc=$?"
  ncu -i gpurun_out/prof_$name.ncu-rep --page raw --csv > gpurun_out/prof_${name}_raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_$name.ncu-rep --page source --csv > gpurun_out/prof_${name}_source.csv 2>/dev/null
done
fi
du -sh gpurun_out; ls -la gpurun_out
# stay under the 64 MiB copy-back limit: drop the biggest reports first
while [ $(du -sm gpurun_out | cut -f1) -ge 60 ]; do f=$(ls -S gpurun_out/*.ncu-rep 2>/dev/null | head -1); [ -z "$f" ] && break; rm -f "$f"; done
