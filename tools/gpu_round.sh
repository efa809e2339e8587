#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch list, ncu --set full of the conv_tc kernels.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_v1.json 2> gpurun_out/bench_v1.err; echo "bench rc=$?"; cat gpurun_out/bench_v1.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
python tools/profile_forward.py > gpurun_out/plain_profile.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 79 -c 79 --csv --log-file gpurun_out/launches_fwd_b16.csv python tools/profile_forward.py > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
python tools/profile_forward.py > gpurun_out/plain_profile2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s ${NCU_SKIP:-110} -c ${NCU_COUNT:-41} -f -o gpurun_out/prof_conv_tc python tools/profile_forward.py > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_full.log
