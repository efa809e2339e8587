#!/bin/bash
# ncu launch lists (gpu__time_duration, tensor pipe %, IPC) of one forward of V1 / V2 / V3 at 16 utterances x 862 frames
mkdir -p gpurun_out
for v in 1 2 3; do
  python tools/profile_forward.py --version $v > gpurun_out/plain_v$v.log 2>&1 || { echo "plain v$v failed"; continue; }
  n=$(awk '{print $NF}' gpurun_out/plain_v$v.log | tail -1)
  ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed.avg.per_cycle_active \
      --clock-control none -s $n -c $n --csv --log-file gpurun_out/launches_v${v}_fwd_b16.csv python tools/profile_forward.py --version $v > gpurun_out/ncu_v$v.log 2>&1
  echo "v$v launches=$n rc=$?"
done
