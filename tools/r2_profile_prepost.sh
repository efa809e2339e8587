#!/bin/bash
# ncu --set full of conv_pre and conv_post (V1 and V2, 16 utterances x 862 frames): what bounds the two small kernels
mkdir -p gpurun_out
for v in 1 2; do
  python tools/profile_forward.py --version $v > gpurun_out/plain_prepost_v$v.log 2>&1 || { echo "plain v$v failed"; tail -3 gpurun_out/plain_prepost_v$v.log; continue; }
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:conv_tc_kernel<__half, \(int\)128, \(int\)16,' -s 1 -c 1 -f -o gpurun_out/prof_post_v$v \
      python tools/profile_forward.py --version $v > gpurun_out/ncu_post_v$v.log 2>&1
  echo "ncu post v$v rc=$?"
  ncu -i gpurun_out/prof_post_v$v.ncu-rep --page raw --csv > gpurun_out/prof_post_v${v}_raw.csv 2>/dev/null
  ncu -i gpurun_out/prof_post_v$v.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/prof_post_v${v}_source.csv.gz
  rm -f gpurun_out/prof_post_v$v.ncu-rep
done
du -sh gpurun_out
