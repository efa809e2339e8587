#!/bin/bash
# ncu --set full of the tensor-core launches of one MPD + MSD forward (32 x 8192): second forward of each model
mkdir -p gpurun_out
python tools/disc_bench.py --once > gpurun_out/plain_disc.log 2>&1 || { echo "plain run failed"; tail -3 gpurun_out/plain_disc.log; exit 1; }
# per model: warm-up forward then the measured one; MPD has 40 conv_tf32 launches per forward, MSD 36
ncu --set full --clock-control none -k regex:conv_tf32_kernel -s 40 -c 40 -f -o gpurun_out/prof_disc_mpd python tools/disc_bench.py --once > gpurun_out/ncu_disc_mpd.log 2>&1
echo "ncu mpd rc=$?"
ncu -i gpurun_out/prof_disc_mpd.ncu-rep --page raw --csv > gpurun_out/prof_disc_mpd_raw.csv 2>/dev/null; rm -f gpurun_out/prof_disc_mpd.ncu-rep
ncu --set full --clock-control none -k regex:conv_tf32_kernel -s 116 -c 36 -f -o gpurun_out/prof_disc_msd python tools/disc_bench.py --once > gpurun_out/ncu_disc_msd.log 2>&1
echo "ncu msd rc=$?"
ncu -i gpurun_out/prof_disc_msd.ncu-rep --page raw --csv > gpurun_out/prof_disc_msd_raw.csv 2>/dev/null; rm -f gpurun_out/prof_disc_msd.ncu-rep
du -sh gpurun_out
