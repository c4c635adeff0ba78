#!/bin/bash
# `--set full` rows of the small kernels that the main pass's launch-count window misses (cfg_ddim, conv_out, im2col).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-step"
$CMD > gpurun_out/plain_small.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_small.log; exit 1; }
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'cfg_ddim|conv3x3_few_out|im2col3x3' -c 4 -o gpurun_out/prof_small2 $CMD > gpurun_out/ncu_small2.log 2>&1
echo "rc=$?"
ncu -i gpurun_out/prof_small2.ncu-rep --page raw --csv > gpurun_out/prof_raw_small2.csv 2>/dev/null
rm -f gpurun_out/prof_small2.ncu-rep
