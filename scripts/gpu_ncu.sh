#!/bin/bash
# ncu evidence: (1) launch list with the device time of every kernel of ONE eager DDIM step (cudaProfilerStart/Stop
# bracket), (2) optionally --set full of selected kernels of that step.  The same command runs first without ncu
# (B200_PROFILING.md).  Usage: gpu_ncu.sh [tag] [kernel-regex for the full capture] [count]
mkdir -p gpurun_out
TAG=${1:-r01}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-step"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_${TAG}.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu1_${TAG}.log 2>&1
echo "launch list rc=$?"
if [ -n "$2" ]; then
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$2" ${4:+-s $4} -c ${3:-30} -o gpurun_out/prof_${TAG} $CMD > gpurun_out/ncu2_${TAG}.log 2>&1
  echo "full rc=$?"; tail -2 gpurun_out/ncu2_${TAG}.log
  # the raw page as CSV always travels; the report itself only when it fits the 64 MiB return limit
  ncu -i gpurun_out/prof_${TAG}.ncu-rep --page raw --csv > gpurun_out/prof_raw_${TAG}.csv 2>/dev/null
  if [ $(stat -c %s gpurun_out/prof_${TAG}.ncu-rep) -gt 40000000 ]; then rm -f gpurun_out/prof_${TAG}.ncu-rep; echo "report dropped (too large), raw CSV kept"; fi
fi
ls -la gpurun_out/ | tail -8
