#!/bin/bash
# ncu evidence of one eager DDIM step (cudaProfilerStart/Stop bracket, bench.py --profile-step): the plain run first
# (B200_PROFILING.md), then (1) the launch list with every kernel's device time, (2) `--set full` captures of the
# memory-bound kernels, the attention kernels and the GEMMs.  Usage: gpu_ncu.sh [tag]
mkdir -p gpurun_out
TAG=${1:-r02}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --profile-step"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_${TAG}.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu1_${TAG}.log 2>&1
echo "launch list rc=$?"
full() {   # name, kernel regex, count
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$2" -c $3 -o gpurun_out/prof_${TAG}_$1 $CMD > gpurun_out/ncu_${TAG}_$1.log 2>&1
  echo "full $1 rc=$?"
  ncu -i gpurun_out/prof_${TAG}_$1.ncu-rep --page raw --csv > gpurun_out/prof_raw_${TAG}_$1.csv 2>/dev/null
  rm -f gpurun_out/prof_${TAG}_$1.ncu-rep
}
full norms 'gn_partial_tok|gn_apply_tok' 4
full slab 'gn_slab_tok' 6
full ada 'ada_pool_main|ada_pool_combine|ada_head' 6
full small 'cfg_ddim|im2col3x3|conv3x3_few_out|zero_bf16|add_bias_residual' 8
full attn 'attention_fwd_kernel' 4
full gemm 'gemm_bf16_tma_kernel' 24
ls -la gpurun_out/ | tail -14
