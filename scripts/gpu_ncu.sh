#!/bin/bash
# ncu evidence: (1) launch list with device time of every kernel of two eager steps, (2) --set full of the top kernels.
# The same command runs first without ncu (B200_PROFILING.md).  Outputs in gpurun_out/.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 3500 -c 1300 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"; tail -2 gpurun_out/ncu1.log
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"attention_fwd_kernel|gemm_bf16_tma_kernel|gn_partial_tok|gn_apply_tok|ada_pool_main|layernorm_bf16" -s 600 -c 40 -o gpurun_out/prof_r01 $CMD > gpurun_out/ncu2.log 2>&1
echo "full rc=$?"; tail -2 gpurun_out/ncu2.log; ls -la gpurun_out/
