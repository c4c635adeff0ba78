#!/bin/bash
# Default bench (with the oracle check + CPU baseline) and the reference arm.  Usage: gpu_bench.sh [extra bench args]
mkdir -p gpurun_out
timeout 1200 python bench.py --dump-kernels gpurun_out/kernels.csv "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -n 5 gpurun_out/bench.err; cat gpurun_out/bench.json | head -c 3000; echo
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "ref rc=$?"; tail -n 3 gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
