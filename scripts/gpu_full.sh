#!/bin/bash
# Round-end style pass: the whole `-m gpu` suite, smoke(), default bench (with the CPU baseline) and the reference arm.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --tb=short > gpurun_out/t_all.log 2>&1
echo "tests rc=$?"; tail -n 6 gpurun_out/t_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py --dump-kernels gpurun_out/kernels_full.csv > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
echo "bench rc=$?"; tail -n 3 gpurun_out/bench_full.err; cat gpurun_out/bench_full.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "ref rc=$?"; cat gpurun_out/bench_ref.json
