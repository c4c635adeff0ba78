#!/bin/bash
# The allocation-free entry point's test alone in a fresh process, then the whole GPU suite (it is collected last).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q --tb=short > gpurun_out/t_noalloc_alone.log 2>&1
echo "alone rc=$?"; tail -n 4 gpurun_out/t_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q --tb=short > gpurun_out/t_all.log 2>&1
echo "all rc=$?"; tail -n 6 gpurun_out/t_all.log
