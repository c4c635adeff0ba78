#!/bin/bash
# attention kernel iteration: parity tests of the attention ops, then the stand-alone probes
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "attention" --tb=short > gpurun_out/t_attn.log 2>&1
echo "attention tests rc=$?"; tail -n 5 gpurun_out/t_attn.log
for b in "$@"; do echo "== $b"; timeout 120 tools/bin/$b 40 4096 4096 16; done
