#!/bin/bash
# Implicit-GEMM conv bring-up: conv + GEMM kernel tests under a short hard timeout, U-Net parity, then A/B benches.
# Usage: gpu_conv.sh [VAR] (default AFB_UPSAMPLE_FOLD; AFB_CONV for the plain 3x3 convolutions)
VAR=${1:-AFB_UPSAMPLE_FOLD}
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -x -k "conv3x3 or linear" > gpurun_out/t_conv.log 2>&1
rc=$?
echo "conv tests rc=$rc"; tail -n 25 gpurun_out/t_conv.log
if [ $rc -ne 0 ]; then exit $rc; fi
timeout -s KILL 600 python -m pytest tests/test_modules_gpu.py -m gpu -q --tb=short -x -k "unet_step or ddim_full" > gpurun_out/t_conv2.log 2>&1
echo "unet tests rc=$?"; tail -n 8 gpurun_out/t_conv2.log
for mode in 0 1 0 1; do
  env $VAR=$mode timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --dump-kernels gpurun_out/kernels_conv$mode.csv > gpurun_out/bench_conv$mode.json 2> gpurun_out/bench_conv$mode.err
  echo "bench $VAR=$mode rc=$?"; tail -n 3 gpurun_out/bench_conv$mode.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_conv$mode.json"))
print("$VAR=$mode ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), d["clocks"])
PY
done
grep "^conv" gpurun_out/kernels_conv1.csv | head -30
