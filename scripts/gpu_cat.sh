#!/bin/bash
# Virtual-concat bring-up: GroupNorm tests, U-Net parity, then A/B of AFB_CAT.
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -x -k "groupnorm" > gpurun_out/t_cat.log 2>&1
rc=$?
echo "gn tests rc=$rc"; tail -n 25 gpurun_out/t_cat.log
if [ $rc -ne 0 ]; then exit $rc; fi
timeout -s KILL 600 python -m pytest tests/test_modules_gpu.py -m gpu -q --tb=short -x -k "unet_step or ddim_full" > gpurun_out/t_cat2.log 2>&1
rc=$?
echo "unet tests rc=$rc"; tail -n 12 gpurun_out/t_cat2.log
if [ $rc -ne 0 ]; then exit $rc; fi
for mode in 0 1 0 1; do
  env AFB_CAT=$mode timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cat$mode.json 2> gpurun_out/bench_cat$mode.err
  echo "bench AFB_CAT=$mode rc=$?"; tail -n 3 gpurun_out/bench_cat$mode.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_cat$mode.json"))
print("AFB_CAT=$mode ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), d["clocks"])
PY
done
