#!/bin/bash
# Short-key-sequence attention configuration: parity tests, then A/B of AFB_ATTN_SMALL with the per-shape kernel table.
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -x -k "attention" > gpurun_out/t_small.log 2>&1
rc=$?
echo "attention tests rc=$rc"; tail -n 25 gpurun_out/t_small.log
if [ $rc -ne 0 ]; then exit $rc; fi
timeout -s KILL 600 python -m pytest tests/test_modules_gpu.py -m gpu -q --tb=short -x -k "unet_step or spatial_transformer" > gpurun_out/t_small2.log 2>&1
rc=$?
echo "module tests rc=$rc"; tail -n 8 gpurun_out/t_small2.log
if [ $rc -ne 0 ]; then exit $rc; fi
for mode in 0 1 0 1; do
  env AFB_ATTN_SMALL=$mode timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --dump-kernels gpurun_out/kernels_small$mode.csv > gpurun_out/bench_small$mode.json 2> gpurun_out/bench_small$mode.err
  echo "bench AFB_ATTN_SMALL=$mode rc=$?"; tail -n 3 gpurun_out/bench_small$mode.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_small$mode.json"))
print("AFB_ATTN_SMALL=$mode ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), "st", round(d["st_path"]["ms_per_step"], 3), d["clocks"])
PY
  grep "attention_fwd" gpurun_out/kernels_small$mode.csv
done
