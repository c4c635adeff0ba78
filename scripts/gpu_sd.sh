#!/bin/bash
# Double-buffered S attention: parity tests, then the bench with the per-shape kernel table.
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -x -k "attention" > gpurun_out/t_sd.log 2>&1
rc=$?
echo "attention tests rc=$rc"; tail -n 25 gpurun_out/t_sd.log
if [ $rc -ne 0 ]; then exit $rc; fi
timeout -s KILL 600 python -m pytest tests/test_modules_gpu.py -m gpu -q --tb=short -x -k "unet_step or spatial_transformer" > gpurun_out/t_sd2.log 2>&1
rc=$?
echo "module tests rc=$rc"; tail -n 8 gpurun_out/t_sd2.log
if [ $rc -ne 0 ]; then exit $rc; fi
for rep in 1 2; do
  timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --dump-kernels gpurun_out/kernels_sd.csv > gpurun_out/bench_sd.json 2> gpurun_out/bench_sd.err
  echo "bench rc=$?"; tail -n 3 gpurun_out/bench_sd.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_sd.json"))
print("ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), "st", round(d["st_path"]["ms_per_step"], 3), d["roofline"]["frac"], d["clocks"])
PY
  grep "attention_fwd" gpurun_out/kernels_sd.csv
done
