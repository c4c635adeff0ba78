#!/bin/bash
# Iteration pass: selected kernel tests, then the bench with a per-shape kernel breakdown.  Usage: gpu_iter.sh "<pytest -k expr>"
mkdir -p gpurun_out
if [ -n "$1" ]; then
  timeout 900 python -m pytest tests -m gpu -q --tb=short -k "$1" > gpurun_out/t_iter.log 2>&1
  echo "tests rc=$?"; tail -n 15 gpurun_out/t_iter.log
fi
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --dump-kernels gpurun_out/kernels.csv > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -n 5 gpurun_out/bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
print("ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), "st_ms", round(d["st_path"]["ms_per_step"], 3), "non_st_ms", round(d["non_st_ms_per_step"], 3), d["clocks"])
for k, v in d["kernels"].items():
    print(f"  {k:18s} {v['ms']:8.3f} ms  {v['launches']:4d}  tflops={v['tflops']}  gbs={v['gbs']}")
PY
head -40 gpurun_out/kernels.csv
