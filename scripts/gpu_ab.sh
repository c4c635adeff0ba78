#!/bin/bash
# A/B of an environment switch: tests with the switch on, then the bench with VAR=0 and VAR=1.  Usage: gpu_ab.sh VAR ["-k expr"]
VAR=${1:-AFB_PDL}
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -m gpu -q --tb=short -x ${2:+-k "$2"} > gpurun_out/t_ab.log 2>&1
rc=$?
echo "tests rc=$rc"; tail -n 12 gpurun_out/t_ab.log
if [ $rc -ne 0 ]; then exit $rc; fi
for mode in 0 1 0 1; do
  env $VAR=$mode timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ab$mode.json 2> gpurun_out/bench_ab$mode.err
  echo "bench $VAR=$mode rc=$?"; tail -n 2 gpurun_out/bench_ab$mode.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_ab$mode.json"))
print("$VAR=$mode ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), d["clocks"])
PY
done
