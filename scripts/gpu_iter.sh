#!/bin/bash
# kernel iteration: the kernel-level GPU tests, then a short bench with the per-shape table; with a second argument
# "ENV=VALUE" the bench runs once more under that setting (A/B).  Usage: gpu_iter.sh [pytest -k expr] [ENV=VALUE]
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_modules_gpu.py -m gpu -x -q --tb=short ${1:+-k "$1"} > gpurun_out/t_iter.log 2>&1
echo "tests rc=$?"; tail -n 4 gpurun_out/t_iter.log
run() {   # tag, env assignment or empty
  env $2 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --dump-kernels gpurun_out/kernels_$1.csv > gpurun_out/bench_$1.json 2> gpurun_out/bench_$1.err
  echo "bench $1 ($2) rc=$?"; tail -n 3 gpurun_out/bench_$1.err
  python - "$1" <<'PY'
import json, sys
d = json.load(open(f"gpurun_out/bench_{sys.argv[1]}.json"))
print("steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), "ms", round(d["ms_per_step"], 3), "st_path ms", round(d["st_path"]["ms_per_step"], 3),
      "parity", d["parity"].get("fixture_rel_l2"), d["clocks"])
for k, v in sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms"]):
    print(f"  {k:22s} {v['ms']:.3f} ms  {v['launches']:3d} launches  {('%.0f TF/s' % v['tflops']) if v['tflops'] else ('%.0f GB/s' % v['gbs'])}")
PY
}
run iter ""
if [ -n "$2" ]; then run iter_ab "$2"; fi
