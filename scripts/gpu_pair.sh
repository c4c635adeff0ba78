#!/bin/bash
# CTA-pair GEMM bring-up: GEMM tests under a short hard timeout first (a hang must not hold the box), then the bench
# with and without pair mode.
mkdir -p gpurun_out
timeout -s KILL 240 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -x -k "linear or gemm" > gpurun_out/t_pair.log 2>&1
rc=$?
echo "gemm tests rc=$rc"; tail -n 25 gpurun_out/t_pair.log
if [ $rc -ne 0 ]; then exit $rc; fi
for mode in 0 1; do
  AFB_GEMM_PAIR=$mode timeout -s KILL 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --dump-kernels gpurun_out/kernels_pair$mode.csv > gpurun_out/bench_pair$mode.json 2> gpurun_out/bench_pair$mode.err
  echo "bench pair=$mode rc=$?"; tail -n 3 gpurun_out/bench_pair$mode.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_pair$mode.json"))
print("pair=$mode ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), d["clocks"])
for k, v in d["kernels"].items():
    print(f"  {k:18s} {v['ms']:8.3f} ms  {v['launches']:4d}  tflops={v['tflops']}")
PY
  grep "^gemm" gpurun_out/kernels_pair$mode.csv | head -30
done
