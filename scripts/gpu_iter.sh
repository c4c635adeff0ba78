#!/bin/bash
# kernel iteration: the kernel-level GPU tests, then a short bench with the per-shape table.  Usage: gpu_iter.sh [pytest -k expr]
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_modules_gpu.py -m gpu -x -q --tb=short ${1:+-k "$1"} > gpurun_out/t_iter.log 2>&1
echo "tests rc=$?"; tail -n 4 gpurun_out/t_iter.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --dump-kernels gpurun_out/kernels_iter.csv > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err
echo "bench rc=$?"; tail -n 3 gpurun_out/bench_iter.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_iter.json"))
print("steps/s", round(d["value"], 2), "e2e", round(d["e2e"]["value"], 2), "ms", round(d["ms_per_step"], 3), "st_path ms", round(d["st_path"]["ms_per_step"], 3),
      "parity", d["parity"].get("fixture_rel_l2"), d["clocks"])
for k, v in sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms"]):
    print(f"  {k:22s} {v['ms']:.3f} ms  {v['launches']:3d} launches  {('%.0f TF/s' % v['tflops']) if v['tflops'] else ('%.0f GB/s' % v['gbs'])}")
PY
AFB_GEMM_AREUSE=0 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --dump-kernels gpurun_out/kernels_iter_off.csv > gpurun_out/bench_iter_off.json 2> gpurun_out/bench_iter_off.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_iter_off.json"))
print("AREUSE=0: steps/s", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "st_path ms", round(d["st_path"]["ms_per_step"], 3), "gemm ms", d["kernels"]["gemm_bf16"]["ms"])
PY
