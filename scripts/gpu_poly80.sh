#!/bin/bash
# d=80 attention: share of exponentials on the FMA pipe (build-time switch), measured with the attention probe-free
# route: rebuild the library with the flag on the box and read the per-shape kernel table of the bench.
mkdir -p gpurun_out
for v in 0 4 3; do
  python - <<PY
import adaface_b200.build as B
B.NVCC_FLAGS.append("-DAFB_ATTN_POLY_EVERY_80=$v")
print(B.build(force=True))
PY
  timeout -s KILL 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "attention" 2>&1 | tail -1
  timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --dump-kernels gpurun_out/kernels_p80_$v.csv > gpurun_out/bench_p80_$v.json 2> gpurun_out/bench_p80_$v.err
  echo "POLY80=$v"; grep "attention_fwd,b16q1024k1024d80\|attention_fwd,b16q1024k77d80" gpurun_out/kernels_p80_$v.csv
done
