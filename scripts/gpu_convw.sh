#!/bin/bash
# Which image sizes should take the native 3x3 convolution?  Bench cuDNN everywhere vs native for sets of widths.
mkdir -p gpurun_out
run() {
  env "$@" timeout -s KILL 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_w.json 2> gpurun_out/bench_w.err
  python - "$*" <<'PY'
import json, sys
d = json.load(open("gpurun_out/bench_w.json"))
print(sys.argv[1], "ms/step", round(d["ms_per_step"], 3), "steps/s", round(d["value"], 2), d["clocks"]["sm_mhz"])
PY
}
for rep in 1 2; do
run AFB_CONV=0
run AFB_CONV=1 AFB_CONV_W=16
run AFB_CONV=1 AFB_CONV_W=16,32
run AFB_CONV=1 AFB_CONV_W=16,32,64
run AFB_CONV=1 AFB_CONV_W=16,8
done
