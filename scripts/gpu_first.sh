#!/bin/bash
# End-to-end GPU pass: kernel + module tests, smoke, bench.  Every stage under its own timeout; logs in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -k "groupnorm" > gpurun_out/t_gn.log 2>&1
echo "gn rc=$?"; tail -n 8 gpurun_out/t_gn.log
timeout 900 python -m pytest tests/test_modules_gpu.py -m gpu -q --tb=short -k "not slow and not unet and not full_loop" > gpurun_out/t_modules.log 2>&1
echo "modules rc=$?"; tail -n 12 gpurun_out/t_modules.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -n 5 gpurun_out/smoke.log
timeout 900 python -m pytest tests/test_modules_gpu.py -m gpu -q --tb=short -s -k "unet or full_loop" > gpurun_out/t_unet.log 2>&1
echo "unet rc=$?"; tail -n 30 gpurun_out/t_unet.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -n 5 gpurun_out/bench.err; cat gpurun_out/bench.json
