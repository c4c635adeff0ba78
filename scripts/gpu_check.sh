#!/bin/bash
# Runs each GPU test group in its own process under `timeout`, so that one hung kernel cannot take the
# rest of the call with it.  Logs go to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
status=0
for grp in "cfg_ddim or casts or unsupported" "groupnorm or layernorm" "linear" "attention" "ada_pool"; do
  name=$(echo "$grp" | tr ' ' '_')
  timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --tb=short -k "$grp" > "gpurun_out/k_${name}.log" 2>&1
  rc=$?
  echo "group [$grp] rc=$rc"; tail -n 25 "gpurun_out/k_${name}.log"
  [ $rc -ne 0 ] && status=1
done
exit $status
