#!/bin/bash
# Multi-GPU measurements on one box (gpurun --gpus N): the sharded-sampling job (BASELINE.json configs[2], all-gather inside
# the timed region), the inversion step (configs[3], flat-gradient all-reduce inside the timed region), the reference arm
# under torchrun.  Usage: gpu_multi.sh [steps]
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
STEPS=${1:-20}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 bench.py --gpus $NG --steps $STEPS --warmup 5 --check-gather > gpurun_out/bench_${NG}gpu.json 2> gpurun_out/bench_${NG}gpu.err
echo "bench $NG GPUs rc=$?"; tail -n 3 gpurun_out/bench_${NG}gpu.err; head -c 1800 gpurun_out/bench_${NG}gpu.json; echo
timeout 900 $TR --master-port 29513 bench.py --gpus $NG --workload config4 --steps 10 --warmup 3 > gpurun_out/bench_inversion_${NG}gpu.json 2> gpurun_out/bench_inversion_${NG}gpu.err
echo "inversion $NG GPUs rc=$?"; tail -n 3 gpurun_out/bench_inversion_${NG}gpu.err; head -c 1500 gpurun_out/bench_inversion_${NG}gpu.json; echo
timeout 600 $TR --master-port 29512 bench.py --gpus $NG --steps 2 --warmup 1 --impl reference --no-full-batch > gpurun_out/bench_ref_${NG}gpu.json 2> gpurun_out/bench_ref_${NG}gpu.err
echo "ref $NG rc=$?"; head -c 600 gpurun_out/bench_ref_${NG}gpu.json; echo
