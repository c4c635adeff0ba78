#!/bin/bash
# Round-end style pass: the whole `-m gpu` suite, smoke(), default bench (with the oracle check / CPU baseline) and the
# reference arm.  With N > 1 GPUs (gpurun --gpus N) also the sharded-job bench at N ranks.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --tb=short -s > gpurun_out/t_all.log 2>&1
echo "tests rc=$?"; grep -E "rel-L2|cosine|passed|failed|error" gpurun_out/t_all.log | tail -n 25
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py --write-step0 --dump-kernels gpurun_out/kernels_full.csv > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
echo "bench rc=$?"; tail -n 3 gpurun_out/bench_full.err; head -c 1500 gpurun_out/bench_full.json; echo
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -gt 1 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $NG --steps 20 --warmup 5 --check-gather > gpurun_out/bench_${NG}gpu.json 2> gpurun_out/bench_${NG}gpu.err
  echo "bench $NG GPUs rc=$?"; tail -n 5 gpurun_out/bench_${NG}gpu.err; head -c 2500 gpurun_out/bench_${NG}gpu.json; echo
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NG --steps 20 --warmup 5 --impl reference --no-full-batch > gpurun_out/bench_ref_${NG}gpu.json 2> gpurun_out/bench_ref_${NG}gpu.err
  echo "ref $NG rc=$?"; head -c 600 gpurun_out/bench_ref_${NG}gpu.json; echo
else
  timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
  echo "ref rc=$?"; head -c 800 gpurun_out/bench_ref.json; echo
fi
