#!/bin/bash
# GEMM phase probe (tools/gemm_probe.cu built here with and without -DAFB_GEMM_PROFILE): where the MMA issuer and the
# epilogue warps of the first CTAs spend their cycles, per shape, wide (2 x 160-column) tiles on and off.
mkdir -p gpurun_out
{
for shp in "16384 640 640 0 0 1 1" "4096 1280 1280 0 0 1 1" "65536 320 320 0 0 1 1" "16384 640 640 0 1 0 0" "16384 640 2560 0 0 1 0"; do
  for wide in 1 0; do
    echo "== $shp wide=$wide"
    AFB_GEMM_WIDE=$wide tools/bin/gemm_probe $shp
    AFB_GEMM_WIDE=$wide tools/bin/gemm_probe_prof $shp | head -4
  done
done
} > gpurun_out/gemm_probe.log 2>&1
cat gpurun_out/gemm_probe.log
