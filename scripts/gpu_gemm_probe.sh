#!/bin/bash
# GEMM phase probe (tools/gemm_probe.cu built here with and without -DAFB_GEMM_PROFILE): where the MMA issuer and the
# epilogue warps of the first CTAs spend their cycles, per shape; narrow tiles, wide tiles (8 epilogue warps) and, when
# tools/bin/gemm_probe16 exists (-DAFB_GEMM_WIDE_EPI_WARPS=16), wide tiles with 16 epilogue warps.
mkdir -p gpurun_out
{
for shp in "16384 640 640 0 0 1 0" "16384 640 640 0 1 0 0" "4096 1280 1280 0 0 1 0" "16384 1920 640 0 1 0 0" "16384 640 2560 0 0 1 0" "65536 320 320 0 0 1 0" "65536 960 320 0 1 0 0"; do
  echo "== $shp narrow"
  AFB_GEMM_WIDE=0 tools/bin/gemm_probe $shp
  AFB_GEMM_WIDE=0 tools/bin/gemm_probe_prof $shp | sed -n 2,3p
  echo "== $shp wide, 8 epilogue warps"
  AFB_GEMM_WIDE_MIN_K=320 tools/bin/gemm_probe $shp
  AFB_GEMM_WIDE_MIN_K=320 tools/bin/gemm_probe_prof $shp | sed -n 2,3p
  if [ -x tools/bin/gemm_probe16 ]; then
    echo "== $shp wide, 16 epilogue warps"
    AFB_GEMM_WIDE_MIN_K=320 tools/bin/gemm_probe16 $shp
    AFB_GEMM_WIDE_MIN_K=320 tools/bin/gemm_probe_prof16 $shp | sed -n 2,3p
  fi
done
} > gpurun_out/gemm_probe.log 2>&1
cat gpurun_out/gemm_probe.log
