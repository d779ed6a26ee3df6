#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 120 python scripts/gelu_gemm_once.py || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'lifter_gemm_kernel' -s 2 -c 1 -f -o gpurun_out/r02_vit_gelu_gemm_v2 python scripts/gelu_gemm_once.py > gpurun_out/r02_gelu_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r02_gelu_ncu.log; ls -la gpurun_out/r02_vit_gelu_gemm_v2.ncu-rep
