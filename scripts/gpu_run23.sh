#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export LIFTER_VIT_WGRAD_STREAM=0
timeout 120 python scripts/vit_one_step.py || exit 1
STEPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'vit_attention_bwd_kernel|vit_ln_bwd_kernel|vit_colsum16_kernel|vit_attention_kernel' -c 12 -f -o gpurun_out/r02_vit_kernels_v1 python scripts/vit_one_step.py > gpurun_out/r02_vit_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r02_vit_ncu.log; ls -la gpurun_out/r02_vit_kernels_v1.ncu-rep
