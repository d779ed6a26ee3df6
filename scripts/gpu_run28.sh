#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v4.json 2> gpurun_out/r02_vit_bench_v4.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v4.json | cut -c1-700
# ncu launch list of the bench command (profiles contract) and --set full pages of the rewritten kernels
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_vit_train_launches_v1.csv python bench.py --workload vit --steps 2 --warmup 3 > gpurun_out/r02_vit_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
export LIFTER_VIT_WGRAD_STREAM=0
STEPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'vit_attention_bwd_kernel|vit_ln_bwd_kernel|vit_attention_kernel' -c 8 -f -o gpurun_out/r02_vit_kernels_v2 python scripts/vit_one_step.py > gpurun_out/r02_vit_ncu2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r02_vit_ncu2.log; ls -la gpurun_out/r02_vit_kernels_v2.ncu-rep
