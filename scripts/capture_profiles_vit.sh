#!/bin/bash
# Runs on the GPU box (gpurun): MyViT bench line (no profiler), then the ncu passes of B200_PROFILING.md on the same script.
# usage: scripts/capture_profiles_vit.sh <tag>      (outputs under gpurun_out/)
tag=${1:-v1}
out=gpurun_out
mkdir -p $out
timeout 300 python scripts/vit_bench.py > $out/r01_vit_bench_$tag.json 2> $out/vit_bench_$tag.err; echo "vit bench rc=$?"
cat $out/r01_vit_bench_$tag.json | cut -c1-900
# launch list (cold-cache, serialised: shares only) of a 8,192-pose run
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 300 --csv \
  --log-file $out/r01_vit_launches_$tag.csv python scripts/vit_bench.py 8192 > $out/ncu_vit_$tag.log 2>&1; echo "ncu list rc=$?"
# one full capture of the GELU GEMM (third GEMM launch of a pass: qkv, attention output, MLP up) and one of the attention kernel
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lifter_gemm_kernel -s 2 -c 1 \
  -o $out/prof_vit_gelu_gemm_$tag -f python scripts/vit_bench.py 8192 > $out/ncu_vit_full_$tag.log 2>&1; echo "ncu gemm rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:vit_attention_kernel -s 6 -c 1 \
  -o $out/prof_vit_attention_$tag -f python scripts/vit_bench.py 8192 > $out/ncu_vit_full_att_$tag.log 2>&1; echo "ncu att rc=$?"
ls -la $out | tail -12
