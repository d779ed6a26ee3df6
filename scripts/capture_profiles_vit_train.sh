#!/bin/bash
# Runs on the GPU box (gpurun): ncu passes of B200_PROFILING.md on the MyViT training bench (the bench line itself is taken
# without a profiler by `python scripts/vit_train_bench.py`).   usage: scripts/capture_profiles_vit_train.sh <tag>
tag=${1:-v1}
out=gpurun_out
mkdir -p $out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 700 -c 260 --csv \
  --log-file $out/r01_vit_train_launches_$tag.csv python scripts/vit_train_bench.py 8192 > $out/ncu_vit_train_$tag.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:vit_attention_bwd_kernel -s 4 -c 1 \
  -o $out/prof_vit_attention_bwd_$tag -f python scripts/vit_train_bench.py 8192 > $out/ncu_vit_train_full_$tag.log 2>&1; echo "ncu att bwd rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:vit_ln_bwd_kernel -s 4 -c 1 \
  -o $out/prof_vit_ln_bwd_$tag -f python scripts/vit_train_bench.py 8192 > $out/ncu_vit_train_full_ln_$tag.log 2>&1; echo "ncu ln bwd rc=$?"
