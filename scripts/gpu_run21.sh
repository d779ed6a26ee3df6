#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_vit_gpu.py -m gpu -q -s -p no:cacheprovider --timeout=200 --timeout-method=thread -k "fused_trainer or training_loop" > gpurun_out/r02_tests_21.log 2>&1
echo "rc=$?"; grep -h "passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_21.log | head -30
timeout 300 python scripts/vit_train_bench.py > gpurun_out/r02_vit_train_v2.json 2> gpurun_out/r02_vit_train_v2.err; echo "vit train bench rc=$?"; grep '^{' gpurun_out/r02_vit_train_v2.json | cut -c1-900; tail -3 gpurun_out/r02_vit_train_v2.err
