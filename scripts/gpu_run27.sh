#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_27.log 2>&1
echo "rc=$?"; grep -h "passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_27.log | head -30
timeout 300 python scripts/vit_train_bench.py > gpurun_out/r02_vit_train_v7.json 2> gpurun_out/r02_vit_train_v7.err; echo "vit train bench rc=$?"; grep '^{' gpurun_out/r02_vit_train_v7.json | cut -c1-520; grep '^#' gpurun_out/r02_vit_train_v7.json | head -20; tail -3 gpurun_out/r02_vit_train_v7.err
