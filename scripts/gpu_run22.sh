#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py -m gpu -q -s -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_22.log 2>&1
echo "rc=$?"; grep -h "worst err/tol\|passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_22.log | head -30
timeout 300 python scripts/vit_train_bench.py > gpurun_out/r02_vit_train_v3.json 2> gpurun_out/r02_vit_train_v3.err; echo "vit train bench rc=$?"; grep '^{' gpurun_out/r02_vit_train_v3.json | cut -c1-700; grep '^#' gpurun_out/r02_vit_train_v3.json | head -8; tail -3 gpurun_out/r02_vit_train_v3.err
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v2.json 2> gpurun_out/r02_vit_bench_v2.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v2.json | cut -c1-300; grep '^#' gpurun_out/r02_vit_bench_v2.json | head -6
