#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py tests/test_gemm_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_35.log 2>&1
echo "rc=$?"; grep -h "passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_35.log | head -30
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v7.json 2> gpurun_out/r02_vit_bench_v7.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v7.json | cut -c1-520; grep '^#' gpurun_out/r02_vit_bench_v7.json | head -6; grep -o '"roofline": {[^}]*}' gpurun_out/r02_vit_bench_v7.json | cut -c1-300
