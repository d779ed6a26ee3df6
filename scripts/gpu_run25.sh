#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_25.log 2>&1
echo "rc=$?"; grep -h "passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_25.log | head -30
timeout 300 python scripts/vit_train_bench.py > gpurun_out/r02_vit_train_v5.json 2> gpurun_out/r02_vit_train_v5.err; echo "vit train bench rc=$?"; grep '^{' gpurun_out/r02_vit_train_v5.json | cut -c1-420; grep '^#' gpurun_out/r02_vit_train_v5.json | head -14; tail -3 gpurun_out/r02_vit_train_v5.err
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v3.json 2> gpurun_out/r02_vit_bench_v3.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v3.json | cut -c1-300; grep '^#' gpurun_out/r02_vit_bench_v3.json | head -8
timeout 400 python bench.py --workload vit --steps 20 --warmup 5 > gpurun_out/r02_bench_vit_v1.json 2> gpurun_out/r02_bench_vit_v1.err; echo "bench vit rc=$?"; cut -c1-1500 gpurun_out/r02_bench_vit_v1.json; tail -3 gpurun_out/r02_bench_vit_v1.err
timeout 400 python bench.py --workload vit_eval --steps 10 --warmup 3 > gpurun_out/r02_bench_vit_eval_v1.json 2> gpurun_out/r02_bench_vit_eval_v1.err; echo "bench vit_eval rc=$?"; cut -c1-1200 gpurun_out/r02_bench_vit_eval_v1.json; tail -3 gpurun_out/r02_bench_vit_eval_v1.err
