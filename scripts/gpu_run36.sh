#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py tests/test_ingest_eval_gpu.py -m gpu -q -s -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_36.log 2>&1
echo "rc=$?"; grep -h "worst err/tol\|passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_36.log | head -30
timeout 300 python scripts/vit_train_bench.py > gpurun_out/r02_vit_train_v8.json 2> gpurun_out/r02_vit_train_v8.err; echo "vit train bench rc=$?"; grep '^{' gpurun_out/r02_vit_train_v8.json | cut -c1-420; grep '^#' gpurun_out/r02_vit_train_v8.json | grep -i "layernorm\|embed_bwd\|ln_bwd\|attention"
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v8.json 2> gpurun_out/r02_vit_bench_v8.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v8.json | cut -c1-520; grep '^#' gpurun_out/r02_vit_bench_v8.json | grep -i "layernorm\|embed"
