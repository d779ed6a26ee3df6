#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py tests/test_gemm_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_26.log 2>&1
echo "rc=$?"; grep -h "passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_26.log | head -30
timeout 300 python scripts/vit_train_bench.py > gpurun_out/r02_vit_train_v6.json 2> gpurun_out/r02_vit_train_v6.err; echo "vit train bench rc=$?"; grep '^{' gpurun_out/r02_vit_train_v6.json | cut -c1-520; grep '^#' gpurun_out/r02_vit_train_v6.json | head -8; tail -3 gpurun_out/r02_vit_train_v6.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_26.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r02_smoke_26.log
