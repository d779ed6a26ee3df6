#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_vit_gpu.py tests/test_gemm_gpu.py tests/test_ingest_eval_gpu.py "tests/test_bench_shapes_gpu.py::test_graph_replay_revalidates_operands_and_survives_batch_size_changes" -m gpu -q -s -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_20.log 2>&1
echo "rc=$?"; grep -h "worst err/tol\|passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_20.log | head -30
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v1.json 2> gpurun_out/r02_vit_bench_v1.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v1.json | cut -c1-400
