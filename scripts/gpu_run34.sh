#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_vit_gpu.py tests/test_gemm_gpu.py tests/test_eval_gpu.py tests/test_bench_shapes_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_34.log 2>&1
echo "rc=$?"; grep -h "passed\|failed\|^FAILED\|^E  " gpurun_out/r02_tests_34.log | head -30
timeout 300 python scripts/vit_bench.py > gpurun_out/r02_vit_bench_v6.json 2> gpurun_out/r02_vit_bench_v6.err; echo "vit bench rc=$?"; grep '^{' gpurun_out/r02_vit_bench_v6.json | cut -c1-520; grep '^#' gpurun_out/r02_vit_bench_v6.json | head -6
timeout 300 python bench.py --workload eval --steps 20 --warmup 5 > gpurun_out/r02_bench_eval_v2.json 2> gpurun_out/r02_bench_eval_v2.err; echo "eval rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_eval_v2.json"))
print("eval ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "roof", round(d["roofline"]["frac"], 3), d["roofline"]["us_per_launch"])
PY
timeout 300 python bench.py --workload vit --steps 20 --warmup 5 > gpurun_out/r02_bench_vit_v2.json 2> gpurun_out/r02_bench_vit_v2.err; echo "vit rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_vit_v2.json"))
print("vit ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "roof", round(d["roofline"]["frac"], 3), d["roofline"]["us_per_launch"])
PY
