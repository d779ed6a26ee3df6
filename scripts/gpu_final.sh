#!/usr/bin/env bash
# round-end check as the driver runs it: smoke(), pytest -m gpu -x, default bench, reference arm
set -u
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('__SMOKE_OK__')" 2>&1 | tail -6
timeout 1200 python -m pytest tests -x -q -m gpu -p no:cacheprovider --timeout=300 --timeout-method=thread > gpurun_out/r02_tests_final.log 2>&1
echo "pytest -x rc=$?"; grep -h "passed\|failed\|^FAILED\|^ERROR\|Timeout" gpurun_out/r02_tests_final.log | head
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_train_final.json 2> gpurun_out/r02_bench_train_final.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_bench_train_final.json"))
print("train ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "sustained", round(d["sustained"]["ms_per_step"], 4), d["sustained"]["clocks"],
      "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches_per_step"], "clocks", d["clocks"])
print("inference", round(d["inference"]["value"]), "e2e", round(d["inference"]["e2e"]["value"]), "cpu", round(d["cpu_baseline"]["value"]), d["cpu_baseline"]["kind"], "c1", round(d["cpu_baseline"]["c1"]["value"]))
PY
