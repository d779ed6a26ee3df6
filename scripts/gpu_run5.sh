#!/usr/bin/env bash
# round 2, GPU call 5 (2 GPUs): data-parallel test matrix at world 2, device-guard test, dp_check inside bench, warm trace
set -u
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_dp_gpu.py tests/test_bench_shapes_gpu.py::test_model_on_another_device_than_the_current_one -m gpu -q -p no:cacheprovider --timeout=400 --timeout-method=thread > gpurun_out/r02_tests_5_dp.log 2>&1
echo "dp pytest rc=$?"; grep -h "passed\|failed\|^FAILED\|^ERROR\|rank 0\]" gpurun_out/r02_tests_5_dp.log | head -30
for w in train phase5 wide; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload $w --steps 50 --warmup 5 \
      > gpurun_out/r02_bench_${w}_dp2.json 2> gpurun_out/r02_bench_${w}_dp2.err
  echo "bench $w dp2 rc=$?"; tail -c 500 gpurun_out/r02_bench_${w}_dp2.err | tail -4
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/r02_bench_${w}_dp2.json") if l.startswith("{")][-1])
    print("$w", "ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "sus", round(d["sustained"]["ms_per_step"], 4), d["sustained"]["clocks"], "e2e", round(d["e2e"]["value"]))
    print("   dp_check", {k: v for k, v in d["dp_check"].items() if k not in ("tolerance",)})
except Exception as e:
    print("$w unreadable", e)
PY
done
timeout 300 python scripts/trace_train.py > gpurun_out/r02_trace_train.txt 2>&1; cat gpurun_out/r02_trace_train.txt
