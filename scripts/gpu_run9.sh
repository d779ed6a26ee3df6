#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_train_gpu.py tests/test_bench_shapes_gpu.py tests/test_layer_kernel_gpu.py tests/test_configs_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread 2>&1 | tail -4
for v in "LIFTER_APPLY_BLOCK=1" "LIFTER_APPLY_SHALLOW=0" "LIFTER_APPLY_SHALLOW=1"; do
  env $v LIFTER_BENCH_NO_CPU=1 timeout 300 python bench.py --steps 200 --warmup 5 > gpurun_out/r02_bench_train_9.json 2> gpurun_out/r02_bench_train_9.err
  python -c "
import json;d=json.load(open('gpurun_out/r02_bench_train_9.json'));print('$v ms/step',round(d['ms_per_step'],4),'sustained',round(d['sustained']['ms_per_step'],4))" || tail -3 gpurun_out/r02_bench_train_9.err
done
for v in "LIFTER_APPLY_SHALLOW=0" "LIFTER_APPLY_SHALLOW=1"; do
env $v LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r02_train_launches_9.csv \
    python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_9.log 2>&1
python scripts/launch_summary.py gpurun_out/r02_train_launches_9.csv 2>&1 | grep -i "apply\|total device"
done
