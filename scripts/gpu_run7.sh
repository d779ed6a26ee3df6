#!/usr/bin/env bash
# round 2, GPU call 7: whole-layer forward kernel -- its tests (hard timeout: a grid barrier that cannot complete traps after
# ~4 s, the wrapper is a second line of defence), the suite, A/B of the training step, launch list
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_layer_kernel_gpu.py -m gpu -q -x -p no:cacheprovider --timeout=120 --timeout-method=thread > gpurun_out/r02_tests_7_layer.log 2>&1
rc=$?; echo "layer tests rc=$rc"; grep -h "passed\|failed\|^FAILED\|^ERROR\|^E  " gpurun_out/r02_tests_7_layer.log | head -20
if [ $rc -ne 0 ]; then
  LIFTER_FUSE_LAYER=0 timeout 300 python -m pytest tests/test_layer_kernel_gpu.py -m gpu -q -x -p no:cacheprovider --timeout=120 --timeout-method=thread 2>&1 | tail -5
fi
for v in 1 0; do
  LIFTER_FUSE_LAYER=$v LIFTER_BENCH_NO_CPU=1 timeout 300 python bench.py --steps 200 --warmup 5 > gpurun_out/r02_bench_train_7_layer$v.json 2> gpurun_out/r02_bench_train_7_layer$v.err
  python -c "
import json;d=json.load(open('gpurun_out/r02_bench_train_7_layer$v.json'));print('FUSE_LAYER=$v ms/step',round(d['ms_per_step'],4),'sustained',round(d['sustained']['ms_per_step'],4),'launches',d['gpu_launches_per_step'],'e2e',round(d['e2e']['ms_per_step'],4))" || tail -5 gpurun_out/r02_bench_train_7_layer$v.err
done
if [ $rc -eq 0 ]; then
  timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_7.log 2>&1
  echo "suite rc=$?"; grep -h "passed\|failed\|^FAILED\|^ERROR\|Timeout" gpurun_out/r02_tests_7.log | head -20
  LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_train_launches_7.csv \
      python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_7.log 2>&1
  python scripts/launch_summary.py gpurun_out/r02_train_launches_7.csv > gpurun_out/r02_train_launches_7_summary.txt 2>&1; head -16 gpurun_out/r02_train_launches_7_summary.txt
fi
