#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for v in 0 1; do
  LIFTER_GEMM_NARROW=$v LIFTER_BENCH_NO_CPU=1 timeout 300 python bench.py --steps 200 --warmup 5 > gpurun_out/r02_bench_train_10_$v.json 2> gpurun_out/r02_bench_train_10_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/r02_bench_train_10_$v.json'));print('NARROW=$v ms/step',round(d['ms_per_step'],4),'sustained',round(d['sustained']['ms_per_step'],4),'gemm us',round(d['roofline']['us_per_launch'],2),'frac',round(d['roofline']['frac'],3))" || tail -3 gpurun_out/r02_bench_train_10_$v.err
  LIFTER_GEMM_NARROW=$v LIFTER_BENCH_NO_CPU=1 timeout 300 python bench.py --workload eval --steps 50 --warmup 5 > gpurun_out/r02_bench_eval_10_$v.json 2> gpurun_out/r02_bench_eval_10_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/r02_bench_eval_10_$v.json'));print('NARROW=$v eval ms/step',round(d['ms_per_step'],4),'gemm us',round(d['roofline']['us_per_launch'],2),'frac',round(d['roofline']['frac'],3))" || tail -3 gpurun_out/r02_bench_eval_10_$v.err
done
LIFTER_GEMM_NARROW=1 timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_bench_shapes_gpu.py -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread 2>&1 | tail -3
