#!/usr/bin/env bash
# round 2, GPU call 3: full suite (phase-5 trainer, heat maps, bench shapes), phase5 / wide bench lines, GEMM full captures
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider > gpurun_out/r02_tests_3.log 2>&1
echo "pytest rc=$?"; grep -h "passed\|failed\|^FAILED\|^ERROR" gpurun_out/r02_tests_3.log | head -30
for w in phase5 wide; do
  timeout 600 python bench.py --workload $w --steps 50 --warmup 5 > gpurun_out/r02_bench_${w}_3.json 2> gpurun_out/r02_bench_${w}_3.err
  echo "bench $w rc=$?"; tail -c 400 gpurun_out/r02_bench_${w}_3.err
  python -c "
import json;d=json.load(open('gpurun_out/r02_bench_${w}_3.json'));print('$w ms/step',round(d['ms_per_step'],4),'value',round(d['value']),'sustained',round(d['sustained']['ms_per_step'],4), d['sustained']['clocks'],'e2e',round(d['e2e']['ms_per_step'],4),'launches',d['gpu_launches_per_step'],'tflops',round(d['model_tflops'],1),'cpu',d['cpu_baseline'] and round(d['cpu_baseline']['value']))"
done
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02_reference_arm_3.json 2>/dev/null; echo "ref rc=$?"; cut -c1-300 gpurun_out/r02_reference_arm_3.json
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'lifter_gemm_kernel<\(int\)256, \(bool\)1' \
    --launch-skip 40 --launch-count 1 -f -o gpurun_out/r02_train_gemm python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_3_train_full.log 2>&1
echo "ncu train full rc=$?"
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'lifter_gemm_kernel<\(int\)256, \(bool\)1' \
    --launch-skip 14 --launch-count 1 -f -o gpurun_out/r02_eval_gemm python bench.py --workload eval --steps 3 --warmup 3 > gpurun_out/r02_ncu_3_eval_full.log 2>&1
echo "ncu eval full rc=$?"
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:bn_apply_fwd_kernel -s 12 -c 1 \
    -f -o gpurun_out/r02_apply python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_3_apply_full.log 2>&1
for n in train_gemm eval_gemm apply; do
  ncu -i gpurun_out/r02_${n}.ncu-rep --page raw --csv > gpurun_out/r02_${n}_ncu_full_raw.csv 2>/dev/null
done
ls -la gpurun_out/*.ncu-rep gpurun_out/*full_raw.csv
