#!/usr/bin/env bash
# round 2, GPU call 2: per-kernel times of the default (separate-launch) step, A/B of the grid-barrier kernels,
# heat-map kernel, full captures of the final GEMM at M = 8,192 and M = 65,536
set -u
mkdir -p gpurun_out
T="tests/test_bench_shapes_gpu.py::test_config2_gradients_at_bench_shape tests/test_train_gpu.py::test_train_step_matches_reference_golden tests/test_heatmap_gpu.py"
timeout 600 python -m pytest $T -q --timeout=300 -p no:cacheprovider -s > gpurun_out/r02_tests_2a.log 2>&1; echo "default rc=$?"
LIFTER_FUSE_BN_FWD=1 LIFTER_FUSE_BN_BWD=1 timeout 600 python -m pytest $T -q --timeout=300 -p no:cacheprovider -s > gpurun_out/r02_tests_2b.log 2>&1; echo "fused rc=$?"
grep -h "vs rounding model\|passed\|failed\|dx vs model" gpurun_out/r02_tests_2a.log gpurun_out/r02_tests_2b.log | head -60
for cfg in "0 0" "1 0" "0 1"; do
  set -- $cfg
  LIFTER_FUSE_BN_FWD=$1 LIFTER_FUSE_BN_BWD=$2 LIFTER_BENCH_NO_CPU=1 timeout 300 python bench.py --steps 200 --warmup 5 > gpurun_out/r02_bench_train_2_f$1b$2.json 2> gpurun_out/r02_bench_train_2_f$1b$2.err
  python -c "
import json;d=json.load(open('gpurun_out/r02_bench_train_2_f$1b$2.json'));print('fwd=$1 bwd=$2 ms/step',round(d['ms_per_step'],4),'sustained',round(d['sustained']['ms_per_step'],4),'launches',d['gpu_launches_per_step'])"
done
python scripts/heatmap_bench.py > gpurun_out/r02_heatmap_bench.jsonl 2>&1; cat gpurun_out/r02_heatmap_bench.jsonl
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_train_launches_2.csv \
    python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_2.log 2>&1
python scripts/launch_summary.py gpurun_out/r02_train_launches_2.csv > gpurun_out/r02_train_launches_2_summary.txt 2>&1
head -24 gpurun_out/r02_train_launches_2_summary.txt
# full captures of the dominant GEMM, final build: training shape (M = 8,192) and inference shape (M = 65,536)
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'lifter_gemm_kernel<256, 1' \
    --launch-skip 40 --launch-count 1 -f -o gpurun_out/r02_train_gemm python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_2_train_full.log 2>&1
echo "ncu train full rc=$?"
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:'lifter_gemm_kernel<256, 1' \
    --launch-skip 14 --launch-count 1 -f -o gpurun_out/r02_eval_gemm python bench.py --workload eval --steps 3 --warmup 3 > gpurun_out/r02_ncu_2_eval_full.log 2>&1
echo "ncu eval full rc=$?"
for n in train eval; do
  ncu -i gpurun_out/r02_${n}_gemm.ncu-rep --page raw --csv > gpurun_out/r02_${n}_gemm_ncu_full_raw.csv 2>/dev/null
done
ls -la gpurun_out/*.ncu-rep gpurun_out/*full_raw.csv
