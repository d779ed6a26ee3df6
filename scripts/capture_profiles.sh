#!/bin/bash
# Runs on the GPU box (gpurun): full GPU test suite, default bench, then the ncu passes of B200_PROFILING.md.
# usage: scripts/capture_profiles.sh <tag>      (outputs under gpurun_out/)
tag=${1:-v3}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_$tag.log
tail -3 $out/pytest_$tag.log
timeout 300 python bench.py > $out/bench_train_$tag.json 2> $out/bench_train_$tag.err; echo "bench rc=$?"
timeout 300 python bench.py --workload eval > $out/bench_eval_$tag.json 2> $out/bench_eval_$tag.err; echo "bench eval rc=$?"
timeout 300 python bench.py --workload wide --steps 50 > $out/bench_wide_$tag.json 2> $out/bench_wide_$tag.err; echo "bench wide rc=$?"
timeout 300 python bench.py --impl reference --steps 5 > $out/bench_ref_$tag.json 2>/dev/null; echo "bench ref rc=$?"
cat $out/bench_train_$tag.json $out/bench_wide_$tag.json
# launch lists (cold-cache, serialised: shares only)
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
  --log-file $out/r01_train_launches_$tag.csv python bench.py --steps 3 --warmup 3 > $out/ncu_train_$tag.log 2>&1
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
  --log-file $out/r01_eval_launches_$tag.csv python bench.py --workload eval --steps 3 --warmup 3 > $out/ncu_eval_$tag.log 2>&1
# one full capture of the dominant kernel (the 2-SM 256-wide GEMM) and one of the elementwise kernels
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:lifter_gemm_kernel -s 30 -c 1 \
  -o $out/prof_gemm_train_$tag -f python bench.py --steps 3 --warmup 3 > $out/ncu_full_$tag.log 2>&1
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:bn_apply_fwd_kernel -s 10 -c 1 \
  -o $out/prof_apply_$tag -f python bench.py --steps 3 --warmup 3 > $out/ncu_full_apply_$tag.log 2>&1
ls -la $out | tail -20
