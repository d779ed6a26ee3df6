#!/usr/bin/env bash
# round 2, GPU call 4 (after the machine swap): no-alloc test alone, full suite, bench lines, launch list, full captures
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider --timeout=120 --timeout-method=thread > gpurun_out/r02_tests_4_noalloc_alone.log 2>&1
echo "no-alloc alone rc=$?"; tail -3 gpurun_out/r02_tests_4_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider --timeout=200 --timeout-method=thread > gpurun_out/r02_tests_4.log 2>&1
echo "pytest rc=$?"; grep -h "passed\|failed\|^FAILED\|^ERROR\|Timeout" gpurun_out/r02_tests_4.log | head -30
timeout 600 python bench.py --steps 200 --warmup 5 > gpurun_out/r02_bench_train_4.json 2> gpurun_out/r02_bench_train_4.err; echo "bench train rc=$?"
for w in eval phase5 wide; do
  timeout 600 python bench.py --workload $w --steps 50 --warmup 5 > gpurun_out/r02_bench_${w}_4.json 2> gpurun_out/r02_bench_${w}_4.err; echo "bench $w rc=$?"
done
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r02_reference_arm_4.json 2>/dev/null; echo "ref rc=$?"
python - <<'PY'
import json
for w in ("train", "eval", "phase5", "wide"):
    try:
        d = json.load(open(f"gpurun_out/r02_bench_{w}_4.json"))
        s = d["sustained"]
        print(w, "ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "sus", round(s["ms_per_step"], 4), s["clocks"]["sm_mhz"], s["clocks"]["reasons"],
              "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches_per_step"], "roof", round(d["roofline"]["frac"], 3),
              "step_frac", {k: round(v, 3) for k, v in d["step_roofline_frac"].items()}, "cpu", d["cpu_baseline"] and round(d["cpu_baseline"]["value"]))
    except Exception as e:
        print(w, "unreadable", e)
PY
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_phase5_launches.csv \
    python bench.py --workload phase5 --steps 3 --warmup 3 > gpurun_out/r02_ncu_4_phase5.log 2>&1
python scripts/launch_summary.py gpurun_out/r02_phase5_launches.csv > gpurun_out/r02_phase5_launches_summary.txt 2>&1; head -12 gpurun_out/r02_phase5_launches_summary.txt
K='regex:lifter_gemm_kernel<\(int\)256, \(bool\)1'
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "$K" \
    --launch-skip 40 --launch-count 1 -f -o gpurun_out/r02_train_gemm python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_4_train_full.log 2>&1; echo "ncu train rc=$?"
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "$K" \
    --launch-skip 14 --launch-count 1 -f -o gpurun_out/r02_eval_gemm python bench.py --workload eval --steps 3 --warmup 3 > gpurun_out/r02_ncu_4_eval_full.log 2>&1; echo "ncu eval rc=$?"
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:bn_apply_fwd_kernel -s 12 -c 1 \
    -f -o gpurun_out/r02_apply python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_4_apply_full.log 2>&1; echo "ncu apply rc=$?"
for n in train_gemm eval_gemm apply; do
  ncu -i gpurun_out/r02_${n}.ncu-rep --page raw --csv > gpurun_out/r02_${n}_ncu_full_raw.csv 2>/dev/null
done
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out/*full_raw.csv
