#!/usr/bin/env bash
# round 2, GPU call 1: full GPU suite, training bench with and without the grid-barrier BatchNorm kernels, launch list
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout=300 -p no:cacheprovider > gpurun_out/r02_tests_1.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_tests_1.log
tail -30 gpurun_out/r02_tests_1.log
timeout 600 python bench.py --steps 100 --warmup 5 > gpurun_out/r02_bench_train_1.json 2> gpurun_out/r02_bench_train_1.err
echo "bench rc=$?"; tail -c 600 gpurun_out/r02_bench_train_1.err
LIFTER_FUSE_BN=0 LIFTER_BENCH_NO_CPU=1 timeout 300 python bench.py --steps 100 --warmup 5 > gpurun_out/r02_bench_train_1_nofuse.json 2> gpurun_out/r02_bench_train_1_nofuse.err
echo "bench nofuse rc=$?"
python - <<'PY'
import json
for f in ("r02_bench_train_1", "r02_bench_train_1_nofuse"):
    try:
        d = json.load(open(f"gpurun_out/{f}.json"))
        print(f, "ms/step", round(d["ms_per_step"], 4), "value", round(d["value"]), "sustained", d["sustained"] and round(d["sustained"]["ms_per_step"], 4),
              "e2e", round(d["e2e"]["ms_per_step"], 4), "launches", d["gpu_launches_per_step"], "roof", d["roofline"] and round(d["roofline"]["frac"], 3),
              "lib", d["roofline"] and d["roofline"].get("library_point"))
    except Exception as e:
        print(f, "unreadable", e)
PY
LIFTER_BENCH_NO_CPU=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_train_launches_1.csv \
    python bench.py --steps 3 --warmup 3 > gpurun_out/r02_ncu_1.log 2>&1
echo "ncu rc=$?"
python scripts/launch_summary.py gpurun_out/r02_train_launches_1.csv > gpurun_out/r02_train_launches_1_summary.txt 2>&1
head -30 gpurun_out/r02_train_launches_1_summary.txt
