#!/usr/bin/env bash
# round 2, 8-GPU call: DP test at world 8, breakdown of the step, bench lines with dp_check for train / phase5 / wide
set -u
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
: > gpurun_out/r02_dp_exchange_ab_dp$N.jsonl
for v in "default" "LIFTER_DP_EXCHANGE_SERIAL=1" "default" "LIFTER_DP_EXCHANGE_SERIAL=1" "LIFTER_DP_DEBUG_NO_EXCHANGE=1"; do
  if [ "$v" = "default" ]; then e=""; else e="$v"; fi
  env $e VARIANT="$v" timeout 300 $TR scripts/dp_breakdown.py 2>/dev/null | grep '^{' >> gpurun_out/r02_dp_exchange_ab_dp$N.jsonl
done
cut -c1-110 gpurun_out/r02_dp_exchange_ab_dp$N.jsonl
for w in train; do
  timeout 600 $TR bench.py --gpus $N --workload $w --steps 50 --warmup 5 > gpurun_out/r02_bench_${w}_dp$N.json 2> gpurun_out/r02_bench_${w}_dp$N.err
  echo "bench $w dp$N rc=$?"
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/r02_bench_${w}_dp$N.json") if l.startswith("{")][-1])
    print("$w", "ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "sus", round(d["sustained"]["ms_per_step"], 4), d["sustained"]["clocks"], "e2e", round(d["e2e"]["value"]))
    print("   dp_check", {k: v for k, v in d["dp_check"].items() if k not in ("tolerance", "losses_dp", "losses_single_process")})
except Exception as e:
    print("$w unreadable", e)
PY
done
