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
timeout 600 python -m pytest "tests/test_dp_gpu.py" -k "8-bench" -m gpu -q -p no:cacheprovider --timeout=400 --timeout-method=thread 2>&1 | tail -2
