#!/usr/bin/env bash
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dp_gpu.py -m gpu -q -p no:cacheprovider --timeout=400 --timeout-method=thread 2>&1 | tail -3
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611"
: > gpurun_out/r02_dp_breakdown4_dp$N.jsonl
for v in "default" "LIFTER_DP_DEBUG_NO_EXCHANGE=1"; do
  if [ "$v" = "default" ]; then e=""; else e="$v"; fi
  env $e VARIANT="$v" timeout 300 $TR scripts/dp_breakdown.py 2>/dev/null | grep '^{' >> gpurun_out/r02_dp_breakdown4_dp$N.jsonl
done
cut -c1-120 gpurun_out/r02_dp_breakdown4_dp$N.jsonl
