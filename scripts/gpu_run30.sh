#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export LIFTER_DP_TIMEOUT_S=30
for g in 1 0; do
DP_CASE=vit DP_USE_GRAPH=$g timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$g tests/dp_worker.py > gpurun_out/r02_dp_vit_$g.log 2>&1; echo "dp vit graph=$g rc=$?"; grep "rank\|Error\|error" gpurun_out/r02_dp_vit_$g.log | head -8
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 10 --warmup 3 --workload vit > gpurun_out/r02_bench_vit_dp2.json 2> gpurun_out/r02_bench_vit_dp2.err; echo "vit dp2 rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r02_bench_vit_dp2.json") if l.startswith("{")][-1])
print("ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches_per_step"], "dp_check", {k: d["dp_check"][k] for k in ("ok", "loss_rel_err", "update_rel_diff", "replicas_identical")})
PY
tail -3 gpurun_out/r02_bench_vit_dp2.err
