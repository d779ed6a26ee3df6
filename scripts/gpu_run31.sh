#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export LIFTER_DP_TIMEOUT_S=60
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 5 --workload vit > gpurun_out/r02_bench_vit_dp8.json 2> gpurun_out/r02_bench_vit_dp8.err; echo "vit dp8 rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r02_bench_vit_dp8.json") if l.startswith("{")][-1])
print("ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "sustained", round(d["sustained"]["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]), "dp_check", {k: d["dp_check"][k] for k in ("ok", "loss_rel_err", "update_rel_diff", "replicas_identical")})
PY
DP_CASE=vit timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29534 tests/dp_worker.py 2>&1 | grep "rank" | head -4
