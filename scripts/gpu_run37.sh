#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export LIFTER_DP_TIMEOUT_S=60
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r02_bench_train_dp4.json 2> gpurun_out/r02_bench_train_dp4.err; echo "train dp4 rc=$?"
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 4 --steps 20 --warmup 5 --workload vit > gpurun_out/r02_bench_vit_dp4.json 2> gpurun_out/r02_bench_vit_dp4.err; echo "vit dp4 rc=$?"
python - <<'PY'
import json
for f in ("gpurun_out/r02_bench_train_dp4.json", "gpurun_out/r02_bench_vit_dp4.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "dp_check", d["dp_check"]["ok"])
    except Exception as e:
        print(f, "no line", e)
PY
