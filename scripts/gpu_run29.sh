#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_train_dp2_final.json 2> gpurun_out/r02_bench_train_dp2_final.err; echo "dp2 bench rc=$?"
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/r02_bench_train_dp2_final.json") if l.startswith("{")][-1])
print("ms", round(d["ms_per_step"], 4), "value", round(d["value"]), "dp_check", {k: d["dp_check"][k] for k in ("ok", "loss_rel_err", "update_rel_diff", "replicas_identical")})
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 10 --warmup 3 --workload vit > gpurun_out/r02_bench_vit_dp2.json 2> gpurun_out/r02_bench_vit_dp2.err; echo "vit x2 rc=$?"; cut -c1-300 gpurun_out/r02_bench_vit_dp2.json; tail -2 gpurun_out/r02_bench_vit_dp2.err
timeout 600 python -m pytest tests/test_dp_gpu.py -m gpu -q -p no:cacheprovider --timeout=300 --timeout-method=thread -k "world2 or w2 or 2" 2>&1 | tail -3
