"""Probe of the data-parallel step at bench shapes (prints progress; run under `timeout`)."""
import importlib
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.getcwd())
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3d_poseestimation_b200")
trainer = importlib.import_module("3d_poseestimation_b200.trainer")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(1)
model = pkg.LinearModel(34, 51, linear_size=1024, num_stage=2, p_dropout=0.5, BN=True).to(dev)
tr = trainer.LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adam", seed=2, world_size=world, rank=rank)
x, y = torch.randn(B, 34, device=dev), torch.randn(B, 51, device=dev)
for i in range(3):
    tr.step(x, y)
    torch.cuda.synchronize()
    print(f"[{rank}] step {i} done, loss {tr.loss_value():.4f}", flush=True)
for n in (5, 20):
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        tr.step(x, y)
    e1.record()
    torch.cuda.synchronize()
    print(f"[{rank}] {n} back-to-back steps: {e0.elapsed_time(e1) / n * 1e3:.1f} us/step", flush=True)
if len(sys.argv) > 2 and sys.argv[2] == "trace":
    # per-kernel device times (run with LIFTER_DP_GRAPH=0 LIFTER_WGRAD_STREAM=0: eager, single stream)
    import ctypes as C
    from collections import defaultdict
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    lib = cabi.lib()
    tot, cnt = defaultdict(float), defaultdict(int)
    N = 5
    for _ in range(N):
        dist.barrier()
        torch.cuda.synchronize()
        lib.lifter_trace_begin(cabi.stream_ptr())
        tr.step(x, y)
        buf = C.create_string_buffer(1 << 16)
        lib.lifter_trace_end(buf, len(buf))
        for line in buf.value.decode().splitlines():
            name, us = line.rsplit(" ", 1)
            tot[name] += float(us)
            cnt[name] += 1
    if rank == 0:
        ssum = sum(tot.values()) / N
        print(f"traced step: {ssum:.1f} us over {sum(cnt.values()) // N} launches")
        for k in sorted(tot, key=lambda k: -tot[k]):
            print(f"{k:36s} n/step={cnt[k] / N:5.1f} total/step={tot[k] / N:8.1f} us  avg={tot[k] / cnt[k]:7.2f} us")
g = torch.randn(4296768, device=dev)
for _ in range(3):
    dist.all_reduce(g)
torch.cuda.synchronize()
e0.record()
for _ in range(20):
    dist.all_reduce(g)
e1.record()
torch.cuda.synchronize()
print(f"[{rank}] NCCL all-reduce of {g.numel() * 4 / 1e6:.1f} MB: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us", flush=True)
dist.barrier()
dist.destroy_process_group()
