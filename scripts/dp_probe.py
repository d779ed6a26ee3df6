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
