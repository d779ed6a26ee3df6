"""Where the end-to-end step time of run_epoch goes: host enqueue time vs device time (one GPU)."""
import importlib
import os
import sys
import time

import torch

sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("3d_poseestimation_b200")
trainer = importlib.import_module("3d_poseestimation_b200.trainer")
dev = torch.device("cuda")
torch.manual_seed(1)
model = pkg.LinearModel(34, 51, linear_size=1024, num_stage=2, p_dropout=0.5, BN=True).to(dev)
tr = trainer.LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adam", seed=2)
B = 8192
xh, yh = torch.randn(B, 34).pin_memory(), torch.randn(B, 51).pin_memory()
xd, yd = xh.to(dev), yh.to(dev)
for _ in range(10):
    tr.step(xd, yd)
torch.cuda.synchronize()
n = 200
t0 = time.perf_counter()
for _ in range(n):
    tr.step(xd, yd)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"step(): host enqueue {(t1 - t0) / n * 1e6:.1f} us/step, total {(t2 - t0) / n * 1e6:.1f} us/step")
tr.run_epoch([(xh, yh)] * 20)
for mode in ("full", "no_loss_read"):
    if mode == "no_loss_read":
        orig = torch.Tensor.copy_
    t0 = time.perf_counter()
    losses = tr.run_epoch([(xh, yh)] * n)
    t2 = time.perf_counter()
    print(f"run_epoch[{mode}]: {(t2 - t0) / n * 1e6:.1f} us/step")
    break
# the pieces on their own
cur = torch.cuda.current_stream()
pinned = torch.empty(1).pin_memory()
t0 = time.perf_counter()
for _ in range(n):
    tr.step(xd, yd)
    pinned.copy_(tr.loss_out[:1], non_blocking=True)
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"step() + per-step D2H of the loss on the same stream: {(t2 - t0) / n * 1e6:.1f} us/step")
copy = torch.cuda.Stream()
xs, ys = torch.empty_like(xd), torch.empty_like(yd)
t0 = time.perf_counter()
for _ in range(n):
    with torch.cuda.stream(copy):
        xs.copy_(xh, non_blocking=True)
        ys.copy_(yh, non_blocking=True)
    tr.step(xd, yd)
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"step() + concurrent uploads on a copy stream (no dependency): {(t2 - t0) / n * 1e6:.1f} us/step")
