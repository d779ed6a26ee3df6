"""Per-kernel device times of one training step with warm caches (debug trace of the C ABI)."""
import ctypes as C
import importlib
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("3d_poseestimation_b200")
cabi = importlib.import_module("3d_poseestimation_b200._cabi")
trainer = importlib.import_module("3d_poseestimation_b200.trainer")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
dev = torch.device("cuda")
torch.manual_seed(1)
model = pkg.LinearModel(34, 51, linear_size=1024, num_stage=2, p_dropout=0.5, BN=True).to(dev)
tr = trainer.LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adam", seed=2, use_graph=False)
trg = trainer.LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adam", seed=2, use_graph=True)
x, y = torch.randn(B, 34, device=dev), torch.randn(B, 51, device=dev)
for _ in range(5):
    tr.step(x, y)
torch.cuda.synchronize()
lib = cabi.lib()
tot, cnt, order = defaultdict(float), defaultdict(int), []
N = 5
for _ in range(N):
    lib.lifter_trace_begin(cabi.stream_ptr())
    tr.step(x, y)
    buf = C.create_string_buffer(1 << 16)
    n = lib.lifter_trace_end(buf, len(buf))
    for line in buf.value.decode().splitlines():
        name, us = line.rsplit(" ", 1)
        if name not in tot:
            order.append(name)
        tot[name] += float(us)
        cnt[name] += 1
s = sum(tot.values()) / N
print(f"traced step: {s:.1f} us over {sum(cnt.values()) // N} launches (events serialise the launches)")
for k in sorted(tot, key=lambda k: -tot[k]):
    print(f"{k:36s} n/step={cnt[k] / N:5.1f} total/step={tot[k] / N:8.1f} us  avg={tot[k] / cnt[k]:7.2f} us  share={tot[k] / N / s * 100:5.1f}%")
# untraced step time for comparison
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    tr.step(x, y)
e1.record()
torch.cuda.synchronize()
print(f"untraced: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us/step")
import time
t0 = time.perf_counter()
for _ in range(20):
    tr.step(x, y)
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"host enqueue time: {(t1 - t0) / 20 * 1e6:.1f} us/step")
for _ in range(3):
    trg.step(x, y)
torch.cuda.synchronize()
e0.record()
for _ in range(20):
    trg.step(x, y)
e1.record()
torch.cuda.synchronize()
print(f"CUDA-graph step: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us/step")
