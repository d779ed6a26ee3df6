"""Phase timeline of the forward GEMMs / whole-layer kernels of one training forward (debug hook: %globaltimer stamps).
usage: python scripts/layer_timeline.py [B]      (LIFTER_FUSE_LAYER=0 for the plain GEMMs)"""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
H, S = 1024, 2
L = 2 * S + 1
tl = torch.zeros(L, 148, 32, dtype=torch.int64, device="cuda")
os.environ["LIFTER_DEBUG_TIMELINE_PTR"] = hex(tl.data_ptr())
pkg = importlib.import_module("3d_poseestimation_b200")
cabi = importlib.import_module("3d_poseestimation_b200._cabi")
ag = importlib.import_module("3d_poseestimation_b200._autograd")
torch.manual_seed(1)
m = pkg.LinearModel(34, 51, linear_size=H, num_stage=S, p_dropout=0.5, BN=True).cuda().train()
x = torch.randn(B, 34, device="cuda")
for _ in range(3):
    ag.run_forward(m, x, True)
torch.cuda.synchronize()
tl.zero_()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ag.run_forward(m, x, True)
e1.record()
torch.cuda.synchronize()
print(f"forward of {B} rows: {e0.elapsed_time(e1) * 1e3:.1f} us (one call, includes launch latency)")
t = tl.cpu().numpy().astype(np.float64)
names = {0: "start", 1: "setup done", 2: "first MMA", 4: "MMA tile0 done", 5: "epi sees tile0", 6: "epi tile0 done", 7: "MMA tile1 done",
         8: "epi sees tile1", 9: "epi tile1 done", 27: "pass1 done", 28: "after barrier 1", 29: "finalize share done",
         30: "after barrier 2", 31: "pass 2 issued"}
for l in range(L):
    a = t[l]
    live = a[:, 0] > 0
    if not live.any():
        continue
    t0 = a[live, 0].min()
    print(f"layer {l}: {int(live.sum())} CTAs stamped; first start = 0")
    for slot, nm in names.items():
        v = a[live, slot]
        v = v[v > 0]
        if len(v):
            print(f"   {nm:22s} min {((v.min() - t0) / 1e3):7.2f}  median {((np.median(v) - t0) / 1e3):7.2f}  max {((v.max() - t0) / 1e3):7.2f} us  (n={len(v)})")
    if l + 1 < L and (t[l + 1][:, 0] > 0).any():
        print(f"   next layer's first start: {((t[l + 1][t[l + 1][:, 0] > 0, 0].min() - t0) / 1e3):7.2f} us")
