"""Device-resident inference time of 65,536 poses against the row-chunk size (chunks keep activations in L2)."""
import importlib
import os
import sys

import torch

sys.path.insert(0, os.getcwd())
pkg = importlib.import_module("3d_poseestimation_b200")
dev = torch.device("cuda")
torch.manual_seed(1)
m = pkg.LinearModel(34, 51, linear_size=1024, num_stage=2, p_dropout=0.5, BN=True).to(dev).eval()
B = 65536
x = torch.randn(B, 34, device=dev)
out = torch.empty(B, 51, device=dev)
for chunk in (0, 9472, 18944, 28416, 37888, 32768, 65536):
    for _ in range(5):
        m.infer(x, out=out, chunk_rows=chunk)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 40
    e0.record()
    for _ in range(n):
        m.infer(x, out=out, chunk_rows=chunk)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"chunk_rows={chunk:6d}: {ms * 1e3:7.1f} us  {B / ms / 1e3:7.1f} M poses/s", flush=True)
