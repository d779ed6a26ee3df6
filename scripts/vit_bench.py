"""Throughput of the MyViT eval-mode forward on one B200: poses/s device-resident with CUDA events,
model TFLOP/s (algorithmic FLOP per pose: every Linear + the two attention products), and a per-kernel time split from
the library's trace hook.  Default 65,536 poses, the batch of BASELINE.json configs[1]."""
import ctypes as C
import importlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3d_poseestimation_b200")
cabi = importlib.import_module("3d_poseestimation_b200._cabi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
torch.manual_seed(0)
m = pkg.MyViT().cuda().eval()
H, N, nb, dh = 256, 17, 2, 64
flop_pose = 2 * N * (2 * H + nb * (H * 3 * H + H * H + 2 * H * 4 * H) + H * (H // 2) + (H // 2) * 3) + nb * 4 * N * N * H
x = torch.randn(B, 17, 2, device="cuda")
for chunk in (1024, 2048, 4096, 8192):
    for _ in range(3):
        m.infer(x, chunk_poses=chunk)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        m.infer(x, chunk_poses=chunk)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(json.dumps({"workload": f"MyViT eval forward, {B} poses, chunk {chunk}", "ms": round(ms, 3), "poses_per_s": round(B / ms * 1e3),
                      "flop_per_pose": flop_pose, "model_tflops": round(B * flop_pose / ms / 1e9, 1)}))
# per-kernel split of one chunk (events serialise the launches: shares, not absolute times)
cabi.check(cabi.lib().lifter_trace_begin(cabi.stream_ptr()))
m.infer(x[:2048], chunk_poses=2048)
buf = C.create_string_buffer(1 << 16)
n = cabi.lib().lifter_trace_end(buf, len(buf))
agg = {}
for line in buf.value.decode().splitlines():
    k, v = line.rsplit(" ", 1)
    agg[k] = agg.get(k, 0.0) + float(v)
tot = sum(agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
    print(f"{k:28s} {v:9.1f} us  {100 * v / tot:5.1f} %")
