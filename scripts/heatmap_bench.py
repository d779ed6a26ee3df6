"""Heat-map target kernel timed alone: GB/s against the 1 MB-per-key-point store floor."""
import importlib
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pl = importlib.import_module("3d_poseestimation_b200.pose_losses")
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pk = bench.peaks()
for poses in (8, 64, 256):
    kp = torch.rand(poses, 17, 3, device="cuda") * 2 - 1
    outs = [torch.empty(poses, 17, 64, 64, 64, device="cuda") for _ in range(2 if poses >= 64 else 8)]   # rotate past L2
    for o in outs:
        pl.keypoints_to_heatmap_3D(kp, out=o, check=False)
    torch.cuda.synchronize()
    n = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        pl.keypoints_to_heatmap_3D(kp, out=outs[i % len(outs)], check=False)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    by = poses * 17 * 64 ** 3 * 4
    print(json.dumps({"kernel": "heatmap_kernel", "poses": poses, "bytes": by, "us_per_launch": us,
                      "roofline": {"bound": "hbm", "achieved": by / us / 1e3, "peak": pk["hbm"], "unit": "GB/s",
                                   "frac": by / us / 1e3 / pk["hbm"], "note": "store-only kernel against the read+write copy rate"}}))
