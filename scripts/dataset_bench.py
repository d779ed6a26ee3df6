"""HBM-roofline reading of the dataset-side kernels: keypoint_stats (4 B read per element) and
process_data's apply pass (4 B read + 4 B write per element), CUDA events on the current stream, inputs rotated so that
no launch finds its data in the 126 MB L2.  Prints one JSON line per (kernel, size)."""
import importlib
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = importlib.import_module("3d_poseestimation_b200.pose_losses")
peak = 6534.1
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def timed(fn, n_sets, reps=20):
    for i in range(3):
        fn(i % n_sets)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i % n_sets)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3          # us per call (stats = 2 launches, apply = 1)


for frames, dim in ((390_000, 3), (390_000, 2), (4_000_000, 3)):
    nbytes = frames * 17 * dim * 4
    n_sets = max(2, int(512e6 // nbytes) + 1)
    xs = [torch.randn(frames, 17, dim, device="cuda") for _ in range(n_sets)]
    stats = P.keypoint_stats(xs[0])
    t_stats = timed(lambda i: P.keypoint_stats(xs[i]), n_sets)
    t_apply = timed(lambda i: P.process_data(xs[i], stats=stats, is_train=False, standardize=True), n_sets)
    t_both = timed(lambda i: P.process_data(xs[i], is_train=True, standardize=True), n_sets)
    for name, t, b in (("keypoint_stats (+finalize)", t_stats, nbytes), ("process_keypoints", t_apply, 2 * nbytes),
                       ("process_data(is_train, standardize)", t_both, 3 * nbytes)):
        print(json.dumps({"kernel": name, "frames": frames, "dims": dim, "us": round(t, 2), "algorithmic_bytes": b,
                          "achieved_gbs": round(b / t / 1e3, 1), "peak_gbs": peak, "frac": round(b / t / 1e3 / peak, 3)}))
    del xs
