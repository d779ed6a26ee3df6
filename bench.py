#!/usr/bin/env python
"""Benchmark of the B200-native lifter hot path (contract: see the task statement / DESIGN.md §Measurement).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload train|eval] [--impl ours|reference]

The default workload is BASELINE.json's metric, "lifted poses/sec (fwd+bwd)": the data-parallel training step of
configs[2] (8,192 poses per GPU); with one GPU the line also carries an ``inference`` object with the eval-mode
numbers of configs[1] (65,536 poses).  ``--workload eval`` makes configs[1] the headline instead.

* ``value``  : lifted poses/s over all ranks, inputs resident in HBM, CUDA-event time, max over ranks.
* ``e2e``    : the same metric through the public module API with HOST (pinned) inputs: H2D copy of the
               step's keypoints, the step, D2H of the step's result -- all inside the timed region.
* ``roofline``: the dominant kernel (the 1024-wide tcgen05 GEMM) timed alone with CUDA events, against the
               measured bf16 peak of MEASURED_PEAKS.json.
* ``cpu_baseline`` / ``--impl reference``: the oracle's PyTorch-fp32 CPU restatement of the reference module
               (the reference itself is Python and does not travel to the GPU box) on the host cores.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "3d_poseestimation_b200"

FLOP_PER_POSE_INFER = 8_562_688          # SURVEY §8d, H=1024, S=2, padding excluded
FLOP_PER_POSE_TRAIN = 25_618_432         # fwd + bwd (no input gradient)
I_DIM, O_DIM, H, S = 34, 51, 1024, 2
EVAL_BATCH = 65536                       # BASELINE.json configs[1]
TRAIN_BATCH = 8192                       # BASELINE.json configs[2], per GPU
WIDE = (16 * 34, 2048, 4)                # BASELINE.json configs[4]: i_dim, linear_size, num_stage
FLOP_PER_POSE_WIDE_TRAIN = 206_409_728


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sustained=1400.0, source="fallback")


class ClockSampler:
    """Samples nvidia-smi SM clocks and throttle reasons while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self.stop, self.thread = index, [], threading.Event(), None

    def _run(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.index)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([s.strip() for s in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------ CPU arm
def synth_batch(batch, seed):
    """Synthetic H36M-shaped keypoints from the repo statistics: normalised 2D in, normalised 3D target."""
    import numpy as np
    import torch
    st = dict(np.load(os.path.join(ROOT, "tests", "golden", "h36m_stats.npz")))
    m2, s2 = torch.from_numpy(st["mean_train_2d"]).float(), torch.from_numpy(st["std_train_2d"]).float()
    g = torch.Generator().manual_seed(seed)
    x = (m2 + s2 * torch.randn(batch, 17, 2, generator=g)).clamp(0, 1)
    xn = ((x - m2) / s2).reshape(batch, 34).contiguous()
    yn = torch.randn(batch, 51, generator=g)                       # target in normalised units: N(0,1) by construction
    return xn, yn


def cpu_arm(workload, steps, warmup, budget_s=25.0):
    """Times the oracle (CPU restatement of the reference LinearModel, PyTorch fp32) on the host cores."""
    import torch

    from oracle import adam_oracle, lifter_oracle as lo
    from oracle.philox import dropout_keep_mask
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    I_DIM, H, S = WIDE if workload == "wide" else (globals()["I_DIM"], globals()["H"], globals()["S"])
    params, buffers = lo.numpy_init(I_DIM, O_DIM, H, S, seed=1, randomize_bn=False)
    if workload == "eval":
        batch = 8192                                            # bounded sample of the 65,536-pose workload
        x, _ = synth_batch(batch, 0)

        def step():
            with torch.no_grad():
                lo.lifter_forward(params, buffers, x, num_stage=S, training=False)
        sample = f"{batch} of {EVAL_BATCH} poses per step, eval forward"
    else:
        batch = 512                                             # bounded sample of the 8,192-pose step
        x, y = synth_batch(batch, 0)
        if workload == "wide":
            x = torch.cat([synth_batch(batch, 100 * f)[0] for f in range(16)], dim=1).contiguous()
        masks = [dropout_keep_mask(batch, H, s, 0, 2, 0.5) for s in range(2 * S + 1)]
        m = {k: torch.zeros_like(v) for k, v in params.items()}
        v = {k: torch.zeros_like(w) for k, w in params.items()}
        state = {"t": 0}

        def step():
            r = lo.lifter_forward_backward(params, buffers, x, lambda o: ((o - y) ** 2).mean(), num_stage=S,
                                           training=True, p_dropout=0.5, masks=masks)
            state["t"] += 1
            for k in params:
                params[k], m[k], v[k] = adam_oracle.adam_step(params[k], r["grads"][k], m[k], v[k], state["t"], 1e-4)
        sample = f"{batch} of {TRAIN_BATCH} poses per step, fwd+MSE+bwd+Adam"
    for _ in range(max(1, min(warmup, 3))):
        step()
    times, t_begin = [], time.perf_counter()
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s:
            break
    ms = statistics.median(times) * 1e3
    return {"value": batch / (ms / 1e3), "unit": "poses/s", "cores": cores, "kind": "port", "sample": sample,
            "ms_per_step": ms, "steps_timed": len(times)}


# ------------------------------------------------------------------------------------------------ GPU arm
def gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    pkg = importlib.import_module(PKG)
    cabi = importlib.import_module(PKG + "._cabi")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cabi.check(cabi.lib().lifter_device_check())
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def measure(workload, steps, warmup, with_clocks):
        """One workload on this rank's GPU: device-resident throughput, end-to-end throughput, dominant kernel."""
        torch.manual_seed(1)
        model = pkg.LinearModel(I_DIM, O_DIM, linear_size=H, num_stage=S, p_dropout=0.5, BN=True).to(dev)
        if workload == "eval":
            batch = EVAL_BATCH
            model.eval()
            xh, _ = synth_batch(batch, rank)
            xh = xh.pin_memory()
            xd = xh.to(dev)
            out = torch.empty(batch, O_DIM, device=dev)
            outh = torch.empty(batch, O_DIM).pin_memory()
            launches_per_step = 1 + (1 + 2 * S + 1)           # prep_input + one GEMM per Linear (65,536 rows = one chunk)

            def step():
                model.infer(xd, out=out)

            def run_e2e(n):                                  # host poses in, host predictions out, every step
                model.infer_host_batches([xh] * n, outs=[outh] * n)      # one pipeline across the batches, one wait
                torch.cuda.current_stream().synchronize()
            h2d, d2h = xh.numel() * 4, outh.numel() * 4
            dtype = "fp16 operands, fp32 accumulate"
            cfg = {"workload": f"configs[1]: eval-mode inference, batch {batch} synthetic H36M 2D poses per GPU "
                               f"(17x2 -> 17x3, 1024-wide, 2 residual blocks)", "batch_per_gpu": batch,
                   "parallelism": f"replicas x{world}" if world > 1 else "single GPU",
                   "l2": "activations of one step (3 x 134 MB) exceed the 126 MB L2; no flush needed"}
            flop = FLOP_PER_POSE_INFER
        else:
            if workload == "wide":            # BASELINE.json configs[4]: 16-frame window in, 2048 wide, 4 blocks
                model = pkg.LinearModel(WIDE[0], O_DIM, linear_size=WIDE[1], num_stage=WIDE[2], p_dropout=0.5,
                                        BN=True).to(dev)
            trainer_mod = importlib.import_module(PKG + ".trainer")
            batch = TRAIN_BATCH
            tr = trainer_mod.LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adam", world_size=world, rank=rank,
                                           seed=2)
            xh, yh = synth_batch(batch, rank)
            if workload == "wide":            # 16 iid frames concatenated
                xh = torch.cat([synth_batch(batch, 100 * f + rank)[0] for f in range(16)], dim=1).contiguous()
            xh, yh = xh.pin_memory(), yh.pin_memory()
            xd, yd = xh.to(dev), yh.to(dev)
            launches_per_step = tr.launches_per_step(batch)

            def step():
                tr.step(xd, yd)

            def run_e2e(n):
                # the epoch loop over host batches: per step, an H2D copy of (x, y) and a D2H read of the loss
                losses = tr.run_epoch([(xh, yh)] * n)
                assert losses.numel() == n and bool(torch.isfinite(losses).all())
            h2d, d2h = (xh.numel() + yh.numel()) * 4, 4
            dtype = "bf16 operands, fp32 accumulate/master weights"
            par = f"dp{world}"
            if world > 1:
                par += (" (in-kernel SyncBN exchange + fused reduce-scatter/Adam/all-gather over NVLink symmetric memory)"
                        if tr._dp_adam else " (NCCL all-reduce fallback)")
            shape = "1024-wide, 2 residual blocks" if workload == "train" else \
                "widened lifter (16 x 34 inputs, 2048-wide, 4 residual blocks; configs[4])"
            cfg = {"workload": f"configs[{2 if workload == 'train' else 4}]: training step fwd + MSE + bwd + Adam, dropout 0.5, batch {batch}/GPU, "
                               f"synced BatchNorm, {shape}", "batch_per_gpu": batch,
                   "global_batch": batch * world, "parallelism": par,
                   "cuda_graph": bool(tr.use_graph),
                   "l2": "one step touches ~0.4 GB of saved activations and gradients per GPU (> 126 MB L2), so "
                         "consecutive steps do not find their inputs in L2; no explicit flush"}
            flop = FLOP_PER_POSE_TRAIN if workload == "train" else FLOP_PER_POSE_WIDE_TRAIN

        def timed(fn, steps, warmup):
            for _ in range(warmup):
                fn()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn()
            e1.record()
            barrier()
            ms = e0.elapsed_time(e1) / steps
            if world > 1:
                t = torch.tensor([ms], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = t.item()
            return ms

        step()                                                 # builds buffers / captures the step graph untimed
        barrier()
        if with_clocks:
            with ClockSampler(local_rank) as clk:
                ms = timed(step, steps, warmup)
                in_region = len(clk.samples)
                # nvidia-smi answers every ~0.1 s: a timed region shorter than ~1.5 s is extended with identical untimed
                # steps (same load).  The count derives from the max-over-ranks step time, so every rank runs the same
                # number of (collective) steps.
                for _ in range(max(0, int(1500.0 / max(ms, 1e-3)) - steps)):
                    step()
                torch.cuda.synchronize()
            clocks = clk.summary()
            clocks["samples_in_timed_region"] = in_region
        else:
            ms, clocks = timed(step, steps, warmup), None

        # end to end through the public API with host buffers (wall clock around synchronising steps)
        run_e2e(max(8, warmup))                                # also past the point where run_epoch's slots get their graphs
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(3, steps // 2)
        run_e2e(n_e2e)
        barrier()
        e2e_ms = (time.perf_counter() - t0) / n_e2e * 1e3
        if world > 1:
            t = torch.tensor([e2e_ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = t.item()

        roof = None
        if rank == 0:   # dominant kernel alone: the H x H tcgen05 GEMM at this workload's per-launch shape
            roof = gemm_roofline(cabi, dev, batch, cabi.FMT_F16 if workload == "eval" else cabi.FMT_BF16, pk,
                                 WIDE[1] if workload == "wide" else H)
        value = batch * world / (ms / 1e3)
        return {"metric": "lifted poses/sec" + (" (eval inference)" if workload == "eval" else " (fwd+bwd+Adam)"),
                "value": value, "unit": "poses/s", "n_gpus": world, "steps": steps, "warmup": warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype,
                "data": "synthetic (H36M-shaped keypoints from the repo's mean/std statistics, random-init weights)",
                "config": cfg, "clocks": clocks,
                "e2e": {"value": batch * world / (e2e_ms / 1e3), "unit": "poses/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms},
                "gpu_launches": launches_per_step * steps,
                "model_flop_per_pose": flop,
                "model_tflops": value * flop / 1e12 / world,
                "roofline": roof,
                "roofline_hbm": hbm_rooflines(cabi, dev, pk, batch) if (rank == 0 and workload != "eval") else None}

    line = measure(args.workload, args.steps, args.warmup, True)
    profiling = bool(os.environ.get("LIFTER_BENCH_NO_CPU"))     # under ncu: skip the untimed extras
    if args.workload == "train" and world == 1 and not profiling:
        # the same lifter in eval mode (BASELINE.json configs[1]), reported beside the headline training number
        inf = measure("eval", args.steps, args.warmup, False)
        line["inference"] = {k: inf[k] for k in ("metric", "value", "unit", "ms_per_step", "dtype", "e2e", "roofline",
                                                 "model_tflops")}
        line["inference"]["workload"] = inf["config"]["workload"]
    if rank == 0:
        line["cpu_baseline"] = None if profiling else cpu_arm(args.workload, steps=8, warmup=1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def ncu_traffic(rows, width):
    """DRAM bytes of one launch of the dominant kernel from the committed `ncu --set full` capture (profiles/), which
    was taken on the training bench (M=8192, N=K=1024); null for any other shape."""
    path = os.path.join(ROOT, "profiles", "r01_train_gemm_ncu_full_raw_v6.csv")
    if rows != TRAIN_BATCH or width != H or not os.path.exists(path):
        return {"traffic": None}
    import csv
    with open(path, newline="") as f:
        rows_ = list(csv.reader(f))
    head, units, vals = rows_[0], rows_[1], rows_[2]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total = 0.0
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = head.index(name)
        total += float(vals[i]) * scale[units[i]]
    return {"traffic": total, "traffic_unit": "bytes/launch",
            "traffic_source": "profiles/r01_train_gemm_ncu_full_raw_v6.csv (dram read + write; the 16 MB output was "
                              "still in L2 when the kernel ended)",
            "algorithmic_bytes": float(rows * width * 2 * 2 + width * width * 2)}


def gemm_roofline(cabi, dev, rows, fmt, pk, H=H):
    """CUDA-event time of the dominant kernel (rows x 1024 x 1024 tcgen05 GEMM with the fused epilogue)."""
    import ctypes as C

    import torch
    dt = torch.float16 if fmt == cabi.FMT_F16 else torch.bfloat16
    nbuf = 6 if rows <= 16384 else 2                   # rotate operands so A is not L2-warm from the previous launch
    As = [torch.randn(rows, H, device=dev).to(dt) for _ in range(nbuf)]
    W = (torch.randn(H, H, device=dev) * 0.03).to(dt)
    D = torch.empty(rows, H, device=dev, dtype=dt)
    scale, shift = torch.ones(H, device=dev), torch.zeros(H, device=dev)
    stream = cabi.stream_ptr()

    def launch(i):
        d = cabi.GemmDesc()
        d.A, d.lda, d.B, d.ldb = As[i % nbuf].data_ptr(), H, W.data_ptr(), H
        d.M, d.N, d.K, d.fmt, d.out_mode, d.D, d.ldd = rows, H, H, fmt, cabi.OUT_16, D.data_ptr(), H
        d.splits, d.relu, d.scale, d.shift = 1, 1, scale.data_ptr(), shift.data_ptr()
        cabi.check(cabi.lib().lifter_gemm(C.byref(d), stream))
    for i in range(5):
        launch(i)
    torch.cuda.synchronize()
    n = 30
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        launch(i)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    flops = 2.0 * rows * H * H
    achieved = flops / (us * 1e-6) / 1e12
    return {"bound": "tensor", "kernel": f"lifter_gemm_kernel<256, 2-SM> M={rows} N=K={H} (+affine+ReLU epilogue)",
            "achieved": achieved, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": achieved / pk["bf16"],
            "peak_source": pk["source"] + " burst bf16 (kernel timed alone)", "us_per_launch": us, **ncu_traffic(rows, H)}


def hbm_rooflines(cabi, dev, pk, batch):
    """The HBM-bound kernels of the step timed alone with CUDA events: fused Adam over the flat buffer
    (28 B/param + 2 B operand copy) and the fused loss + gradient + metric kernel (612 B/pose)."""
    import torch
    lib = cabi.lib()
    n = 4_296_768                                     # flat buffer of the 1024-wide, 2-block lifter (tensors padded to 32)
    # three full sets of (p, g, m, v) = 206 MB, rotated: no launch finds its operands in the 126 MB L2
    sets = [[torch.randn(n, device=dev) * 0.01 for _ in range(3)] + [torch.rand(n, device=dev) * 1e-4] for _ in range(3)]
    st = cabi.stream_ptr()
    tick = [0]

    def time_us(fn, reps=30):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps * 1e3
    out = []
    def adam():
        p, g, m, v = sets[tick[0] % 3]
        tick[0] += 1
        cabi.check(lib.lifter_adam_step(cabi.ptr(p), cabi.ptr(g), cabi.ptr(m), cabi.ptr(v), n, 1e-4, 0.9, 0.999, 1e-8,
                                        0.0, 0, 1, 1.0, st))
    us = time_us(adam)
    gbs = 28.0 * n / (us * 1e-6) / 1e9
    out.append({"bound": "hbm", "kernel": f"adam_kernel over {n} fp32 parameters (28 B/param = 120 MB per launch; "
                "operand sets rotated past the L2 capacity)", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                "us_per_launch": us, "traffic": None})
    nrot = 40                                           # 40 x 5 MB of (pred, target, dpred): past the L2 capacity too
    preds = [torch.randn(batch, O_DIM, device=dev) for _ in range(nrot)]
    tgts = [torch.randn(batch, O_DIM, device=dev) for _ in range(nrot)]
    dpreds = [torch.empty(batch, O_DIM, device=dev) for _ in range(nrot)]
    scratch = torch.empty(cabi.check(lib.lifter_loss_scratch_bytes(batch, 17)), dtype=torch.uint8, device=dev)
    res = torch.zeros(18, device=dev)
    def loss():
        i = tick[0] % nrot
        tick[0] += 1
        cabi.check(lib.lifter_loss(cabi.ptr(preds[i]), cabi.ptr(tgts[i]), batch, 17, cabi.LOSS_MSE, batch, None,
                                   cabi.ptr(dpreds[i]), cabi.ptr(scratch), cabi.ptr(res), 0, st))
    us = time_us(loss)
    gbs = 612.0 * batch / (us * 1e-6) / 1e9
    out.append({"bound": "hbm", "kernel": f"fused_loss_kernel + loss_finalize_kernel, {batch} poses (612 B/pose = "
                f"{612 * batch / 1e6:.1f} MB: launch-latency-bound at this size)", "achieved": gbs, "peak": pk["hbm"],
                "unit": "GB/s", "frac": gbs / pk["hbm"], "us_per_launch": us, "traffic": None})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("LIFTER_BENCH_WORKLOAD", "train"), choices=["train", "eval", "wide"])
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_arm(args.workload, steps=args.steps, warmup=args.warmup, budget_s=120.0)
        line = {"impl": "reference", "metric": "lifted poses/sec" + (" (eval inference)" if args.workload == "eval" else " (fwd+bwd+Adam)"),
                "value": r["value"], "unit": "poses/s", "n_gpus": args.gpus, "steps": r["steps_timed"], "warmup": args.warmup,
                "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "fp32", "data": "synthetic (same generator as the GPU arm)",
                "config": {"workload": {"eval": "configs[1] eval inference", "train": "configs[2] training step",
                                        "wide": "configs[4] widened-lifter training step"}[args.workload] +
                           "; CPU restatement of the reference module on the host cores, bounded sample: " + r["sample"]},
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": r["value"], "unit": "poses/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return
    gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
