#!/usr/bin/env python
"""Benchmark of the B200-native lifter hot path (contract: see the task statement / DESIGN.md §7).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload train|eval|wide|phase5|vit|vit_eval] [--impl ours|reference]

The default workload is BASELINE.json's metric, "lifted poses/sec (fwd+bwd)": the data-parallel training step of
configs[2] (8,192 poses per GPU); with one GPU the line also carries an ``inference`` object with the eval-mode
numbers of configs[1] (65,536 poses).  ``--workload eval`` makes configs[1] the headline, ``--workload phase5`` is
configs[3] (the phase-5 loop body around the lifter), ``--workload wide`` configs[4] (2,048 wide, 4 blocks).

* ``value``     : lifted poses/s over all ranks, inputs resident in HBM, CUDA-event time, max over ranks.
* ``e2e``       : the same metric through the public module API with HOST (pinned) inputs: H2D copy of the
                  step's keypoints, the step, D2H of the step's result -- all inside the timed region.
* ``sustained`` : the same step repeated for >= 1.5 s (CUDA-event time, nvidia-smi clock samples INSIDE the region).
* ``roofline``  : the dominant kernel (the H x H tcgen05 GEMM) timed alone with CUDA events, against the measured
                  bf16 peak of MEASURED_PEAKS.json; ``library_point`` is cuBLAS (torch.matmul) at the same shape.
* ``dp_check``  : N > 1 only.  After the timed loops every rank replays 3 steps from a fixed seed, data parallel, and
                  the same 3 steps single-process on the concatenated global batch: losses, parameters and replica
                  identity are compared; a failed check makes the process exit non-zero.
* ``cpu_baseline`` / ``--impl reference``: the UNMODIFIED reference module (staged from /root/reference into
                  oracle/_ref/ by oracle/make_ref.sh) on the host cores, at the GPU arm's configuration.
"""
from __future__ import annotations

import argparse
import hashlib
import importlib
import importlib.util
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
# configs[3]: four lifter forwards per pose (pred / gt keypoints, each plus its flipped twin) and the backward of the
# two that reach the loss, with the gradient into the predicted 2D input (SURVEY §3.3, §8d "25,688,064 if input grad")
FLOP_PER_POSE_PHASE5 = 4 * FLOP_PER_POSE_INFER + 2 * (25_688_064 - FLOP_PER_POSE_INFER)
# my_HybrIK/utils.py:131-172: (fx, fy, cx, cy) in pixels of the four Human3.6M cameras
H36M_INTRINSICS = [(1145.0494384765625, 1143.7811279296875, 512.54150390625, 515.4514770507812),
                   (1149.6756591796875, 1147.5916748046875, 508.8486328125, 508.0649108886719),
                   (1149.1407470703125, 1148.7989501953125, 519.8158569335938, 501.40264892578125),
                   (1145.5113525390625, 1144.77392578125, 514.9682006835938, 501.88201904296875)]
WORKLOADS = {"eval": "configs[1]", "train": "configs[2]", "phase5": "configs[3]", "wide": "configs[4]",
             "vit": "SURVEY 8 f1 (MyViT training)", "vit_eval": "SURVEY 8 f1 (MyViT eval)"}
# MyViT(chw=(1, 17, 2), n_blocks=2, hidden_d=256, n_heads=4, out_d=3), baselineModel.py:307-362: multiply-adds x 2 per pose
VIT_H, VIT_N, VIT_NB = 256, 17, 2
FLOP_PER_POSE_VIT_INFER = 2 * VIT_N * (2 * VIT_H + VIT_NB * (VIT_H * 3 * VIT_H + VIT_H * VIT_H + 2 * VIT_H * 4 * VIT_H)
                                       + VIT_H * (VIT_H // 2) + (VIT_H // 2) * 3) + VIT_NB * 4 * VIT_N * VIT_N * VIT_H
FLOP_PER_POSE_VIT_TRAIN = 3 * FLOP_PER_POSE_VIT_INFER      # dgrad + wgrad ~ 2 x forward (no data gradient into the 2-wide input)
VIT_EVAL_CHUNK = 32768                   # poses per pass of the eval kernel sequence (scripts/vit_bench.py sweep: 10.6 ms at 8,192, 9.6 at 32,768, 9.8 at 65,536)


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

    def summary(self, first=0, last=None):
        samples = self.samples[first:last]
        if not samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"], "samples": 0}
        sm = [float(s[0]) for s in samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in samples for n, v in zip(names, s[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(samples)}


# ------------------------------------------------------------------------------------------------ synthetic data
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


def synth_vit(batch, seed):
    """Key points (B, 17, 2) and targets (B, 17, 3) for the MyViT lifter: the H36M-shaped rows of synth_batch, per joint."""
    x, y = synth_batch(batch, seed)
    return x.reshape(batch, VIT_N, 2).contiguous(), y.reshape(batch, VIT_N, 3).contiguous()


def synth_wide(batch, seed):
    """configs[4]: 16 iid frames concatenated into a 544-wide input."""
    import torch
    x = torch.cat([synth_batch(batch, 100 * f + seed)[0] for f in range(16)], dim=1).contiguous()
    return x, synth_batch(batch, seed)[1]


def synth_phase5(batch, seed):
    """configs[3]: what the phase-5 loop hands the lifter per step (backup_june8.py:128-138): the 2D model's prediction
    ``y1_hat`` and the 3D model's ``y2_hat`` (stand-ins: ground truth + noise, the image CNNs are out of scope), the
    ground truth ``y1`` / ``y2``, and the intrinsics of a camera drawn per pose (my_HybrIK/utils.py:131-172) for the
    parity-unpinned reprojection term."""
    import torch
    y1, y2 = synth_batch(batch, seed)
    g = torch.Generator().manual_seed(seed + 7919)
    y1_hat = y1 + 0.05 * torch.randn(y1.shape, generator=g)
    y2_hat = y2 + 0.10 * torch.randn(y2.shape, generator=g)
    cam = torch.randint(0, 4, (batch,), generator=g)
    intr = torch.tensor(H36M_INTRINSICS, dtype=torch.float32)[cam].contiguous()      # (fx, fy, cx, cy) per pose
    return y1_hat.contiguous(), y2_hat.contiguous(), y1, y2, intr


# ------------------------------------------------------------------------------------------------ CPU arm
def load_reference():
    """The UNMODIFIED reference modules staged by oracle/make_ref.sh (oracle/_ref/, git-ignored, travels with the
    snapshot).  Returns (baselineModel, losses, info) or None when they are not there."""
    d = os.path.join(ROOT, "oracle", "_ref")
    paths = {n: os.path.join(d, n + ".py") for n in ("baselineModel", "losses")}
    if not all(os.path.exists(p) for p in paths.values()):
        return None
    mods, shas = {}, {}
    for n, p in paths.items():
        spec = importlib.util.spec_from_file_location("_reference_" + n, p)
        mods[n] = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mods[n])
        shas[n + ".py"] = hashlib.sha256(open(p, "rb").read()).hexdigest()
    want = {}
    rec = os.path.join(ROOT, "oracle", "ref_sha256.txt")
    if os.path.exists(rec):
        want = {ln.split()[1]: ln.split()[0] for ln in open(rec) if ln.strip()}
    return mods["baselineModel"], mods["losses"], {"sha256": shas, "matches_recorded": all(want.get(k) == v for k, v in shas.items())}


def _time_cpu(step, steps, warmup, budget_s):
    for _ in range(max(1, warmup)):
        step()
    times, t_begin = [], time.perf_counter()
    for _ in range(max(1, steps)):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_begin > budget_s:
            break
    return statistics.median(times) * 1e3, len(times)


def cpu_arm(workload, steps, warmup, budget_s=25.0, with_c1=True):
    """The reference's own CPU implementation of the workload on the host cores (PyTorch fp32, every host thread):
    the imported ``LinearModel`` (baselineModel.py:50-102) driven by the loop body of train_1.py:73-100
    (``MSELoss(mean)``, ``torch.optim.Adam``, lr 1e-4) at the GPU arm's batch.  Falls back to the oracle port (kind
    "port") only when oracle/_ref/ is absent."""
    import warnings

    import torch
    warnings.filterwarnings("ignore")
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    ref = load_reference()
    if ref is None:
        return cpu_arm_port(workload, steps, warmup, budget_s)
    rm, rl, info = ref
    i_dim, h, s = WIDE if workload == "wide" else (I_DIM, H, S)
    torch.manual_seed(1)
    if workload in ("vit", "vit_eval"):
        model = rm.MyViT(chw=(1, VIT_N, 2), n_blocks=VIT_NB, hidden_d=VIT_H, n_heads=4, out_d=3)
    else:
        model = rm.LinearModel(i_dim=i_dim, o_dim=O_DIM, p_dropout=0.5, linear_size=h, num_stage=s, BN=True)
    out = {"unit": "poses/s", "cores": cores, "threads": torch.get_num_threads(), "kind": "reference",
           "reference_files": info, "torch": torch.__version__}
    if workload in ("vit", "vit_eval"):
        # a BOUNDED sample: the full 8,192 / 65,536-pose step of the transformer takes minutes per step on the host cores
        batch = 1024 if workload == "vit" else 4096
        x, y = synth_vit(batch, 0)
        if workload == "vit_eval":
            model.eval()

            def step():
                with torch.no_grad():
                    model(x)
            sample = f"{batch} of the poses per step: the imported MyViT, eval forward under no_grad"
        else:
            model.train()
            opt = torch.optim.Adam(model.parameters(), lr=1e-4)
            loss_fn = torch.nn.MSELoss(reduction="mean")

            def step():
                opt.zero_grad()
                loss_fn(model(x), y).backward()
                opt.step()
            sample = f"{batch} of the poses per step: the imported MyViT, fwd + MSELoss + backward + Adam.step (train_1.py:73-100)"
    elif workload == "eval":
        batch = EVAL_BATCH
        x, _ = synth_batch(batch, 0)
        model.eval()

        def step():
            with torch.no_grad():
                model(x)
        sample = f"full configs[1] step: {batch} poses, eval forward under no_grad (train_1.py:112-118)"
    elif workload == "phase5":
        from oracle import losses_oracle as L          # flip_pose: my_HybrIK/utils.py imports matplotlib (absent)
        batch = TRAIN_BATCH
        y1_hat, y2_hat, y1, y2, _ = synth_phase5(batch, 0)       # (the reference has no reprojection term to feed)
        model.train()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4)
        tri = rl.TriangleLoss()
        B = batch

        def step():
            opt.zero_grad()
            a = y1_hat.reshape(B, 17, 2).clone().requires_grad_(True)
            b3 = y2_hat.reshape(B, 17, 3).clone().requires_grad_(True)
            lp = model(a).reshape(B, 17, 3)
            lg = model(y1.reshape(B, 17, 2)).reshape(B, 17, 3)
            lp = (L.flip_pose(model(L.flip_pose(a)).reshape(B, 17, 3)) + lp) / 2
            lg = (L.flip_pose(model(L.flip_pose(y1.reshape(B, 17, 2))).reshape(B, 17, 3)) + lg) / 2
            loss = tri(a, b3, lg, lp, y1.reshape(B, 17, 2), y2.reshape(B, 17, 3))[0]
            loss.backward()
            opt.step()
        sample = (f"full configs[3] step: {batch} poses, 4 lifter forwards + flip averaging + TriangleLoss + backward "
                  "+ AdamW (backup_june8.py:120-175; flip_pose from the oracle restatement)")
    else:
        batch = TRAIN_BATCH
        x, y = synth_wide(batch, 0) if workload == "wide" else synth_batch(batch, 0)
        model.train()
        opt = torch.optim.Adam(model.parameters(), lr=1e-4)
        loss_fn = torch.nn.MSELoss(reduction="mean")

        def step():
            opt.zero_grad()
            loss = loss_fn(model(x), y)
            loss.backward()
            opt.step()
        sample = f"full {WORKLOADS[workload]} step: {batch} poses, fwd + MSELoss + backward + Adam.step (train_1.py:73-100)"
    ms, n = _time_cpu(step, steps, min(warmup, 2), budget_s)
    out.update({"value": batch / (ms / 1e3), "sample": sample, "ms_per_step": ms, "steps_timed": n, "batch": batch})
    if with_c1 and workload in ("train", "eval"):
        # SURVEY §8d / BASELINE.md §2, C1: batch 64 fwd + MSE + bwd + Adam.step, 5 warm-up + 50 timed, median
        torch.manual_seed(1)
        m1 = rm.LinearModel(i_dim=I_DIM, o_dim=O_DIM, p_dropout=0.5, linear_size=H, BN=True).train()
        o1 = torch.optim.Adam(m1.parameters(), lr=1e-4)
        x1, y1_ = synth_batch(64, 0)
        lf = torch.nn.MSELoss(reduction="mean")

        def c1():
            o1.zero_grad()
            lf(m1(x1), y1_).backward()
            o1.step()
        ms1, n1 = _time_cpu(c1, 50, 5, 20.0)
        out["c1"] = {"config": "configs[0]: batch 64 fwd+MSE+bwd+Adam on CPU, 5 warm-up + 50 timed, median",
                     "ms_per_step": ms1, "value": 64 / (ms1 / 1e3), "unit": "poses/s", "steps_timed": n1}
    return out


def cpu_arm_port(workload, steps, warmup, budget_s=25.0):
    """Fallback when oracle/_ref/ is absent: the oracle's functional restatement of the reference (kind "port")."""
    import torch

    from oracle import adam_oracle, lifter_oracle as lo
    from oracle.philox import dropout_keep_mask
    cores = os.cpu_count() or 1
    if workload in ("vit", "vit_eval"):
        from oracle import vit_oracle as V
        torch.set_num_threads(cores)
        batch = 1024 if workload == "vit" else 4096
        x, y = synth_vit(batch, 0)
        prm = {k: v.clone().requires_grad_(workload == "vit" and k != "pos_embed") for k, v in V.numpy_init(1).items()}
        if workload == "vit":
            opt = torch.optim.Adam([v for v in prm.values() if v.requires_grad], lr=1e-4)

            def step():
                opt.zero_grad()
                torch.nn.functional.mse_loss(V.vit_forward(prm, x), y).backward()
                opt.step()
        else:
            def step():
                with torch.no_grad():
                    V.vit_forward(prm, x)
        ms, n = _time_cpu(step, steps, min(warmup, 2), budget_s)
        return {"value": batch / (ms / 1e3), "unit": "poses/s", "cores": cores, "kind": "port", "ms_per_step": ms, "steps_timed": n,
                "batch": batch, "sample": f"{batch} of the poses per step, oracle/vit_oracle.py (PyTorch fp32 CPU)"}
    i_dim, h, s = WIDE if workload == "wide" else (I_DIM, H, S)
    params, buffers = lo.numpy_init(i_dim, O_DIM, h, s, seed=1, randomize_bn=False)
    if workload == "eval":
        batch = EVAL_BATCH
        x, _ = synth_batch(batch, 0)

        def step():
            with torch.no_grad():
                lo.lifter_forward(params, buffers, x, num_stage=s, training=False)
        sample = f"full configs[1] step: {batch} poses, eval forward"
    else:
        batch = TRAIN_BATCH
        x, y = synth_wide(batch, 0) if workload == "wide" else synth_batch(batch, 0)
        masks = [dropout_keep_mask(batch, h, k, 0, 2, 0.5) for k in range(2 * s + 1)]
        m = {k: torch.zeros_like(v) for k, v in params.items()}
        v = {k: torch.zeros_like(w) for k, w in params.items()}
        state = {"t": 0}

        def step():
            r = lo.lifter_forward_backward(params, buffers, x, lambda o: ((o - y) ** 2).mean(), num_stage=s,
                                           training=True, p_dropout=0.5, masks=masks)
            state["t"] += 1
            for k in params:
                params[k], m[k], v[k] = adam_oracle.adam_step(params[k], r["grads"][k], m[k], v[k], state["t"], 1e-4)
        sample = f"full {WORKLOADS.get(workload, workload)} step: {batch} poses, fwd+MSE+bwd+Adam (oracle port; phase5 runs the plain step)"
    ms, n = _time_cpu(step, steps, min(warmup, 2), budget_s)
    return {"value": batch / (ms / 1e3), "unit": "poses/s", "cores": cores, "kind": "port", "sample": sample,
            "ms_per_step": ms, "steps_timed": n, "batch": batch}


# ------------------------------------------------------------------------------------------------ GPU arm
def gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    pkg = importlib.import_module(PKG)
    cabi = importlib.import_module(PKG + "._cabi")
    trainer_mod = importlib.import_module(PKG + ".trainer")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cabi.check(cabi.lib().lifter_device_check())
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk = peaks()
    profiling = bool(os.environ.get("LIFTER_BENCH_NO_CPU"))     # under ncu: skip the untimed extras

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world > 1:
            t = torch.tensor([v], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return v

    def build_model(workload):
        torch.manual_seed(1)
        if workload in ("vit", "vit_eval"):
            return pkg.MyViT(chw=(1, VIT_N, 2), n_blocks=VIT_NB, hidden_d=VIT_H, n_heads=4, out_d=3).to(dev)
        if workload == "wide":
            return pkg.LinearModel(WIDE[0], O_DIM, linear_size=WIDE[1], num_stage=WIDE[2], p_dropout=0.5, BN=True).to(dev)
        return pkg.LinearModel(I_DIM, O_DIM, linear_size=H, num_stage=S, p_dropout=0.5, BN=True).to(dev)

    def build_trainer(workload, model, lr=1e-4, dp=True):
        if workload == "vit":             # no BatchNorm: the fused reduce-scatter / Adam / all-gather launch is the only exchange
            kw = dict(world_size=world, rank=rank) if (dp and world > 1) else {}
            return importlib.import_module(PKG + ".vit_trainer").VitTrainer(model.train(), lr=lr, optimizer="adam", **kw)
        kw = dict(world_size=world, rank=rank, require_symmetric=True) if (dp and world > 1) else {}
        if workload == "phase5":
            import numpy as np
            st = dict(np.load(os.path.join(ROOT, "tests", "golden", "h36m_stats.npz")))
            m3 = torch.from_numpy(st["mean_train_3d"]).float().clone()
            root = m3[0].clone()
            m3[0] = 0                         # zero_centre (H36_dataset.py:288-289): the root's mean is the absolute root
            return trainer_mod.Phase5Trainer(model, lr=lr, loss="triangle", optimizer="adamw", seed=2, w_reproj=0.1,
                                             w_bone=0.1, stats3d=(m3, torch.from_numpy(st["std_train_3d"]).float()),
                                             stats2d=(torch.from_numpy(st["mean_train_2d"]).float(),
                                                      torch.from_numpy(st["std_train_2d"]).float()), root=root, **kw)
        return trainer_mod.LifterTrainer(model, lr=lr, loss="mse", optimizer="adam", seed=2, **kw)

    def host_batch(workload, batch, seed):
        if workload in ("vit", "vit_eval"):
            return synth_vit(batch, seed)
        if workload == "phase5":
            return synth_phase5(batch, seed)
        return synth_wide(batch, seed) if workload == "wide" else synth_batch(batch, seed)

    def measure(workload, steps, warmup, with_clocks):
        """One workload on this rank's GPU: device-resident throughput, end-to-end throughput, dominant kernel."""
        model = build_model(workload)
        tr = None
        if workload == "vit_eval":
            batch = EVAL_BATCH
            model.eval()
            model.chunk_poses = VIT_EVAL_CHUNK
            xh = synth_vit(batch, rank)[0].pin_memory()
            xd = xh.to(dev)
            xs = torch.empty_like(xd)
            outh = torch.empty(batch, VIT_N, 3).pin_memory()

            def step():
                model.infer(xd)

            def run_e2e(n):                                  # host key points in, host predictions out, every step
                for _ in range(n):
                    xs.copy_(xh, non_blocking=True)
                    outh.copy_(model.infer(xs), non_blocking=True)
                torch.cuda.current_stream().synchronize()
            h2d, d2h = xh.numel() * 4, outh.numel() * 4
            dtype = "fp16 GEMM operands, fp32 residual stream / accumulate / LayerNorm / softmax"
            cfg = {"workload": f"SURVEY 8 f1: MyViT(chw=(1,17,2), n_blocks=2, hidden_d=256, n_heads=4) eval forward, batch {batch} "
                               f"synthetic H36M 2D poses per GPU, {VIT_EVAL_CHUNK} poses per pass of the kernel sequence",
                   "batch_per_gpu": batch, "parallelism": f"replicas x{world}" if world > 1 else "single GPU",
                   "l2": "one pass streams > 1 GB of activations (> 126 MB L2); no flush needed"}
            flop = FLOP_PER_POSE_VIT_INFER
        elif workload == "eval":
            batch = EVAL_BATCH
            model.eval()
            xh, _ = synth_batch(batch, rank)
            xh = xh.pin_memory()
            xd = xh.to(dev)
            out = torch.empty(batch, O_DIM, device=dev)
            outh = torch.empty(batch, O_DIM).pin_memory()

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
            batch = TRAIN_BATCH
            tr = build_trainer(workload, model)
            host = [t.pin_memory() for t in host_batch(workload, batch, rank)]
            devt = [t.to(dev) for t in host]

            def step():
                tr.step(*devt)

            def run_e2e(n):
                # the epoch loop over host batches: per step, an H2D copy of the inputs and a D2H read of the loss
                losses = tr.run_epoch([tuple(host)] * n)
                assert losses.numel() == n and bool(torch.isfinite(losses).all())
            h2d, d2h = sum(t.numel() * t.element_size() for t in host), 4
            dtype = "bf16 operands, fp32 accumulate/master weights"
            par = f"dp{world}"
            if workload == "vit":
                par = (f"dp{world} (fused reduce-scatter/Adam/all-gather over NVLink symmetric memory; no BatchNorm, so no other "
                       "exchange)") if world > 1 else "single GPU"
            elif world > 1:
                par += " (in-kernel SyncBN exchange + fused reduce-scatter/Adam/all-gather over NVLink symmetric memory)"
            what = {"train": "training step fwd + MSE + bwd + Adam, dropout 0.5, batch {b}/GPU, synced BatchNorm, 1024-wide, "
                             "2 residual blocks",
                    "wide": "training step fwd + MSE + bwd + Adam, dropout 0.5, batch {b}/GPU, synced BatchNorm, widened lifter "
                            "(16 x 34 inputs, 2048-wide, 4 residual blocks)",
                    "phase5": "phase-5 loop body around the lifter (backup_june8.py:120-175), {b} poses/GPU: 4 train-mode lifter "
                              "forwards (predicted / ground-truth 2D, each with its flipped twin), flip averaging, TriangleLoss "
                              "+ pinhole-reprojection and bone-length terms (parity unpinned) on synthetic H36M intrinsics, "
                              "backward into the lifter and the predicted 2D / 3D inputs, AdamW",
                    "vit": "MyViT(chw=(1,17,2), n_blocks=2, hidden_d=256, n_heads=4) training step fwd + MSE + bwd + Adam, {b} "
                           "poses/GPU, VitTrainer (flat parameters, one optimiser launch, CUDA-graph replay)"}[workload].format(b=batch)
            cfg = {"workload": f"{WORKLOADS[workload]}: {what}", "batch_per_gpu": batch, "global_batch": batch * world,
                   "parallelism": par, "cuda_graph": bool(tr.use_graph),
                   "l2": "one step touches >= 0.4 GB of saved activations and gradients per GPU (> 126 MB L2), so "
                         "consecutive steps do not find their inputs in L2; no explicit flush"}
            flop = {"train": FLOP_PER_POSE_TRAIN, "wide": FLOP_PER_POSE_WIDE_TRAIN, "phase5": FLOP_PER_POSE_PHASE5,
                    "vit": FLOP_PER_POSE_VIT_TRAIN}[workload]
            if workload == "vit":
                dtype = "bf16 operands / activations / stored gradients, fp32 accumulate, fp32 parameters and parameter gradients"

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
            return max_over_ranks(e0.elapsed_time(e1) / steps)

        step()                                                 # builds buffers / captures the step graph untimed
        barrier()
        sustained = None
        if with_clocks:
            with ClockSampler(local_rank) as clk:
                t_wait = time.perf_counter()                    # the first nvidia-smi query on a fresh box can take > 1 s:
                while not clk.samples and time.perf_counter() - t_wait < 8.0:     # let it answer once before timing
                    time.sleep(0.05)
                barrier()
                start = len(clk.samples)
                ms = timed(step, steps, warmup)
                in_region = len(clk.samples) - start
                # nvidia-smi answers every ~0.1 s, so a 20-step region sees no sample.  The same step therefore runs on
                # for >= 2 s as a second, separately timed region ("sustained").  Its step count derives from the
                # max-over-ranks step time, so every rank runs the same number of (exchanging) steps.
                n_sus = max(steps, int(2000.0 / max(ms, 1e-3)))
                first = len(clk.samples)
                ms_sus = timed(step, n_sus, 0)
                last = len(clk.samples)
            clocks = clk.summary(start)                         # samples taken under load only
            clocks["samples_in_timed_region"] = in_region
            sc = clk.summary(first, last)
            sc["samples_in_timed_region"] = sc.pop("samples")
            sustained = {"ms_per_step": ms_sus, "steps": n_sus, "seconds": ms_sus * n_sus / 1e3,
                         "value": batch * world / (ms_sus / 1e3), "unit": "poses/s", "clocks": sc}
        else:
            ms, clocks = timed(step, steps, warmup), None

        # end to end through the public API with host buffers (wall clock around synchronising steps)
        run_e2e(max(8, warmup))                                # also past the point where run_epoch's slots get their graphs
        barrier()
        t0 = time.perf_counter()
        n_e2e = max(3, steps // 2)
        run_e2e(n_e2e)
        barrier()
        e2e_ms = max_over_ranks((time.perf_counter() - t0) / n_e2e * 1e3)

        roof = None
        if rank == 0:   # dominant kernel alone: the H x H tcgen05 GEMM at this workload's per-launch shape
            if workload in ("vit", "vit_eval"):
                roof = vit_gemm_roofline(cabi, dev, (VIT_EVAL_CHUNK if workload == "vit_eval" else batch) * VIT_N,
                                         cabi.FMT_F16 if workload == "vit_eval" else cabi.FMT_BF16, pk)
            else:
                roof = gemm_roofline(cabi, dev, batch, cabi.FMT_F16 if workload == "eval" else cabi.FMT_BF16, pk,
                                     WIDE[1] if workload == "wide" else H, profiling)
        if workload == "vit_eval":
            n0 = cabi.lib().lifter_launch_count()
            model.infer(xd)
            launches = int(cabi.lib().lifter_launch_count() - n0)
        else:
            launches = tr.graph_launches() if tr is not None else model.infer_launches(batch)
        value = batch * world / (ms / 1e3)
        return {"metric": "lifted poses/sec" + (" (eval inference)" if workload in ("eval", "vit_eval") else " (fwd+bwd+Adam)"),
                "value": value, "unit": "poses/s", "n_gpus": world, "steps": steps, "warmup": warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": dtype,
                "data": "synthetic (H36M-shaped keypoints from the repo's mean/std statistics, random-init weights)",
                "config": cfg, "clocks": clocks, "sustained": sustained,
                "e2e": {"value": batch * world / (e2e_ms / 1e3), "unit": "poses/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms},
                "gpu_launches": launches * steps, "gpu_launches_per_step": launches,
                "model_flop_per_pose": flop,
                "model_tflops": value * flop / 1e12 / world,
                "step_roofline_frac": {"burst": value * flop / 1e12 / world / pk["bf16"],
                                       "sustained": value * flop / 1e12 / world / pk["bf16_sustained"]},
                "roofline": roof,
                "roofline_hbm": hbm_rooflines(cabi, dev, pk, batch) if (rank == 0 and workload == "train" and not profiling) else None}

    def dp_check(workload, k=3):
        """N data-parallel ranks == one process on the concatenated global batch (SURVEY §8c multi-GPU oracle), on the
        exact path that was just timed (graph replay, in-kernel SyncBN exchange, per-layer reduce-scatter/Adam/all-gather
        on the side stream)."""
        batch, lr = TRAIN_BATCH, 1e-4                                     # the learning rate of the timed run
        glob = [[torch.cat([t for t in col]) for col in zip(*[host_batch(workload, batch, 1000 * s + r) for r in range(world)])]
                for s in range(k)]
        # data parallel
        m = build_model(workload)
        p0 = m._flat.clone()
        tr = build_trainer(workload, m, lr=lr)
        losses = []
        for s in range(k):
            tr.step(*[t[rank * batch:(rank + 1) * batch].to(dev) for t in glob[s]])
            losses.append(tr.loss_value())
        torch.cuda.synchronize()
        dp_flat, dp_rm = m._flat.clone(), m._rm.clone()
        del tr
        # single process, global batch (every rank computes it: same answer everywhere, nothing to ship)
        ref = build_model(workload)
        tr_ref = build_trainer(workload, ref, lr=lr, dp=False)
        ref_losses = []
        for s in range(k):
            tr_ref.step(*[t.to(dev) for t in glob[s]])
            ref_losses.append(tr_ref.loss_value())
        torch.cuda.synchronize()
        loss_err = max(abs(a - b) / abs(b) for a, b in zip(losses, ref_losses))
        pd = ((dp_flat - ref._flat).norm() / ref._flat.norm()).item()
        # the UPDATE the k steps made, not the parameters (Adam at lr 1e-4 moves a weight by 3e-4 in three steps: any two
        # runs agree on the parameters).  Adam's first steps are sign-like, so a gradient entry within bf16 noise of zero
        # flips its whole step when the summation order changes (N shards vs one batch): measured 0.03-0.12; a rank's
        # gradient missing from the exchange, or local instead of global BatchNorm statistics, gives >= 0.5
        ud = ((dp_flat - ref._flat).norm() / (ref._flat - p0).norm()).item()
        rmd = ((dp_rm - ref._rm).abs().max() / ref._rm.abs().max()).item()
        other = dp_flat.clone()
        dist.broadcast(other, src=0)
        same = torch.tensor([1.0 if torch.equal(other, dp_flat) else 0.0], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        vals = torch.tensor([loss_err, pd, ud, rmd], device=dev, dtype=torch.float64)
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        loss_err, pd, ud, rmd = [float(v) for v in vals]
        tol = {"loss_rel_err": 1e-3, "param_rel_diff": 1e-3, "update_rel_diff": 0.2, "running_mean_rel_diff": 5e-3}
        ok = loss_err < tol["loss_rel_err"] and pd < tol["param_rel_diff"] and ud < tol["update_rel_diff"] and \
            rmd < tol["running_mean_rel_diff"] and bool(same.item())
        return {"world": world, "steps": k, "batch_per_gpu": batch, "lr": lr, "loss_rel_err": loss_err, "param_rel_diff": pd,
                "update_rel_diff": ud, "running_mean_rel_diff": rmd, "replicas_identical": bool(same.item()),
                "tolerance": tol, "ok": ok, "losses_dp": losses, "losses_single_process": ref_losses}

    def dp_check_vit(k=3):
        """The same check for the transformer lifter: N data-parallel VitTrainer ranks == one process on the concatenated batch."""
        batch, lr = TRAIN_BATCH, 1e-4
        glob = [[torch.cat(col) for col in zip(*[host_batch("vit", batch, 1000 * s + r) for r in range(world)])] for s in range(k)]
        m = build_model("vit")
        tr = build_trainer("vit", m, lr=lr)
        p0 = tr.flat.clone()
        losses = []
        for s in range(k):
            tr.step(*[t[rank * batch:(rank + 1) * batch].to(dev) for t in glob[s]])
            losses.append(tr.loss_value())
        torch.cuda.synchronize()
        dp_flat = tr.flat.clone()
        ref = build_model("vit")
        tr_ref = build_trainer("vit", ref, lr=lr, dp=False)
        ref_losses = []
        for s in range(k):
            tr_ref.step(*[t.to(dev) for t in glob[s]])
            ref_losses.append(tr_ref.loss_value())
        torch.cuda.synchronize()
        loss_err = max(abs(a - b) / abs(b) for a, b in zip(losses, ref_losses))
        pd = ((dp_flat - tr_ref.flat).norm() / tr_ref.flat.norm()).item()
        ud = ((dp_flat - tr_ref.flat).norm() / (tr_ref.flat - p0).norm()).item()
        other = dp_flat.clone()
        dist.broadcast(other, src=0)
        same = torch.tensor([1.0 if torch.equal(other, dp_flat) else 0.0], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        vals = torch.tensor([loss_err, pd, ud], device=dev, dtype=torch.float64)
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        loss_err, pd, ud = [float(v) for v in vals]
        tol = {"loss_rel_err": 1e-3, "param_rel_diff": 1e-3, "update_rel_diff": 0.2}
        ok = loss_err < tol["loss_rel_err"] and pd < tol["param_rel_diff"] and ud < tol["update_rel_diff"] and bool(same.item())
        return {"world": world, "steps": k, "batch_per_gpu": batch, "lr": lr, "loss_rel_err": loss_err, "param_rel_diff": pd,
                "update_rel_diff": ud, "replicas_identical": bool(same.item()), "tolerance": tol, "ok": ok,
                "losses_dp": losses, "losses_single_process": ref_losses}

    line = measure(args.workload, args.steps, args.warmup, True)
    if args.workload == "train" and world == 1 and not profiling:
        # the same lifter in eval mode (BASELINE.json configs[1]), reported beside the headline training number
        inf = measure("eval", args.steps, args.warmup, False)
        line["inference"] = {k: inf[k] for k in ("metric", "value", "unit", "ms_per_step", "dtype", "e2e", "roofline",
                                                 "model_tflops", "gpu_launches_per_step")}
        line["inference"]["workload"] = inf["config"]["workload"]
    ok = True
    if world > 1 and args.workload == "vit" and not profiling:
        line["dp_check"] = dp_check_vit()
        ok = line["dp_check"]["ok"]
    elif world > 1 and args.workload not in ("eval", "vit_eval") and not profiling:
        line["dp_check"] = dp_check(args.workload)
        ok = line["dp_check"]["ok"]
    if rank == 0:
        line["cpu_baseline"] = None if (profiling or world > 1) else cpu_arm(args.workload, steps=4, warmup=1, budget_s=20.0)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        sys.exit(1)


def ncu_traffic(rows, width):
    """DRAM bytes of one launch of the dominant kernel from the committed `ncu --set full` capture (profiles/), which
    was taken on the training bench (M=8192, N=K=1024) or the eval bench (M=65536); null for any other shape."""
    name = {TRAIN_BATCH: "train", EVAL_BATCH: "eval"}.get(rows)
    cands = [os.path.join(ROOT, "profiles", f"r02_{name}_gemm_ncu_full_raw.csv"),
             os.path.join(ROOT, "profiles", "r01_train_gemm_ncu_full_raw_v6.csv") if name == "train" else ""]
    path = next((p for p in cands if p and os.path.exists(p)), None)
    if name is None or width != H or path is None:
        return {"traffic": None}
    import csv
    with open(path, newline="") as f:
        rows_ = list(csv.reader(f))
    head, units, vals = rows_[0], rows_[1], rows_[2]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total = 0.0
    for nm in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = head.index(nm)
        total += float(vals[i]) * scale[units[i]]
    return {"traffic": total, "traffic_unit": "bytes/launch",
            "traffic_source": os.path.relpath(path, ROOT) + " (dram read + write of one launch)",
            "algorithmic_bytes": float(rows * width * 2 * 2 + width * width * 2)}


def gemm_roofline(cabi, dev, rows, fmt, pk, H=H, profiling=False):
    """CUDA-event time of the dominant kernel (rows x H x H tcgen05 GEMM with the fused epilogue), and cuBLAS
    (torch.matmul, library reference point -- not on any product path) at the same shape, timed the same way."""
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

    def time_us(fn, n=30):
        for i in range(5):
            fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3
    us = time_us(launch)
    flops = 2.0 * rows * H * H
    achieved = flops / (us * 1e-6) / 1e12
    out = {"bound": "tensor", "kernel": f"lifter_gemm_kernel<256, 2-SM> M={rows} N=K={H} (+affine+ReLU epilogue)",
           "achieved": achieved, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": achieved / pk["bf16"],
           "peak_source": pk["source"] + " burst bf16 (kernel timed alone)", "us_per_launch": us, **ncu_traffic(rows, H)}
    if not profiling:
        Wt = W.t()
        us_lib = time_us(lambda i: torch.matmul(As[i % nbuf], Wt, out=D))
        out["library_point"] = {"what": "cuBLAS via torch.matmul at the same shape and operand type, no epilogue; context "
                                        "only, not on any product path", "us_per_launch": us_lib,
                                "achieved": flops / (us_lib * 1e-6) / 1e12, "unit": "TFLOP/s"}
    return out


def vit_gemm_roofline(cabi, dev, T, fmt, pk):
    """The dominant GEMM of the transformer lifter alone: the MLP-up projection with bias + GELU in the epilogue,
    u [T][256] x W1 [1024][256]^T -> m [T][1024] (baselineModel.py:294-296), operands rotated past L2."""
    import ctypes as C

    import torch
    Hd, sets = VIT_H, 4
    dt = torch.float16 if fmt == cabi.FMT_F16 else torch.bfloat16
    A = [torch.randn(T, Hd, device=dev).to(dt) for _ in range(sets)]
    W = (torch.randn(4 * Hd, Hd, device=dev) / 16).to(dt)
    bias = torch.randn(4 * Hd, device=dev)
    D = [torch.empty(T, 4 * Hd, device=dev, dtype=dt) for _ in range(sets)]
    descs = []
    for i in range(sets):
        d = cabi.GemmDesc()
        d.A, d.lda, d.B, d.ldb = A[i].data_ptr(), Hd, W.data_ptr(), Hd
        d.M, d.N, d.K, d.fmt, d.out_mode, d.D, d.ldd, d.relu, d.splits = T, 4 * Hd, Hd, fmt, cabi.OUT_16, D[i].data_ptr(), 4 * Hd, 2, 1
        d.shift = bias.data_ptr()
        descs.append(d)
    lib, st = cabi.lib(), cabi.stream_ptr()
    for i in range(6):
        cabi.check(lib.lifter_gemm(C.byref(descs[i % sets]), st))
    torch.cuda.synchronize()
    reps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        cabi.check(lib.lifter_gemm(C.byref(descs[i % sets]), st))
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    tflop = 2.0 * T * 4 * Hd * Hd / 1e12
    return {"bound": "tensor", "kernel": f"lifter_gemm_kernel<256, 2-SM, GELU> M={T} N=1024 K=256 (+bias+GELU epilogue)",
            "achieved": tflop / (us / 1e6), "peak": pk["bf16"], "unit": "TFLOP/s", "frac": tflop / (us / 1e6) / pk["bf16"],
            "us_per_launch": us, "traffic": None, "algorithmic_bytes": 2.0 * (T * Hd + 4 * Hd * Hd + T * 4 * Hd),
            "peak_source": pk.get("source"),
            "note": "K = 256: four K-blocks per tile, so the epilogue (bias + GELU + 16-bit store), not the MMA, bounds this launch; "
                    "the step as a whole is a mix of such GEMMs and HBM-bound row kernels (profiles/README.md)"}


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
    out.append({"bound": "hbm", "kernel": f"fused loss + gradient + metric, {batch} poses (612 B/pose = "
                f"{612 * batch / 1e6:.1f} MB: launch-latency-bound at this size)", "achieved": gbs, "peak": pk["hbm"],
                "unit": "GB/s", "frac": gbs / pk["hbm"], "us_per_launch": us, "traffic": None})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("LIFTER_BENCH_WORKLOAD", "train"), choices=sorted(WORKLOADS))
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_arm(args.workload, steps=args.steps, warmup=args.warmup, budget_s=100.0)
        what = {"eval": "configs[1] eval inference, batch 65536", "train": "configs[2] training step, batch 8192",
                "wide": "configs[4] widened-lifter training step, batch 8192",
                "phase5": "configs[3] phase-5 loop body around the lifter, batch 8192",
                "vit": "SURVEY 8 f1 MyViT training step (GPU arm: batch 8192)",
                "vit_eval": "SURVEY 8 f1 MyViT eval inference (GPU arm: batch 65536)"}[args.workload]
        line = {"impl": "reference", "metric": "lifted poses/sec" + (" (eval inference)" if args.workload in ("eval", "vit_eval") else " (fwd+bwd+Adam)"),
                "value": r["value"], "unit": "poses/s", "n_gpus": args.gpus, "steps": r["steps_timed"], "warmup": args.warmup,
                "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "fp32", "data": "synthetic (same generator as the GPU arm)",
                "config": {"workload": what + "; the reference's own PyTorch CPU path on the host cores (one process: the "
                                              "reference has no data parallelism, n_gpus is the arm it is compared with): " + r["sample"],
                           "batch_per_gpu": r["batch"]},
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample") if k in r},
                "e2e": {"value": r["value"], "unit": "poses/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        for k in ("c1", "reference_files", "threads", "torch"):
            if k in r:
                line[k] = r[k]
        print(json.dumps(line), flush=True)
        return
    gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
