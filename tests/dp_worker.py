"""torchrun worker: data-parallel training steps (in-kernel SyncBN over symmetric memory + gradient all-reduce)
must reproduce the single-process run on the concatenated global batch.  Launched by tests/test_dp_gpu.py."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def vit_case(pkg, rank, world, dev):
    """The transformer lifter: N data-parallel VitTrainer ranks (fused NVLink reduce-scatter / Adam / all-gather as the only
    exchange) == one process on the concatenated batch; replicas start from different seeds and end bit-identical; the
    batch size changes mid-run (eager step + a second captured graph) and so does the learning rate."""
    VT = importlib.import_module("3d_poseestimation_b200.vit_trainer").VitTrainer
    use_graph = os.environ.get("DP_USE_GRAPH", "1") != "0"
    sizes = [256, 256, 256, 192, 192, 256, 256, 256]
    rs = np.random.RandomState(0)
    xs = [torch.from_numpy(rs.normal(0, 1, (world * B, 17, 2)).astype(np.float32)) for B in sizes]
    ys = [torch.tanh(x.sum(-1, keepdim=True) * torch.tensor([0.3, -0.2, 0.5])) for x in xs]

    def build(seed):
        torch.manual_seed(seed)
        return pkg.MyViT().to(dev).train()
    ref = build(3)
    tr_ref = VT(ref, lr=3e-4, optimizer="adamw", use_graph=use_graph)
    p0 = tr_ref.flat.clone()
    ref_losses = []
    for s, B in enumerate(sizes):
        if s == 5:
            tr_ref.set_lr(1e-4)
        ref_losses.append(float(tr_ref.step(xs[s].to(dev), ys[s].to(dev))))
    m = build(3 if rank == 0 else 100 + rank)                         # the trainer must broadcast rank 0's parameters
    tr = VT(m, lr=3e-4, optimizer="adamw", use_graph=use_graph, world_size=world, rank=rank)
    losses = []
    for s, B in enumerate(sizes):
        if s == 5:
            tr.set_lr(1e-4)
        tr.step(xs[s][rank * B:(rank + 1) * B].to(dev), ys[s][rank * B:(rank + 1) * B].to(dev))
        losses.append(tr.loss_value())
    torch.cuda.synchronize()
    err = max(abs(a - b) / abs(b) for a, b in zip(losses, ref_losses))
    ud = (tr.flat - tr_ref.flat).norm().item() / (tr_ref.flat - p0).norm().item()
    mine = tr.flat.clone()
    other = mine.clone()
    dist.broadcast(other, src=0)
    same = bool(torch.equal(mine, other))
    # the module's parameters are the symmetric buffer: the eval forward sees the exchanged weights (and their operand copies)
    xe = xs[0][:33].to(dev)
    with torch.no_grad():
        oe, re_ = m.eval()(xe), ref.eval()(xe)
    ev = ((oe - re_).norm() / re_.norm()).item()
    print(f"[rank {rank}] vit: loss rel err {err:.2e}  update rel diff {ud:.2e}  eval rel diff {ev:.2e}  replicas identical {same}",
          flush=True)
    ok = err < 2e-3 and ud < 0.2 and ev < 2e-2 and same
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    pkg = importlib.import_module("3d_poseestimation_b200")
    trainer = importlib.import_module("3d_poseestimation_b200.trainer")
    if os.environ.get("DP_CASE") == "vit":
        sys.exit(vit_case(pkg, rank, world, dev))
    # B rows per rank; the defaults are the small case, DP_CASE=bench is the shape bench.py times (configs[2])
    H, S, B, steps = (1024, 2, 8192, 4) if os.environ.get("DP_CASE") == "bench" else (256, 1, 512, 6)
    use_graph = os.environ.get("DP_USE_GRAPH", "1") != "0"
    rs = np.random.RandomState(0)
    xg = torch.from_numpy(rs.normal(0, 1, (steps, world * B, 34)).astype(np.float32))
    yg = torch.from_numpy(rs.normal(0, 1, (steps, world * B, 51)).astype(np.float32))

    def build():
        torch.manual_seed(3)
        return pkg.LinearModel(34, 51, linear_size=H, num_stage=S, p_dropout=0.5, BN=True).to(dev)
    # single process on the global batch (every rank computes it: cheap, and no cross-rank transfer of the answer)
    ref = build()
    tr_ref = trainer.LifterTrainer(ref, lr=1e-3, loss="mse", optimizer="adam", seed=5, use_graph=use_graph)
    ref_losses = [float(tr_ref.step(xg[s].to(dev), yg[s].to(dev)).item()) for s in range(steps)]
    # data parallel
    m = build()
    if os.environ.get("DP_DIVERGENT_INIT") == "1":                    # replicas built from different seeds: the trainer
        torch.manual_seed(100 + rank)                                 # must broadcast rank 0's
        m2 = pkg.LinearModel(34, 51, linear_size=H, num_stage=S, p_dropout=0.5, BN=True).to(dev)
        if rank != 0:
            m = m2
    tr = trainer.LifterTrainer(m, lr=1e-3, loss="mse", optimizer="adam", seed=5, world_size=world, rank=rank,
                               use_graph=use_graph, require_symmetric=True)
    assert (m._dp is not None) == (world > 1), "symmetric-memory SyncBN exchange was not set up"
    losses = []
    for s in range(steps):
        tr.step(xg[s, rank * B:(rank + 1) * B].to(dev), yg[s, rank * B:(rank + 1) * B].to(dev))
        losses.append(tr.loss_value())
    torch.cuda.synchronize()
    err = max(abs(a - b) / abs(b) for a, b in zip(losses, ref_losses))
    p0 = build()._flat
    # the UPDATE the steps made (Adam's first steps are sign-like: a gradient entry within bf16 noise of zero flips its
    # whole step when the summation order changes, N shards vs one batch: measured 0.01-0.12; a rank missing from the
    # exchange or local BatchNorm statistics give >= 0.5) and the running mean relative to its scale
    ud = (m._flat - ref._flat).norm().item() / (ref._flat - p0).norm().item()
    rm = (m._rm - ref._rm).abs().max().item() / ref._rm.abs().max().item()
    # identical parameters on every rank (deterministic exchange + reduce-scatter)
    mine = m._flat.clone()
    other = mine.clone()
    dist.broadcast(other, src=0)
    same = bool(torch.equal(mine, other))
    print(f"[rank {rank}] loss rel err {err:.2e}  update rel diff {ud:.2e}  running_mean rel diff {rm:.2e}  "
          f"replicas identical {same}", flush=True)
    ok = err < 2e-3 and ud < 0.2 and rm < 5e-3 and same
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
