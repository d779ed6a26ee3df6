"""Where the data-parallel step's extra time goes (run under torchrun): the configs[2] step timed with CUDA events under
the library's debug switches.  One JSON line per variant from rank 0.
    default                      in-kernel SyncBN exchange + per-layer reduce-scatter / Adam / all-gather ranges
    LIFTER_DP_DEBUG_NO_EXCHANGE  local BatchNorm statistics (WRONG numerics, timing probe): cost of the ten exchanges
    LIFTER_FUSE_OPTIM=0          one reduce-scatter / Adam / all-gather launch after backward instead of six ranges
    LIFTER_DP_NCCL_GRADS=1       NCCL all-reduce of the flat gradient + replicated Adam
Also prints whether symmetric memory exposes an NVSwitch multicast mapping on this box."""
import importlib
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
pkg = importlib.import_module("3d_poseestimation_b200")
trainer = importlib.import_module("3d_poseestimation_b200.trainer")
torch.manual_seed(1)
model = pkg.LinearModel(34, 51, linear_size=1024, num_stage=2, p_dropout=0.5, BN=True).to(dev)
tr = trainer.LifterTrainer(model, lr=1e-4, loss="mse", optimizer="adam", seed=2, world_size=world, rank=rank)
x, y = [t.to(dev) for t in bench.synth_batch(8192, rank)]
for _ in range(10):
    tr.step(x, y)
dist.barrier()
torch.cuda.synchronize()
n = 400
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    tr.step(x, y)
e1.record()
dist.barrier()
torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
mc = None
if tr._sym is not None:
    h = tr._sym_hdl
    mc = {"multicast_ptr": int(getattr(h, "multicast_ptr", 0) or 0), "has": hasattr(h, "multicast_ptr")}
if rank == 0:
    print(json.dumps({"world": world, "variant": os.environ.get("VARIANT", "default"), "ms_per_step": float(t.item()),
                      "dp_adam": bool(tr._dp_adam), "symmetric": tr._sym is not None, "multicast": mc,
                      "launches": tr.graph_launches()}), flush=True)
dist.barrier()
dist.destroy_process_group()
