"""MyViT training step on one B200: forward + MSE + backward (+ torch.optim.Adam over
model.parameters(), as train_1.py:39,94-96 drives it) at 8,192 poses per step; one bench-shaped JSON line, then a
per-kernel split of one forward + backward from the library's trace hook (events serialise the launches: shares only)."""
import ctypes as C
import importlib
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("3d_poseestimation_b200")
cabi = importlib.import_module("3d_poseestimation_b200._cabi")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
torch.manual_seed(0)
m = pkg.MyViT().cuda().train()
opt = torch.optim.Adam(m.parameters(), lr=1e-4)
H, N, nb = 256, 17, 2
fwd = 2 * N * (2 * H + nb * (H * 3 * H + H * H + 2 * H * 4 * H) + H * (H // 2) + (H // 2) * 3) + nb * 4 * N * N * H
flop_pose = 3 * fwd            # dgrad + wgrad ~ 2x forward (the 2-wide input layer has no data gradient)
sets = [(torch.randn(B, 17, 2, device="cuda"), torch.randn(B, 17, 3, device="cuda")) for _ in range(4)]


def step(i, with_opt=True):
    x, y = sets[i % 4]
    opt.zero_grad(set_to_none=True)
    loss = torch.nn.functional.mse_loss(m(x), y)
    loss.backward()
    if with_opt:
        opt.step()
    return loss


def timed(fn, reps=20, warm=5):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ms = timed(step)
ms_noopt = timed(lambda i: step(i, False))
# the fused trainer: flat parameters, fused MSE, ONE Adam launch that also refreshes the operands, CUDA-graph replay
VT = importlib.import_module("3d_poseestimation_b200.vit_trainer").VitTrainer
m2 = pkg.MyViT().cuda().train()
tr = VT(m2, lr=1e-4)
ms_fused = timed(lambda i: tr.step(*sets[i % 4]), reps=40, warm=5)
tr_e = VT(pkg.MyViT().cuda().train(), lr=1e-4, use_graph=False)
ms_fused_eager = timed(lambda i: tr_e.step(*sets[i % 4]), reps=40, warm=5)
os.environ["LIFTER_VIT_WGRAD_STREAM"] = "0"          # A/B: weight gradients on the chain's own stream (read at capture)
tr_1 = VT(pkg.MyViT().cuda().train(), lr=1e-4)
ms_fused_one_stream = timed(lambda i: tr_1.step(*sets[i % 4]), reps=40, warm=5)
del os.environ["LIFTER_VIT_WGRAD_STREAM"]
os.environ["LIFTER_VIT_FUSE_GELU_BWD"] = "1"         # A/B: GELU backward in the epilogue of the data-gradient GEMM (default: own pass)
tr_2 = VT(pkg.MyViT().cuda().train(), lr=1e-4)
ms_fused_gelu_pass = timed(lambda i: tr_2.step(*sets[i % 4]), reps=40, warm=5)
del os.environ["LIFTER_VIT_FUSE_GELU_BWD"]
# CPU oracle on a bounded sample
from oracle import vit_oracle as V  # noqa: E402
params = {k: v.detach().cpu().clone().requires_grad_(k != "pos_embed") for k, v in m.state_dict().items()}
copt = torch.optim.Adam([v for k, v in params.items() if k != "pos_embed"], lr=1e-4)
xs, ys = torch.randn(256, 17, 2), torch.randn(256, 17, 3)
torch.set_num_threads(os.cpu_count())


def cpu_step():
    copt.zero_grad()
    torch.nn.functional.mse_loss(V.vit_forward(params, xs), ys).backward()
    copt.step()


cpu_step()
t0, reps = time.perf_counter(), 0
while time.perf_counter() - t0 < 10.0:
    cpu_step()
    reps += 1
cpu_s = (time.perf_counter() - t0) / reps
print(json.dumps({
    "metric": "lifted poses/sec (MyViT fwd+bwd+Adam)", "value": B / ms_fused * 1e3, "unit": "poses/s", "n_gpus": 1,
    "ms_per_step": ms_fused, "ms_per_step_eager_launches": ms_fused_eager, "ms_per_step_one_stream": ms_fused_one_stream, "ms_per_step_gelu_backward_in_dgrad_epilogue": ms_fused_gelu_pass, "gpu_launches": tr.graph_launches(),
    "autograd_torch_optim": {"value": B / ms * 1e3, "ms_per_step": ms, "ms_per_step_without_optimizer": ms_noopt},
    "higher_is_better": True,
    "dtype": "bf16 operands / activations / stored gradients, fp32 accumulate, fp32 parameters and parameter gradients",
    "data": "synthetic", "config": {"workload": f"MyViT(chw=(1,17,2), n_blocks=2, hidden_d=256, n_heads=4) training step, {B} poses, MSE, "
                                                 "VitTrainer (flat Adam, fused MSE, CUDA-graph replay), 4 rotating batches; saved activations of a step exceed L2"},
    "model_flop_per_pose": flop_pose, "model_tflops": B * flop_pose / ms_fused / 1e9,
    "cpu_baseline": {"value": 256 / cpu_s, "unit": "poses/s", "cores": os.cpu_count(), "kind": "port",
                     "sample": "256 poses per step, oracle/vit_oracle.py + autograd + torch.optim.Adam (PyTorch fp32 CPU)"}}))
step(0)
torch.cuda.synchronize()
os.environ["LIFTER_VIT_WGRAD_STREAM"] = "0"        # the split serialises launches anyway: keep them on one stream
cabi.check(cabi.lib().lifter_trace_begin(cabi.stream_ptr()))
step(1, False)
buf = C.create_string_buffer(1 << 17)
cabi.lib().lifter_trace_end(buf, len(buf))
agg, cnt = {}, {}
for line in buf.value.decode().splitlines():
    k, v = line.rsplit(" ", 1)
    agg[k] = agg.get(k, 0.0) + float(v)
    cnt[k] = cnt.get(k, 0) + 1
tot = sum(agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
    print(f"# {k:28s} n={cnt[k]:3d} {v:9.1f} us  {100 * v / tot:5.1f} %")
