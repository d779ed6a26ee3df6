"""MyViT eval-mode forward on one B200: one JSON line in the shape of bench.py's.

value        poses/s with the key points resident in HBM (CUDA events around `reps` forwards, after warm-up)
e2e          the same through pinned HOST buffers: H2D of the key points, forward, D2H of the joints, all in the timed region
roofline     the dominant GEMM (MLP-up projection with the GELU epilogue, M = tokens, N = 1024, K = 256) timed alone against
             the measured bf16 tensor peak of MEASURED_PEAKS.json (burst figure, kernel timed alone)
cpu_baseline the fp32 oracle of the reference module (oracle/vit_oracle.py, PyTorch CPU, all threads) on a bounded sample
Then a per-kernel split of one chunk from the library's trace hook (events serialise the launches: shares only).
"""
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
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
peak = 1665.4
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:
    pass
torch.manual_seed(0)
m = pkg.MyViT().cuda().eval()
H, N, nb = 256, 17, 2
flop_pose = 2 * N * (2 * H + nb * (H * 3 * H + H * H + 2 * H * 4 * H) + H * (H // 2) + (H // 2) * 3) + nb * 4 * N * N * H


def timed(fn, reps=5, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


x = torch.randn(B, 17, 2, device="cuda")
sweep = {c: timed(lambda: m.infer(x, chunk_poses=c)) for c in (2048, 4096, 8192, 16384, 32768, 65536)}
chunk = min(sweep, key=sweep.get)
ms = sweep[chunk]
xh = torch.randn(B, 17, 2).pin_memory()
oh = torch.empty(B, 17, 3).pin_memory()


def host_to_host():
    xd = xh.to("cuda", non_blocking=True)
    oh.copy_(m.infer(xd, chunk_poses=chunk), non_blocking=True)


ms_e2e = timed(host_to_host)

# dominant GEMM alone: u [T][256] x W1 [1024][256] -> GELU -> m [T][1024], operands rotated past L2
T = chunk * N
sets = 4
A = [torch.randn(T, H, device="cuda").half() for _ in range(sets)]
W = (torch.randn(4 * H, H, device="cuda") / 16).half()
bias = torch.randn(4 * H, device="cuda")
D = [torch.empty(T, 4 * H, device="cuda", dtype=torch.float16) for _ in range(sets)]
it = [0]


def gemm():
    i = it[0] % sets
    it[0] += 1
    d = cabi.GemmDesc()
    d.A, d.lda, d.B, d.ldb = A[i].data_ptr(), H, W.data_ptr(), H
    d.M, d.N, d.K, d.fmt, d.out_mode, d.D, d.ldd, d.relu, d.splits = T, 4 * H, H, 0, cabi.OUT_16, D[i].data_ptr(), 4 * H, 2, 1
    d.shift = bias.data_ptr()
    cabi.check(cabi.lib().lifter_gemm(C.byref(d), cabi.stream_ptr()))


us_gemm = timed(gemm, reps=20) * 1e3
gflop = 2.0 * T * 4 * H * H / 1e9
del A, D

# CPU oracle on a bounded sample
sys.path.insert(0, ROOT)
from oracle import vit_oracle as V  # noqa: E402
params = {k: v.detach().cpu() for k, v in m.state_dict().items()}
xs = torch.randn(2048, 17, 2)
torch.set_num_threads(os.cpu_count())
with torch.no_grad():
    V.vit_forward(params, xs[:256])
    t0 = time.perf_counter()
    reps = 0
    while time.perf_counter() - t0 < 10.0:
        V.vit_forward(params, xs)
        reps += 1
    cpu_s = (time.perf_counter() - t0) / reps

print(json.dumps({
    "metric": "lifted poses/sec (MyViT eval inference)", "value": B / ms * 1e3, "unit": "poses/s", "n_gpus": 1, "ms_per_step": ms,
    "higher_is_better": True, "dtype": "fp16 GEMM operands, fp32 residual stream / accumulate / LayerNorm / softmax", "data": "synthetic",
    "config": {"workload": f"MyViT(chw=(1,17,2), n_blocks=2, hidden_d=256, n_heads=4) eval forward, {B} poses, {chunk} poses per pass",
               "chunk_sweep_ms": {str(k): round(v, 3) for k, v in sweep.items()}},
    "model_flop_per_pose": flop_pose, "model_tflops": B * flop_pose / ms / 1e9,
    "e2e": {"value": B / ms_e2e * 1e3, "unit": "poses/s", "h2d_bytes_per_step": B * 17 * 2 * 4, "d2h_bytes_per_step": B * 17 * 3 * 4,
            "ms_per_step": ms_e2e},
    "roofline": {"bound": "tensor", "kernel": f"lifter_gemm_kernel<256, 2-SM, GELU> M={T} N=1024 K=256 (+bias+GELU epilogue)",
                 "achieved": gflop / us_gemm * 1e3, "peak": peak, "unit": "TFLOP/s", "frac": gflop / us_gemm * 1e3 / peak,
                 "us_per_launch": us_gemm, "traffic": None, "algorithmic_bytes": 2.0 * (T * H + 4 * H * H + T * 4 * H),
                 "note": "K = 256: four K-blocks per tile, so the 16-bit store path and the GELU arithmetic of the epilogue bound this launch, not the MMA"},
    "cpu_baseline": {"value": 2048 / cpu_s, "unit": "poses/s", "cores": os.cpu_count(), "kind": "port",
                     "sample": "2048 of the poses, oracle/vit_oracle.py (PyTorch fp32 CPU)"}}))

# per-kernel split of one pass (events serialise the launches: shares, not absolute times)
m.infer(x[:chunk], chunk_poses=chunk)
torch.cuda.synchronize()
cabi.check(cabi.lib().lifter_trace_begin(cabi.stream_ptr()))
m.infer(x[:chunk], chunk_poses=chunk)
buf = C.create_string_buffer(1 << 16)
cabi.lib().lifter_trace_end(buf, len(buf))
agg = {}
for line in buf.value.decode().splitlines():
    k, v = line.rsplit(" ", 1)
    agg[k] = agg.get(k, 0.0) + float(v)
tot = sum(agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
    print(f"# {k:28s} {v:9.1f} us  {100 * v / tot:5.1f} %")
