"""GEMM timing sweep on the B200 (CUDA events): separates per-launch, per-tile and per-K-block cost."""
import ctypes as C
import importlib
import os
import sys

import torch

sys.path.insert(0, os.getcwd())
cabi = importlib.import_module("3d_poseestimation_b200._cabi")
lib = cabi.lib()
dev = torch.device("cuda")


def time_gemm(M, N, K, fmt=1, no_cluster=0, out_mode=0, nbuf=4, iters=20, affine=True, a_mn=0, b_mn=0, splits=1):
    dt = torch.bfloat16 if fmt else torch.float16
    As = [(torch.randn(K, M, device=dev) if a_mn else torch.randn(M, K, device=dev)).to(dt) for _ in range(nbuf)]
    W = ((torch.randn(K, N, device=dev) if b_mn else torch.randn(N, K, device=dev)) * 0.03).to(dt)
    if out_mode == 0:
        D = torch.empty(M, N, device=dev, dtype=dt)
    else:
        D = torch.empty(lib.lifter_gemm_effective_splits(K, splits) * lib.lifter_gemm_slab_rows(M), N, device=dev)
    scale, shift = torch.ones(N, device=dev), torch.zeros(N, device=dev)
    st = cabi.stream_ptr()

    def launch(i):
        d = cabi.GemmDesc()
        A = As[i % nbuf]
        d.A, d.lda, d.B, d.ldb = A.data_ptr(), A.stride(0), W.data_ptr(), W.stride(0)
        d.M, d.N, d.K, d.fmt, d.out_mode, d.D, d.ldd = M, N, K, fmt, out_mode, D.data_ptr(), N
        d.a_mn_major, d.b_mn_major, d.splits, d.no_cluster = a_mn, b_mn, splits, no_cluster
        if affine and out_mode == 0:
            d.relu, d.scale, d.shift = 1, scale.data_ptr(), shift.data_ptr()
        cabi.check(lib.lifter_gemm(C.byref(d), st))
    for i in range(3):
        launch(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        launch(i)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    return us, 2.0 * M * N * K / us / 1e6


if __name__ == "__main__":
    print("M N K cluster | us TFLOP/s")
    for nc in (1, 0):
        for (M, N, K) in [(148 * 128, 256, 64), (148 * 128, 256, 1024), (148 * 128, 256, 4096), (148 * 128, 256, 16384),
                          (148 * 128 * 4, 256, 1024), (9472, 1024, 1024), (8192, 1024, 1024), (9472 * 4, 1024, 1024),
                          (65536, 1024, 1024), (9472, 1024, 4096)]:
            us, tf = time_gemm(M, N, K, no_cluster=nc)
            print(f"{M:7d} {N:5d} {K:6d} cl={'off' if nc else 'on '} | {us:9.2f} us {tf:8.1f} TF")
    print("wgrad-shaped (MN-major both, split 4, fp32 slabs):")
    for nc in (1, 0):
        us, tf = time_gemm(1024, 1024, 8192, a_mn=1, b_mn=1, out_mode=1, splits=4, no_cluster=nc)
        print(f"   1024  1024   8192 cl={'off' if nc else 'on '} | {us:9.2f} us {tf:8.1f} TF")
    print("dgrad-shaped (B MN-major):")
    for nc in (1, 0):
        us, tf = time_gemm(8192, 1024, 1024, b_mn=1, no_cluster=nc)
        print(f"   8192  1024   1024 cl={'off' if nc else 'on '} | {us:9.2f} us {tf:8.1f} TF")
