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


def time_gemm(M, N, K, fmt=1, cluster=0, out_mode=0, nbuf=4, iters=20, affine=True, a_mn=0, b_mn=0, splits=1,
              timeline=False, residual=False, max_stages=0):
    dt = torch.bfloat16 if fmt else torch.float16
    As = [(torch.randn(K, M, device=dev) if a_mn else torch.randn(M, K, device=dev)).to(dt) for _ in range(nbuf)]
    W = ((torch.randn(K, N, device=dev) if b_mn else torch.randn(N, K, device=dev)) * 0.03).to(dt)
    if out_mode == 0:
        D = torch.empty(M, N, device=dev, dtype=dt)
    else:
        D = torch.empty(lib.lifter_gemm_effective_splits(K, splits) * lib.lifter_gemm_slab_rows(M), N, device=dev)
    scale, shift = torch.ones(N, device=dev), torch.zeros(N, device=dev)
    R = torch.randn(M, N, device=dev).to(dt) if residual else None
    tl = torch.zeros(160 * 32, dtype=torch.int64, device=dev) if timeline else None
    st = cabi.stream_ptr()

    def launch(i):
        d = cabi.GemmDesc()
        A = As[i % nbuf]
        d.A, d.lda, d.B, d.ldb = A.data_ptr(), A.stride(0), W.data_ptr(), W.stride(0)
        d.M, d.N, d.K, d.fmt, d.out_mode, d.D, d.ldd = M, N, K, fmt, out_mode, D.data_ptr(), N
        d.a_mn_major, d.b_mn_major, d.splits, d.cluster = a_mn, b_mn, splits, cluster
        d.max_stages = max_stages
        if R is not None:
            d.residual, d.ldr = R.data_ptr(), N
        if tl is not None:
            d.timeline = tl.data_ptr()
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
    if tl is not None:
        show_timeline(tl.cpu().view(160, 32))
    return us, 2.0 * M * N * K / us / 1e6


def show_timeline(t):
    """Phases of the last launch, per CTA, in us relative to the earliest CTA start (median / max over CTAs)."""
    t = t[t[:, 0] > 0].double()
    t0 = t[:, 0].min()
    names = {0: "cta start", 1: "setup done", 2: "first MMA issued"}
    for it in range(4):
        names[4 + 3 * it] = f"tile {it}: last MMA issued"
        names[5 + 3 * it] = f"tile {it}: accumulator ready"
        names[6 + 3 * it] = f"tile {it}: epilogue done"
    for slot in sorted(names):
        col = t[:, slot]
        col = col[col > 0]
        if len(col) == 0:
            continue
        rel = (col - t0) / 1e3
        print(f"      {names[slot]:28s} n={len(col):3d} median {rel.median():7.2f}  min {rel.min():7.2f}  max {rel.max():7.2f} us")


if __name__ == "__main__":
    for cl in (2, 1):
        for st in (2, 3, 4, 5):
            for M in (8192, 65536):
                us, tf = time_gemm(M, 1024, 1024, cluster=cl, max_stages=st, nbuf=2)
                print(f"cl={cl} stages<={st} M={M}: {us:9.2f} us {tf:8.1f} TF")
