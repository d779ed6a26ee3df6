"""One launch sequence of the MLP-up GEMM of the transformer lifter (bias + GELU epilogue) for ncu captures."""
import ctypes as C
import importlib
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
cabi = importlib.import_module("3d_poseestimation_b200._cabi")
T, H = int(sys.argv[1]) if len(sys.argv) > 1 else 16384 * 17, 256
A = torch.randn(T, H, device="cuda").half()
W = (torch.randn(4 * H, H, device="cuda") / 16).half()
bias = torch.randn(4 * H, device="cuda")
D = torch.empty(T, 4 * H, device="cuda", dtype=torch.float16)
d = cabi.GemmDesc()
d.A, d.lda, d.B, d.ldb = A.data_ptr(), H, W.data_ptr(), H
d.M, d.N, d.K, d.fmt, d.out_mode, d.D, d.ldd, d.relu, d.splits = T, 4 * H, H, 0, cabi.OUT_16, D.data_ptr(), 4 * H, 2, 1
d.shift = bias.data_ptr()
for _ in range(3):
    cabi.check(cabi.lib().lifter_gemm(C.byref(d), cabi.stream_ptr()))
torch.cuda.synchronize()
print("ok")
