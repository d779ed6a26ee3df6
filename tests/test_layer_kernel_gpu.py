"""The whole-layer forward kernel (GEMM + column statistics + grid barrier + BatchNorm / ReLU / dropout / residual from
TMEM in one launch; csrc/gemm_sm100.cu FUSED instantiation) against (a) the per-layer C-ABI path, which never takes it
(separate GEMM, reduce, finalize and apply launches -- the arithmetic the oracle tests pinned in round 1), in the same
process on the same inputs, and (b) the CPU oracle.  Shapes: ragged row counts (an odd number of 128-row blocks leaves a
phantom block in the last CTA pair), one and several column blocks, residual layers, with and without dropout."""
import ctypes as C
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import lifter_oracle as lo
from oracle.philox import dropout_keep_mask

pytestmark = pytest.mark.gpu
T = torch.from_numpy


def _forward(model, x, whole, step):
    """Train-mode forward through the single-call entry point (whole = True: takes the layer kernel where it fits) or
    layer by layer (never does).  Returns (out, saved workspace, BN buffers after)."""
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    ag = importlib.import_module("3d_poseestimation_b200._autograd")
    lib, cfg = cabi.lib(), C.byref(model._cfg)
    B = x.shape[0]
    fmt = model._fmt_code(model.train_format)
    saved = torch.zeros(cabi.check(lib.lifter_train_saved_bytes(cfg, B)), dtype=torch.uint8, device=x.device)
    out = torch.empty(B, model.output_size, device=x.device)
    t = ag._make_ctx(model, x, out, saved, fmt, True, step, B, model.row_offset)
    st = cabi.stream_ptr()
    if whole:
        os.environ["LIFTER_FUSE_LAYER"] = "1"               # off by default; read on every call
        try:
            cabi.check(lib.lifter_train_forward(cfg, C.byref(t), st))
        finally:
            del os.environ["LIFTER_FUSE_LAYER"]
    else:
        for l in range(2 * model.num_stage + 1):
            cabi.check(lib.lifter_train_fwd_layer_gemm(cfg, C.byref(t), l, st))
            cabi.check(lib.lifter_train_fwd_layer_finish(cfg, C.byref(t), l, st))
        cabi.check(lib.lifter_train_fwd_output(cfg, C.byref(t), st))
    torch.cuda.synchronize()
    return out, saved, (model._rm.clone(), model._rv.clone(), model._nbt.clone())


@pytest.mark.parametrize("B,H,S,p,row_offset", [(8192, 1024, 2, 0.5, 0), (1100, 512, 1, 0.5, 32), (300, 256, 2, 0.5, 16),
                                                (2000, 1024, 1, 0.0, 0), (4096, 2048, 1, 0.5, 4096)])
def test_whole_layer_kernel_matches_the_separate_launches(pkg, B, H, S, p, row_offset):
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    params, buffers = lo.numpy_init(34, 51, H, S, seed=B % 97)
    x = T(np.random.RandomState(B).normal(0, 1, (B, 34)).astype(np.float32)).cuda()
    res = []
    for whole in (False, True):
        m = pkg.LinearModel(34, 51, linear_size=H, num_stage=S, p_dropout=p, BN=True)
        m.load_state_dict({**params, **buffers})
        m = m.cuda().train()
        m.dropout_seed, m.row_offset = 1234, row_offset
        m._operands(m._fmt_code(m.train_format))
        before = cabi.lib().lifter_launch_count()
        res.append(_forward(m, x, whole, step=3) + (cabi.lib().lifter_launch_count() - before,))
    (o_a, s_a, (rm_a, rv_a, n_a), l_a), (o_b, s_b, (rm_b, rv_b, n_b), l_b) = res
    L = 2 * S + 1
    assert l_b == l_a - 3 * L, f"the whole-layer kernel did not run: {l_a} vs {l_b} launches"   # 4 launches -> 1 per layer
    # same arithmetic, same rounding points: the outputs agree to the last bits (the fp64 column sums are added in another
    # order, which can move a scale / shift by one fp32 ulp and with it the odd 16-bit rounding)
    d = (o_a - o_b).abs().max().item()
    assert d <= 2e-3 * o_a.abs().max().item(), f"out: max diff {d}"
    # the saved workspace holds z, a, the active bits and the statistics of every layer: almost every byte identical
    same = (s_a == s_b).float().mean().item()
    assert same > 0.999, f"saved workspace: {same:.6f} of the bytes identical"
    torch.testing.assert_close(rm_a, rm_b, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(rv_a, rv_b, rtol=1e-5, atol=1e-6)
    assert torch.equal(n_a, n_b) and int(n_b[0]) == 1
    # and against the oracle (bf16 operands: 1e-2 level)
    masks = [dropout_keep_mask(B, H, s, 3, 1234, p, row_offset=row_offset) for s in range(L)] if p > 0 else None
    ref, nb = lo.lifter_forward(params, buffers, x.cpu(), num_stage=S, training=True, p_dropout=p, masks=masks)
    rel = ((o_b.cpu() - ref).norm() / ref.norm()).item()
    assert rel < 2e-2, rel
    np.testing.assert_allclose(rm_b[0].cpu().numpy(), nb["batch_norm1.running_mean"].numpy(), rtol=2e-2, atol=2e-3)


def test_whole_layer_kernel_falls_back_where_it_does_not_fit(pkg):
    """p = 0.25 (8-bit lanes), an unaligned shard offset, or more than two tiles per CTA pair: the separate launches run."""
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    params, buffers = lo.numpy_init(34, 51, 256, 1, seed=4)
    for B, p, off in ((512, 0.25, 0), (512, 0.5, 7), (40000, 0.5, 0)):
        m = pkg.LinearModel(34, 51, linear_size=256, num_stage=1, p_dropout=p, BN=True)
        m.load_state_dict({**params, **buffers})
        m = m.cuda().train()
        m.dropout_seed, m.row_offset = 5, off
        x = torch.randn(B, 34, device="cuda")
        m._operands(m._fmt_code(m.train_format))            # weight packing is not part of the count
        before = cabi.lib().lifter_launch_count()
        o_b, _, _ = _forward(m, x, True, step=0)
        n = cabi.lib().lifter_launch_count() - before
        assert n == 1 + 3 * 3 + 1, n                       # prep_input, (GEMM, reduce + finalize, apply) x 3, output GEMM
        masks = [dropout_keep_mask(B, 256, s, 0, 5, p, row_offset=off) for s in range(3)]
        ref, _ = lo.lifter_forward(params, buffers, x.cpu(), num_stage=1, training=True, p_dropout=p, masks=masks)
        assert ((o_b.cpu() - ref).norm() / ref.norm()).item() < 2e-2
