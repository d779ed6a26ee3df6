"""lifter_loss_static: the no-allocation variant of the fused loss (its per-block partial sums live in a file-scope
``static __device__`` array of the library instead of caller-provided scratch; include/lifter_b200.h).  It has no
counterpart in the reference -- it was added for callers that cannot call
cudaMalloc -- so the checker here is the EXISTING entry point ``lifter_loss`` (itself held to the oracle and the
reference's golden vectors in tests/test_train_gpu.py), called through the same C ABI on the same inputs.

This file is collected last on purpose (``pytest --collect-only``).  It passes alone
in a fresh process (the first call does the per-device address look-up) and after the other tests.
"""
import importlib

import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(cabi, pred, tgt, kind, std, accumulate, out_static, out_ref):
    """One call of each entry point on the same inputs; returns (dpred_static, dpred_ref)."""
    lib = cabi.lib()
    B, J = pred.shape[0], pred.shape[1] // 3
    st = cabi.stream_ptr()
    d_s, d_r = torch.full_like(pred, 3.0), torch.full_like(pred, -3.0)
    scratch = torch.empty(cabi.check(lib.lifter_loss_scratch_bytes(B, J)), dtype=torch.uint8, device=pred.device)
    cabi.check(lib.lifter_loss(cabi.ptr(pred), cabi.ptr(tgt), B, J, kind, B, cabi.ptr(std), cabi.ptr(d_r),
                               cabi.ptr(scratch), cabi.ptr(out_ref), accumulate, st))
    cabi.check(lib.lifter_loss_static(cabi.ptr(pred), cabi.ptr(tgt), B, J, kind, B, cabi.ptr(std), cabi.ptr(d_s),
                                      cabi.ptr(out_static), accumulate, st))
    return d_s, d_r


def test_static_scratch_loss_matches_lifter_loss_on_two_calls_in_a_row():
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    cabi.check(cabi.lib().lifter_device_check())
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator().manual_seed(11)
    J = 17
    std = (torch.rand(3 * J, generator=g) + 0.5).to(dev)
    out_s, out_r = torch.zeros(1 + J, device=dev), torch.zeros(1 + J, device=dev)
    # call 1: a large batch fills most of the shared array; call 2: a SMALLER batch of another loss kind right behind it
    # on the same stream, accumulating its metric onto call 1's -- the array still holds call 1's partial sums, which must
    # not leak into the result; call 3: the largest batch the entry point accepts
    calls = [(50000, cabi.LOSS_MSE, 0), (777, cabi.LOSS_MPJPE, 1), (cabi.LOSS_STATIC_MAX_ROWS, cabi.LOSS_L1, 0)]
    for B, kind, accumulate in calls:
        pred = torch.randn(B, 3 * J, generator=g).to(dev)
        tgt = torch.randn(B, 3 * J, generator=g).to(dev)
        d_s, d_r = _both(cabi, pred, tgt, kind, std, accumulate, out_s, out_r)
        torch.cuda.synchronize()
        # the tolerances tests/test_train_gpu.py::test_fused_loss_matches_oracle holds lifter_loss to (1e-5 on the loss,
        # 1e-4 on the gradient and the metric); same kernels in the same order, so in practice the bits agree
        torch.testing.assert_close(out_s[:1], out_r[:1], rtol=1e-5, atol=0)
        torch.testing.assert_close(out_s[1:], out_r[1:], rtol=1e-4, atol=0)
        torch.testing.assert_close(d_s, d_r, rtol=1e-4, atol=1e-9)
        assert torch.isfinite(out_s).all() and float(out_s[0]) > 0


def test_static_scratch_loss_rejects_what_does_not_fit_and_allocates_nothing():
    cabi = importlib.import_module("3d_poseestimation_b200._cabi")
    lib = cabi.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    J = 17
    x = torch.zeros(8, 3 * J, device=dev)
    out = torch.zeros(1 + J, device=dev)
    st = cabi.stream_ptr()
    with pytest.raises(ValueError, match="exceed the static scratch"):
        cabi.check(lib.lifter_loss_static(cabi.ptr(x), cabi.ptr(x), cabi.LOSS_STATIC_MAX_ROWS + 1, J, 0, 8, None, None,
                                          cabi.ptr(out), 0, st))
    with pytest.raises(ValueError, match="joints unsupported"):
        cabi.check(lib.lifter_loss_static(cabi.ptr(x), cabi.ptr(x), 8, 22, 0, 8, None, None, cabi.ptr(out), 0, st))
    with pytest.raises(ValueError, match="null pointer"):
        cabi.check(lib.lifter_loss_static(None, cabi.ptr(x), 8, J, 0, 8, None, None, cabi.ptr(out), 0, st))
    # no allocation: neither the driver's free memory nor torch's allocator moves over many calls (no gradient, no std)
    cabi.check(lib.lifter_loss_static(cabi.ptr(x), cabi.ptr(x), 8, J, 0, 8, None, None, cabi.ptr(out), 0, st))
    torch.cuda.synchronize()
    free0, reserved0 = torch.cuda.mem_get_info()[0], torch.cuda.memory_reserved()
    for _ in range(200):
        cabi.check(lib.lifter_loss_static(cabi.ptr(x), cabi.ptr(x), 8, J, 0, 8, None, None, cabi.ptr(out), 0, st))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0 and torch.cuda.memory_reserved() == reserved0
    assert float(out[0]) == 0.0
